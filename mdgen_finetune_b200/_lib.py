"""Build + ctypes binding of libmdgen_b200.so (the C ABI in include/mdgen_b200.h).

There is no CPU / PyTorch fallback: if the shared library is missing or fails to load,
every entry point raises.  `build()` compiles csrc/model.cu in-tree for sm_100a with nvcc.
"""
import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(HERE, "libmdgen_b200.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC"]

# flag bits / enums of include/mdgen_b200.h
F_SIM, F_TPS, F_INPAINT, F_DESIGN, F_NO_AA_EMB, F_ABS_POS, F_ABS_TIME, F_NO_ROPE, F_PREPEND_IPA = (
    1 << 0, 1 << 1, 1 << 2, 1 << 3, 1 << 4, 1 << 5, 1 << 6, 1 << 7, 1 << 8)
PREC_BF16, PREC_FP32_CHECK = 0, 1
EXPORTS = ["mdgen_create", "mdgen_destroy", "mdgen_last_error", "mdgen_abi_version", "mdgen_load_weight",
           "mdgen_finalize_weights", "mdgen_set_conditioning", "mdgen_forward", "mdgen_euler_rollout", "mdgen_dirichlet_flow", "mdgen_set_dirichlet_tables", "mdgen_dirichlet_flags", "mdgen_prep_batch", "mdgen_decode_atom14",
           "mdgen_launches_per_forward", "mdgen_launch_count", "mdgen_workspace_bytes", "mdgen_profile_begin", "mdgen_profile_end",
           "mdgen_profile_category_name", "mdgen_debug_set_buffer", "mdgen_test_gemm", "mdgen_test_attention",
           "mdgen_variant_launches", "mdgen_gemm_noalloc", "mdgen_backward"]
NOALLOC_MAX_OPERAND_ELEMS = 1 << 21      # MDGEN_NOALLOC_MAX_OPERAND_ELEMS
# kernel variants counted by mdgen_variant_launches(i) (tests assert that the full-size cases reach the timed kernels)
VARIANTS = ["gemm_ln", "gemm_tc2", "gemm_tc", "attention_tc", "attention_seq4", "graph_replays", "qkv_attn4", "gemm_mlp", "gemm_ln4"]


class MdgenConfig(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "embed_dim", "num_layers", "mha_heads", "ipa_heads", "ipa_head_dim", "ipa_qk", "ipa_v",
        "latent_dim", "crop", "num_frames", "flags", "precision")] + [("time_multiplier", ctypes.c_float)]


def _sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))) + [
        os.path.join(ROOT, "include", "mdgen_b200.h")]


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(s) > t for s in _sources())


def build(force=False, verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a ... -> mdgen_finetune_b200/libmdgen_b200.so"""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [
        "-o", LIB_PATH, os.path.join(CSRC, "model.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB_PATH


_lib = None


def lib():
    """The loaded library.  Raises (never falls back) when it cannot be loaded."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`. "
                           "mdgen_finetune_b200 has no CPU fallback.")
    L = ctypes.CDLL(LIB_PATH)
    vp, i32, i64p, fp = ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p
    L.mdgen_create.argtypes = [ctypes.POINTER(MdgenConfig), ctypes.POINTER(vp)]
    L.mdgen_destroy.argtypes = [vp]
    L.mdgen_last_error.restype = ctypes.c_char_p
    L.mdgen_load_weight.argtypes = [vp, ctypes.c_char_p, fp, ctypes.POINTER(ctypes.c_int64), i32, i32, vp]
    L.mdgen_finalize_weights.argtypes = [vp, vp]
    L.mdgen_set_conditioning.argtypes = [vp, i32, i32, i32] + [fp] * 8 + [i64p, i64p, vp]
    L.mdgen_forward.argtypes = [vp, fp, fp, fp, vp]
    L.mdgen_euler_rollout.argtypes = [vp, fp, i32, fp, fp, vp]
    L.mdgen_dirichlet_flow.argtypes = [vp, fp, fp, vp, ctypes.c_double, ctypes.c_double, ctypes.c_float, ctypes.c_float, i32,
                                       ctypes.POINTER(ctypes.c_int32), vp]
    L.mdgen_set_dirichlet_tables.argtypes = [vp, i32, vp, vp, vp, ctypes.c_float, ctypes.c_float, i32, vp]
    L.mdgen_dirichlet_flags.argtypes = [vp, ctypes.POINTER(ctypes.c_int32), vp]
    L.mdgen_decode_atom14.argtypes = [i32, i32, i32, i32, fp, i32, fp, fp, vp, fp, vp, fp, fp, i32, fp, vp, vp]
    L.mdgen_prep_batch.argtypes = [i32, i32, i32, fp, fp, fp, i32, i32, fp, fp, vp, vp]
    L.mdgen_launches_per_forward.argtypes = [vp]
    L.mdgen_workspace_bytes.argtypes = [vp]
    L.mdgen_workspace_bytes.restype = ctypes.c_int64
    L.mdgen_launch_count.restype = ctypes.c_int64
    L.mdgen_variant_launches.argtypes = [i32]
    L.mdgen_variant_launches.restype = ctypes.c_int64
    L.mdgen_test_gemm.argtypes = [i32, i32, i32, fp, fp, fp, fp, i32, i32, vp]
    L.mdgen_test_attention.argtypes = [i32] * 5 + [fp] * 5 + [i32, i32, vp]
    L.mdgen_backward.argtypes = [vp, fp, fp, fp, fp, i32, ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_void_p), fp, vp]
    L.mdgen_gemm_noalloc.argtypes = [i32, i32, i32, fp, fp, fp, fp, vp]
    L.mdgen_debug_set_buffer.argtypes = [vp, i32]
    L.mdgen_profile_end.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int32), i32]
    L.mdgen_profile_category_name.argtypes = [i32]
    L.mdgen_profile_category_name.restype = ctypes.c_char_p
    for name in EXPORTS:
        if name not in ("mdgen_last_error", "mdgen_workspace_bytes", "mdgen_profile_category_name",
                        "mdgen_launch_count", "mdgen_variant_launches"):
            getattr(L, name).restype = ctypes.c_int32
    _lib = L
    return L


def check(status, what="mdgen"):
    """Maps the C status codes to the reference's exception types (SURVEY 8b error convention)."""
    if status == 0:
        return
    msg = lib().mdgen_last_error().decode("utf-8", "replace")
    if status == -1:
        raise ValueError(f"{what}: {msg}")
    raise RuntimeError(f"{what} failed ({status}): {msg}")


NCAT = 11


def profile_begin():
    check(lib().mdgen_profile_begin(), "mdgen_profile_begin")


def profile_end():
    """{category: (total_ms, launches)} for the launches since profile_begin()."""
    ms = (ctypes.c_float * NCAT)()
    n = (ctypes.c_int32 * NCAT)()
    check(lib().mdgen_profile_end(ms, n, NCAT), "mdgen_profile_end")
    return {lib().mdgen_profile_category_name(i).decode(): (float(ms[i]), int(n[i])) for i in range(NCAT) if n[i]}
