"""TEST / BASELINE INFRASTRUCTURE ONLY -- stages the reference's own `mdgen` package for the GPU box.

/root/reference does not exist on the GPU box, and the reference is pure Python with no setup.py / pyproject (it cannot
be pip-installed).  This recipe copies the `src/mdgen` package, UNMODIFIED, from where it lies under /root/reference into
oracle/_ref/src/mdgen -- a git-ignored build output (like the built .so files it travels with the gpurun snapshot, it is
never committed).  oracle/ref_loader.py then imports the reference from there, so that `bench.py --impl reference` and
bench.py's `cpu_baseline` leg time the REFERENCE'S OWN modules on the box's host cores (cpu_baseline.kind = "reference").

Run by __graft_entry__.build() whenever /root/reference is present:   python oracle/stage_ref.py
"""
import filecmp
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/src/mdgen"
DST = os.path.join(HERE, "_ref", "src", "mdgen")
# what the velocity network, the sampler and the wrapper import; analysis / dataset need pyemma / mdtraj
SKIP = {"analysis.py", "dataset.py", "__pycache__"}


def stage(verbose=True):
    if not os.path.isdir(SRC):
        if verbose:
            print(f"stage_ref: {SRC} not present (GPU box): using the staged copy as is")
        return os.path.isdir(DST)
    n = 0
    for root, dirs, files in os.walk(SRC):
        dirs[:] = [d for d in dirs if d not in SKIP]
        rel = os.path.relpath(root, SRC)
        out = os.path.join(DST, rel) if rel != "." else DST
        os.makedirs(out, exist_ok=True)
        for f in files:
            if f in SKIP or not f.endswith(".py"):
                continue
            s, d = os.path.join(root, f), os.path.join(out, f)
            if not (os.path.exists(d) and filecmp.cmp(s, d, shallow=False)):
                shutil.copyfile(s, d)
            n += 1
    if verbose:
        print(f"stage_ref: {n} reference files staged under {DST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if stage() else 1)
