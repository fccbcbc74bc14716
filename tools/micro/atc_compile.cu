// compile-only harness: `nvcc -Xptxas -v` on the attention kernels alone (5 s instead of the whole library)
//   nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xptxas -v -I mdgen_finetune_b200/csrc -o /tmp/atc tools/micro/atc_compile.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdlib.h>
#include <string>
#include "common.cuh"
#include "attention_tc.cuh"
namespace mdgen { long long* g_dbg_buf = nullptr; int g_dbg_cta = 0; thread_local long long g_launches = 0; Profiler g_prof; void set_error(const std::string&) {} int num_sms() { return 148; } }
int main() { mdgen::AttnParams<mdgen::bf16> p{}; return mdgen::launch_attention_tc(p, 24, 0) + mdgen::launch_attention_simt(p, 24, 0); }
