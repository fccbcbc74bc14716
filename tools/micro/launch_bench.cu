// Per-launch fixed cost on B200 for the launch configurations this library uses, and what programmatic dependent launch
// (PDL) recovers of it.  Measured: ~3.9 us per stream-ordered empty launch, PDL -2.2 us on compute-only kernels; in the
// real rollout (memory-heavy kernels whose tails dominate the boundary) PDL gained nothing and was not adopted.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o launch_bench launch_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_plain(float* p) { if (p && threadIdx.x == 9999) p[0] = 1.f; }
__global__ void __cluster_dims__(2, 1, 1) k_cluster(float* p) { if (p && threadIdx.x == 9999) p[0] = 1.f; }
__global__ void k_pdl(float* p) {
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (p && threadIdx.x == 9999) p[0] = 1.f;
}
// ~20 us of work so that launch overhead of the NEXT kernel can overlap
__global__ void k_work(float* p, int iters, int pdl) {
  if (pdl) asm volatile("griddepcontrol.launch_dependents;");
  if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
  float a = threadIdx.x;
  for (int i = 0; i < iters; ++i) a = a * 1.0001f + 0.5f;
  if (a == 12345.f) p[0] = a;
}
static float time_loop(cudaStream_t st, int n, void (*body)(cudaStream_t, int)) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  body(st, 10); cudaStreamSynchronize(st);
  cudaEventRecord(a, st); body(st, n); cudaEventRecord(b, st); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms * 1e3f / n;
}
static float* g_p;
static const int BIG = 213 * 1024;
static void launch_pdl(void (*kern)(float*, int, int), dim3 g, dim3 b, size_t sm, cudaStream_t st, int iters) {
  cudaLaunchConfig_t cfg = {}; cfg.gridDim = g; cfg.blockDim = b; cfg.dynamicSmemBytes = sm; cfg.stream = st;
  cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaLaunchKernelEx(&cfg, kern, g_p, iters, 1);
}
int main() {
  cudaMalloc(&g_p, 1024);
  cudaStream_t st; cudaStreamCreate(&st);
  cudaFuncSetAttribute(k_plain, cudaFuncAttributeMaxDynamicSharedMemorySize, BIG);
  cudaFuncSetAttribute(k_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, BIG);
  cudaFuncSetAttribute(k_work, cudaFuncAttributeMaxDynamicSharedMemorySize, BIG);
  const int n = 200;
  printf("empty, 0 B smem          : %.2f us/launch\n", time_loop(st, n, [](cudaStream_t s, int m) { for (int i = 0; i < m; ++i) k_plain<<<148, 320, 0, s>>>(g_p); }));
  printf("empty, 213 KB smem       : %.2f us/launch\n", time_loop(st, n, [](cudaStream_t s, int m) { for (int i = 0; i < m; ++i) k_plain<<<148, 320, BIG, s>>>(g_p); }));
  printf("empty, 213 KB, cluster 2 : %.2f us/launch\n", time_loop(st, n, [](cudaStream_t s, int m) { for (int i = 0; i < m; ++i) k_cluster<<<148, 320, BIG, s>>>(g_p); }));
  printf("alternating 0 B / 213 KB : %.2f us/launch\n", time_loop(st, n, [](cudaStream_t s, int m) { for (int i = 0; i < m; ++i) { if (i & 1) k_plain<<<148, 320, BIG, s>>>(g_p); else k_plain<<<2000, 256, 0, s>>>(g_p); } }));
  for (int iters : {2000, 20000}) {
    float t1 = time_loop(st, n, [](cudaStream_t s, int m) { for (int i = 0; i < m; ++i) k_work<<<148, 320, BIG, s>>>(g_p, 0, 0); });
    (void)t1;
  }
  static int g_iters;
  for (int iters : {4000, 40000}) {
    g_iters = iters;
    float a = time_loop(st, n, [](cudaStream_t s, int m) { for (int i = 0; i < m; ++i) k_work<<<148, 320, BIG, s>>>(g_p, g_iters, 0); });
    float b = time_loop(st, n, [](cudaStream_t s, int m) { for (int i = 0; i < m; ++i) launch_pdl(k_work, 148, 320, BIG, s, g_iters); });
    float c = time_loop(st, n, [](cudaStream_t s, int m) { for (int i = 0; i < m; ++i) { if (i & 1) k_work<<<148, 320, BIG, s>>>(g_p, g_iters, 0); else k_work<<<2000, 256, 0, s>>>(g_p, g_iters / 8, 0); } });
    float d = time_loop(st, n, [](cudaStream_t s, int m) { for (int i = 0; i < m; ++i) { if (i & 1) launch_pdl(k_work, 148, 320, BIG, s, g_iters); else launch_pdl(k_work, 2000, 256, 0, s, g_iters / 8); } });
    printf("work iters=%d: plain %.2f us/launch, PDL %.2f us/launch; alternating plain %.2f, alternating PDL %.2f\n", iters, a, b, c, d);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
