// How much does a kernel boundary cost inside a CUDA graph on this GPU, and how much of it does
// programmatic dependent launch (griddepcontrol) take back?  Chain of N small dependent kernels.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o launch_gap launch_gap.cu && ./launch_gap
#include <cuda_runtime.h>
#include <stdio.h>

__global__ void __launch_bounds__(256) step_kernel(float* p, int pdl, int work) {
  if (pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  __shared__ float s[256];
  s[threadIdx.x] = threadIdx.x;  // stand-in for a prologue (barrier init, TMEM alloc ...)
  __syncthreads();
  if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float v = p[i];
  for (int k = 0; k < work; ++k) v = v * 1.0001f + s[(threadIdx.x + k) & 255] * 1e-9f;
  p[i] = v;
}

static float run(int n, int blocks, int pdl, int work, float* buf) {
  cudaStream_t st;
  cudaStreamCreate(&st);
  cudaGraph_t g;
  cudaGraphExec_t ge;
  cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
  for (int i = 0; i < n; ++i) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(blocks);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, step_kernel, buf, pdl, work);
  }
  cudaStreamEndCapture(st, &g);
  if (cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) { printf("instantiate failed\n"); return -1; }
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int w = 0; w < 3; ++w) cudaGraphLaunch(ge, st);
  cudaStreamSynchronize(st);
  cudaEventRecord(e0, st);
  for (int r = 0; r < 10; ++r) cudaGraphLaunch(ge, st);
  cudaEventRecord(e1, st);
  cudaStreamSynchronize(st);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) printf("error: %s\n", cudaGetErrorString(e));
  cudaGraphExecDestroy(ge);
  cudaGraphDestroy(g);
  cudaStreamDestroy(st);
  return ms / 10 / n * 1000.f;  // us per kernel
}

int main() {
  float* buf;
  cudaMalloc(&buf, 148 * 8 * 256 * sizeof(float));
  cudaMemset(buf, 0, 148 * 8 * 256 * sizeof(float));
  const int n = 1000;
  printf("| CTAs | work iters | plain us/kernel | PDL us/kernel |\n|---|---|---|---|\n");
  const int works[3] = {0, 2000, 20000};
  const int grids[3] = {148, 592, 1184};
  for (int wi = 0; wi < 3; ++wi)
    for (int gi = 0; gi < 3; ++gi) {
      const float a = run(n, grids[gi], 0, works[wi], buf);
      const float b = run(n, grids[gi], 1, works[wi], buf);
      printf("| %d | %d | %.2f | %.2f |\n", grids[gi], works[wi], a, b);
    }
  return 0;
}
