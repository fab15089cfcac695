// mufu_probe.cu -- throughput of the transcendental / conversion / packed-math instructions the
// fused epilogue can be built from, per SM (ops per clock), at 4 / 8 / 16 warps per SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int OP>
__device__ __forceinline__ void op8(float (&v)[8], uint32_t (&u)[8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    if (OP == 0) asm volatile("tanh.approx.f32 %0, %0;" : "+f"(v[i]));
    if (OP == 1) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
    if (OP == 2) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
    if (OP == 3) asm volatile("tanh.approx.bf16x2 %0, %0;" : "+r"(u[i]));
    if (OP == 4) asm volatile("tanh.approx.f16x2 %0, %0;" : "+r"(u[i]));
    if (OP == 5) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %1;" : "=r"(u[i]) : "f"(v[i]));
    if (OP == 6) asm volatile("fma.rn.bf16x2 %0, %0, %0, %0;" : "+r"(u[i]));
    if (OP == 7) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(v[i]));
    if (OP == 8) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(u[i]));
    if (OP == 9) asm volatile("tanh.approx.bf16 %0, %0;" : "+h"(*reinterpret_cast<uint16_t*>(&u[i])));
  }
}

template <int OP>
__global__ void k(float* out, long long* cyc, int iters) {
  float v[8]; uint32_t u[8];
  for (int i = 0; i < 8; ++i) { v[i] = 0.001f * (threadIdx.x + i); u[i] = 0x3c003c00u + threadIdx.x + i; }
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) op8<OP>(v, u);
  long long t1 = clock64();
  float s = 0; for (int i = 0; i < 8; ++i) s += v[i] + __uint_as_float(u[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int OP>
void run(const char* name, int elems_per_op) {
  float* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  const int iters = 2000;
  printf("%-24s", name);
  for (int warps : {4, 8, 16}) {
    k<OP><<<148, warps * 32>>>(out, cyc, iters);
    k<OP><<<148, warps * 32>>>(out, cyc, iters);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    double ops = (double)iters * 8 * warps * 32;
    printf("  %2dw: %6.2f instr-lanes/clk/SM (%6.2f elem/clk)", warps, ops / c, ops * elems_per_op / c);
  }
  printf("\n");
  cudaFree(out); cudaFree(cyc);
}

int main() {
  run<0>("tanh.approx.f32", 1);
  run<1>("ex2.approx.f32", 1);
  run<2>("rcp.approx.f32", 1);
  run<3>("tanh.approx.bf16x2", 2);
  run<4>("tanh.approx.f16x2", 2);
  run<9>("tanh.approx.bf16", 1);
  run<8>("ex2.approx.bf16x2", 2);
  run<5>("cvt.rn.bf16x2.f32", 2);
  run<6>("fma.rn.bf16x2", 2);
  run<7>("fma.rn.f32", 1);
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return 0;
}
