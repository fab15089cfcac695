// epi_probe.cu -- how fast can ONE warp per SM sub-partition run the hidden-layer epilogue pattern
// (MUFU.TANH + FFMA per element, one bf16x2 pack per pair, one 16-byte shared store per 8), and what
// slows it: a co-resident spinning warp, a second working warp, the pack instruction.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/epi_probe tools/epi_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ float tanh_approx(float x) { float y; asm volatile("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t pack_cvt(float lo, float hi) { uint32_t r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }
__device__ __forceinline__ uint32_t pack_alu(float lo, float hi) {
  return __byte_perm(__float_as_uint(lo) + 0x8000u, __float_as_uint(hi) + 0x8000u, 0x7632);
}
// swish on the FMA pipe only: x*sigmoid(x) with sigmoid from a clamped odd polynomial of tanh(g)
__device__ __forceinline__ float swish_poly(float g) {
  // tanh(g) ~ g*(c0 + c1 g^2 + c2 g^4 + c3 g^6) on |g|<=~3, clamped to +-1 (coefficients are placeholders:
  // this probe measures cost, not accuracy)
  const float g2 = g * g;
  float p = fmaf(g2, -2.1e-4f, 5.2e-3f);
  p = fmaf(p, g2, -5.3e-2f);
  p = fmaf(p, g2, 0.33f);
  p = fmaf(p, -g2, 1.0f);
  float t = fminf(fmaxf(g * p, -1.f), 1.f);
  return fmaf(g, t, g);
}

template <int MODE>
__device__ __forceinline__ void chunk(float (&r)[32], uint32_t saddr) {
  uint32_t o[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    float a = r[2 * i], b = r[2 * i + 1];
    if (MODE == 0) { o[i] = __float_as_uint(tanh_approx(a)) ^ __float_as_uint(tanh_approx(b)); continue; }
    float ya, yb;
    if (MODE == 5 && (i & 3) == 3) { ya = swish_poly(a); yb = swish_poly(b); }       // 1/4 of the pairs on the FMA pipe
    else { ya = fmaf(a, tanh_approx(a), a); yb = fmaf(b, tanh_approx(b), b); }
    o[i] = (MODE == 4) ? pack_alu(ya, yb) : pack_cvt(ya, yb);
  }
#pragma unroll
  for (int g = 0; g < 4; ++g)
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(saddr + 128u * g), "r"(o[4 * g]), "r"(o[4 * g + 1]),
                 "r"(o[4 * g + 2]), "r"(o[4 * g + 3]) : "memory");
}

// workers = number of warps that run the pattern; the other warps of the block spin on an mbarrier
// (SPIN = 1) or exit (SPIN = 0)
template <int MODE, int SPIN>
__global__ void __launch_bounds__(512, 1) k(float* out, long long* cyc, int iters, int workers) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(&bar);
  if (threadIdx.x == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar_a), "r"(workers * 32));
  __syncthreads();
  if (warp >= workers) {
    if (SPIN) {
      uint32_t ok = 0;
      while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar_a), "r"(0) : "memory");
    }
    return;
  }
  float r[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) r[i] = 0.01f * (float)(lane + i) - 0.3f;
  const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(smem) + warp * 4096 + (lane & 7) * 16 + (lane >> 3) * 1024;
  const uint32_t laddr = (uint32_t)__cvta_generic_to_shared(smem) + 16384 + threadIdx.x * 32;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    // fresh inputs every pass (the real epilogue gets them from TMEM): 8 x 16-byte shared loads
#pragma unroll
    for (int g = 0; g < 8; ++g)
      asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r[4 * g]), "=f"(r[4 * g + 1]), "=f"(r[4 * g + 2]), "=f"(r[4 * g + 3])
                   : "r"(laddr + 16u * g) : "memory");
    chunk<MODE>(r, saddr);
  }
  long long t1 = clock64();
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_a) : "memory");
  if (lane == 0 && blockIdx.x == 0) cyc[warp] = t1 - t0;
  float s = 0;
  for (int i = 0; i < 32; ++i) s += r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE, int SPIN>
void run(const char* name, int workers, int threads) {
  float* out; long long* cyc;
  cudaMalloc(&out, 148 * 256 * 4); cudaMalloc(&cyc, 64);
  const int iters = 2000;
  for (int rep = 0; rep < 2; ++rep) k<MODE, SPIN><<<148, threads, 32768>>>(out, cyc, iters, workers);
  long long c[8]; cudaMemcpy(c, cyc, 64, cudaMemcpyDeviceToHost);
  cudaError_t e = cudaDeviceSynchronize();
  printf("%-62s %7.2f cycles per 32-element warp column (warp 0), %7.2f (last worker)%s\n", name,
         (double)c[0] / (iters * 32.0), (double)c[workers - 1] / (iters * 32.0), e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  run<0, 0>("MUFU.TANH only, 1 warp/SMSP", 4, 128);
  run<1, 0>("tanh+ffma+cvt pack+sts, 1 warp/SMSP", 4, 128);
  run<1, 1>("tanh+ffma+cvt pack+sts, 1 warp/SMSP + 1 spinning warp/SMSP", 4, 256);
  run<1, 1>("same + 2 spinning warps/SMSP", 4, 384);
  run<1, 1>("same + 3 spinning warps/SMSP", 4, 512);
  run<1, 0>("tanh+ffma+cvt pack+sts, 2 warps/SMSP (per-warp time)", 8, 256);
  run<4, 0>("tanh+ffma+ALU pack+sts, 1 warp/SMSP", 4, 128);
  run<4, 0>("tanh+ffma+ALU pack+sts, 2 warps/SMSP (per-warp time)", 8, 256);
  run<5, 0>("3/4 tanh + 1/4 FMA-pipe polynomial, cvt pack, 1 warp/SMSP", 4, 128);
  run<5, 0>("3/4 tanh + 1/4 FMA-pipe polynomial, 2 warps/SMSP (per-warp)", 8, 256);
  return 0;
}
