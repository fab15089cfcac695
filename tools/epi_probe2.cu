// epi_probe2.cu -- which ingredient of the hidden-layer epilogue keeps one warp per SM sub-partition
// from the 8 cycles per warp-wide MUFU.TANH the XU pipe allows.  Feature bits:
//   1 = FFMA per element (swish), 2 = bf16x2 pack per pair, 4 = 16-byte shared store per 8 elements,
//   8 = inputs reloaded from shared memory every pass (else: outputs feed the next pass)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ float tanh_approx(float x) { float y; asm volatile("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t pack_cvt(float lo, float hi) { uint32_t r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }

template <int F>
__global__ void __launch_bounds__(256, 1) k(float* out, long long* cyc, int iters) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float r[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) r[i] = 0.01f * (float)(lane + i) - 0.3f;
  float* sf = reinterpret_cast<float*>(smem);
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) sf[i] = 0.001f * (float)(i & 511) - 0.25f;
  __syncthreads();
  const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(smem) + 32768 + warp * 4096 + (lane & 7) * 16 + (lane >> 3) * 1024;
  const uint32_t laddr = (uint32_t)__cvta_generic_to_shared(smem) + warp * 4096 + lane * 16;
  uint32_t acc = 0;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (F & 8) {
#pragma unroll
      for (int g = 0; g < 8; ++g)
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r[4 * g]), "=f"(r[4 * g + 1]), "=f"(r[4 * g + 2]), "=f"(r[4 * g + 3])
                     : "r"(laddr + 512u * g) : "memory");
    }
    uint32_t o[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float a = r[2 * i], b = r[2 * i + 1];
      float ta = tanh_approx(a), tb = tanh_approx(b);
      float ya = (F & 1) ? fmaf(a, ta, a) : ta, yb = (F & 1) ? fmaf(b, tb, b) : tb;
      if (!(F & 8)) { r[2 * i] = ya; r[2 * i + 1] = yb; }
      o[i] = (F & 2) ? pack_cvt(ya, yb) : (__float_as_uint(ya) ^ __float_as_uint(yb));
    }
    if (F & 4) {
#pragma unroll
      for (int g = 0; g < 4; ++g)
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(saddr + 128u * g), "r"(o[4 * g]), "r"(o[4 * g + 1]),
                     "r"(o[4 * g + 2]), "r"(o[4 * g + 3]) : "memory");
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) acc ^= o[i];
    }
  }
  long long t1 = clock64();
  if (lane == 0 && blockIdx.x == 0) cyc[warp] = t1 - t0;
  float s = __uint_as_float(acc);
  for (int i = 0; i < 32; ++i) s += r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int F>
void run() {
  float* out; long long* cyc;
  cudaMalloc(&out, 148 * 256 * 4); cudaMalloc(&cyc, 64);
  const int iters = 2000;
  printf("features %2d [%s%s%s%s]:", F, (F & 1) ? "ffma " : "", (F & 2) ? "pack " : "", (F & 4) ? "sts " : "", (F & 8) ? "lds" : "feedback");
  for (int threads : {128, 256}) {
    cudaFuncSetAttribute(k<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    for (int rep = 0; rep < 2; ++rep) k<F><<<148, threads, 65536>>>(out, cyc, iters);
    long long c[8]; cudaMemcpy(c, cyc, 64, cudaMemcpyDeviceToHost);
    printf("   %d warp/SMSP: %6.2f cycles per warp-MUFU per SMSP", threads / 128, (double)c[0] / (iters * 32.0) / (threads / 128));
  }
  printf("  %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  cudaFree(out); cudaFree(cyc);
}

int main() {

  run<0>(); run<1>(); run<3>(); run<7>(); run<8>(); run<9>(); run<11>(); run<15>(); run<4>(); run<12>();
  return 0;
}
