// umma_probe.cu -- hardware probe for the tcgen05 encodings the ensemble kernel relies on.
// Each invocation runs ONE case (own process => a trap cannot poison later cases):
//
//   umma_probe ss   <cg> <N> <K> <swap>          A,B from smem (no-swizzle K-major); swap=1
//                                                exchanges the LBO/SBO descriptor fields
//   umma_probe ts   <cg> <N> <K> <order>         A from TMEM (packed bf16x2); order=1 swaps the
//                                                halves inside each 32-bit column
//   umma_probe chain                              2-layer chain, layer-1 accumulator converted
//                                                in place into the layer-2 TMEM A operand
//   umma_probe time <ss|ts> <cg> <N> <K> <reps>   cycles per MMA, back-to-back issue
//
// Prints "PASS"/"FAIL" with max abs error against a double-precision host product.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include <string>
#include "../constrainedmbpo_b200/csrc/sm100_ptx.cuh"

using namespace sm100;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)

static uint16_t f2bf(float f) {
  uint32_t u; memcpy(&u, &f, 4);
  uint32_t r = u + 0x7FFF + ((u >> 16) & 1);
  return (uint16_t)(r >> 16);
}
static float bf2f(uint16_t h) { uint32_t u = (uint32_t)h << 16; float f; memcpy(&f, &u, 4); return f; }

// canonical no-swizzle K-major image: (r,k) -> (r/8)*sbo + (k/8)*lbo + (r%8)*16 + (k%8)*2
static void make_image(const std::vector<uint16_t>& M, int rows, int K, std::vector<uint16_t>& img) {
  const int lbo = 128, sbo = (K / 8) * 128;
  img.assign((size_t)rows * K, 0);
  for (int r = 0; r < rows; ++r)
    for (int k = 0; k < K; ++k) {
      size_t off = (size_t)(r / 8) * sbo + (size_t)(k / 8) * lbo + (r % 8) * 16 + (k % 8) * 2;
      img[off / 2] = M[(size_t)r * K + k];
    }
}

struct Params {
  const uint16_t* a_img;   // SS: per-CTA canonical image [cg][128*K]; TS: row-major [cg*128][K]
  const uint16_t* b_img;   // per-CTA canonical image [cg][(N/cg)*K]
  const uint16_t* b2_img;  // chain only
  float* d;                // [cg*128][N]
  int N, K, swap, order, reps;
  long long* cycles;
  int* status;
};

template <int CG, bool TS>
__global__ void __launch_bounds__(128, 1) probe_kernel(Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t cta = (CG == 2) ? cluster_ctarank() : 0;
  const int N = p.N, K = p.K, Nh = N / CG;
  const uint32_t a_bytes = TS ? 0u : 128u * K * 2u;
  const uint32_t b_bytes = (uint32_t)Nh * K * 2u;
  uint8_t* sA = smem;
  uint8_t* sB = smem + a_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + a_bytes + b_bytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
  const uint32_t bar_load = smem_u32(&bars[0]), bar_mma = smem_u32(&bars[1]);

  if (tid == 0) {
    mbar_init(bar_load, 1);
    mbar_init(bar_mma, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0) {
    mbar_arrive_expect_tx(bar_load, a_bytes + b_bytes);
    if (!TS) bulk_g2s(smem_u32(sA), p.a_img + (size_t)cta * 128 * K, a_bytes, bar_load);
    bulk_g2s(smem_u32(sB), p.b_img + (size_t)cta * Nh * K, b_bytes, bar_load);
  }
  if (warp == 0) {
    tmem_alloc<CG>(smem_u32(tmem_slot), 512);
    tmem_relinquish<CG>();
  }
  mbar_wait(bar_load, 0, p.status, 1);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
  const uint32_t a_col = 256;  // TS: A operand columns [256, 256+K/2)

  if (TS) {
    const uint16_t* arow = p.a_img + ((size_t)cta * 128 + tid) * K;
    for (int j = 0; j < K / 16; ++j) {
      uint32_t r[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        uint32_t lo = arow[16 * j + 2 * i], hi = arow[16 * j + 2 * i + 1];
        r[i] = p.order ? ((lo << 16) | hi) : ((hi << 16) | lo);
      }
      tmem_st8(tmem_base + lane_base + a_col + 8 * j, r);
    }
    tmem_wait_st();
  }
  tc_fence_before();
  if (CG == 2) cluster_sync(); else __syncthreads();
  tc_fence_after();

  long long t0 = 0;
  if (cta == 0 && tid == 0) {
    const uint32_t idesc = make_idesc(128 * CG, N, kFmtBF16);
    uint32_t lbo = 128, sbo = (K / 8) * 128;
    if (p.swap) { uint32_t t = lbo; lbo = sbo; sbo = t; }
    t0 = clock64();
    for (int rep = 0; rep < p.reps; ++rep) {
      for (int j = 0; j < K / 16; ++j) {
        const uint64_t bdesc = make_smem_desc(smem_u32(sB) + j * 256, lbo, sbo);
        const uint32_t acc = (j > 0 || rep > 0) ? 1u : 0u;
        if (TS) {
          mma_bf16_ts<CG>(tmem_base, tmem_base + a_col + 8 * j, bdesc, idesc, acc);
        } else {
          const uint64_t adesc = make_smem_desc(smem_u32(sA) + j * 256, lbo, sbo);
          mma_bf16_ss<CG>(tmem_base, adesc, bdesc, idesc, acc);
        }
      }
    }
    if (CG == 2) mma_commit_2sm(bar_mma, 0x3); else mma_commit(bar_mma);
  }
  mbar_wait(bar_mma, 0, p.status, 2);
  tc_fence_after();
  if (cta == 0 && tid == 0 && p.cycles) *p.cycles = clock64() - t0;

  for (int c = 0; c < N / 16; ++c) {
    uint32_t r[16];
    tmem_ld16(tmem_base + lane_base + c * 16, r);
    tmem_wait_ld();
    float* drow = p.d + ((size_t)cta * 128 + tid) * N + c * 16;
#pragma unroll
    for (int i = 0; i < 16; ++i) drow[i] = __uint_as_float(r[i]);
  }
  tc_fence_before();
  if (CG == 2) cluster_sync(); else __syncthreads();
  if (warp == 0) tmem_dealloc<CG>(tmem_base, 512);
}

// two-layer chain, cta_group::1, TS: D1 = A0 B1^T (K0), A1 = bf16(D1) written IN PLACE over the
// first half of D1's columns, D2 = A1 B2^T into a second region.
__global__ void __launch_bounds__(128, 1) chain_kernel(Params p, int K0) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5;
  const int N = p.N;  // hidden width (K of layer 2)
  const uint32_t b1_bytes = (uint32_t)N * K0 * 2u, b2_bytes = (uint32_t)N * N * 2u;
  uint8_t* sB1 = smem;
  uint8_t* sB2 = smem + b1_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + b1_bytes + b2_bytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
  const uint32_t bar_load = smem_u32(&bars[0]), bar_m1 = smem_u32(&bars[1]), bar_m2 = smem_u32(&bars[2]);
  if (tid == 0) {
    mbar_init(bar_load, 1); mbar_init(bar_m1, 1); mbar_init(bar_m2, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0) {
    mbar_arrive_expect_tx(bar_load, b1_bytes + b2_bytes);
    bulk_g2s(smem_u32(sB1), p.b_img, b1_bytes, bar_load);
    bulk_g2s(smem_u32(sB2), p.b2_img, b2_bytes, bar_load);
  }
  if (warp == 0) { tmem_alloc<1>(smem_u32(tmem_slot), 512); tmem_relinquish<1>(); }
  mbar_wait(bar_load, 0, p.status, 1);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
  const uint32_t colP = 0, colQ = 256, colA0 = 480;
  {  // stage A0 (row-major global) into TMEM
    const uint16_t* arow = p.a_img + (size_t)tid * K0;
    for (int j = 0; j < K0 / 16; ++j) {
      uint32_t r[8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
        r[i] = ((uint32_t)arow[16 * j + 2 * i + 1] << 16) | arow[16 * j + 2 * i];
      tmem_st8(tmem_base + lane_base + colA0 + 8 * j, r);
    }
    tmem_wait_st();
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    const uint32_t idesc = make_idesc(128, N, kFmtBF16);
    for (int j = 0; j < K0 / 16; ++j)
      mma_bf16_ts<1>(tmem_base + colP, tmem_base + colA0 + 8 * j,
                     make_smem_desc(smem_u32(sB1) + j * 256, 128, (K0 / 8) * 128), idesc, j > 0);
    mma_commit(bar_m1);
  }
  mbar_wait(bar_m1, 0, p.status, 2);
  tc_fence_after();
  // in-place conversion: read 32 fp32 columns, write 16 packed columns at half the offset
  for (int c = 0; c < N / 16; ++c) {
    uint32_t r[16], o[8];
    tmem_ld16(tmem_base + lane_base + colP + c * 16, r);
    tmem_wait_ld();
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = pack_bf16x2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
    tmem_st8(tmem_base + lane_base + colP + c * 8, o);
  }
  tmem_wait_st();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    const uint32_t idesc = make_idesc(128, N, kFmtBF16);
    for (int j = 0; j < N / 16; ++j)
      mma_bf16_ts<1>(tmem_base + colQ, tmem_base + colP + 8 * j,
                     make_smem_desc(smem_u32(sB2) + j * 256, 128, (N / 8) * 128), idesc, j > 0);
    mma_commit(bar_m2);
  }
  mbar_wait(bar_m2, 0, p.status, 3);
  tc_fence_after();
  for (int c = 0; c < N / 16; ++c) {
    uint32_t r[16];
    tmem_ld16(tmem_base + lane_base + colQ + c * 16, r);
    tmem_wait_ld();
    float* drow = p.d + (size_t)tid * N + c * 16;
#pragma unroll
    for (int i = 0; i < 16; ++i) drow[i] = __uint_as_float(r[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<1>(tmem_base, 512);
}

template <int CG, bool TS>
static void launch(Params p, size_t smem_bytes) {
  CK(cudaFuncSetAttribute(probe_kernel<CG, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(CG); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem_bytes; cfg.stream = 0;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  CK(cudaLaunchKernelEx(&cfg, probe_kernel<CG, TS>, p));
}

int main(int argc, char** argv) {
  if (argc < 2) { printf("usage: see source header\n"); return 1; }
  std::string mode = argv[1];
  bool timing = mode == "time";
  bool chain = mode == "chain";
  bool ts; int cg, N, K, variant = 0, reps = 1;
  if (chain) { ts = true; cg = 1; N = 208; K = 48; }
  else if (timing) { ts = std::string(argv[2]) == "ts"; cg = atoi(argv[3]); N = atoi(argv[4]); K = atoi(argv[5]); reps = atoi(argv[6]); }
  else { ts = mode == "ts"; cg = atoi(argv[2]); N = atoi(argv[3]); K = atoi(argv[4]); variant = atoi(argv[5]); }
  const int M = 128 * cg, Nh = N / cg;
  srand(1234);
  std::vector<uint16_t> A((size_t)M * K), B((size_t)N * K), B2((size_t)N * N);
  for (auto& v : A) v = f2bf((float)((rand() % 17) - 8) / 8.0f);
  for (auto& v : B) v = f2bf((float)((rand() % 13) - 6) / 16.0f);
  for (auto& v : B2) v = f2bf((float)((rand() % 13) - 6) / 64.0f);
  // device operands
  std::vector<uint16_t> a_dev_host, b_dev_host, img;
  if (ts) a_dev_host = A;
  else for (int c = 0; c < cg; ++c) {
    std::vector<uint16_t> sub(A.begin() + (size_t)c * 128 * K, A.begin() + (size_t)(c + 1) * 128 * K);
    make_image(sub, 128, K, img); a_dev_host.insert(a_dev_host.end(), img.begin(), img.end());
  }
  for (int c = 0; c < cg; ++c) {
    std::vector<uint16_t> sub(B.begin() + (size_t)c * Nh * K, B.begin() + (size_t)(c + 1) * Nh * K);
    make_image(sub, Nh, K, img); b_dev_host.insert(b_dev_host.end(), img.begin(), img.end());
  }
  std::vector<uint16_t> b2_img; if (chain) make_image(B2, N, N, b2_img);
  uint16_t *dA, *dB, *dB2 = nullptr; float* dD; long long* dCyc; int* dStatus;
  CK(cudaMalloc(&dA, a_dev_host.size() * 2)); CK(cudaMalloc(&dB, b_dev_host.size() * 2));
  CK(cudaMalloc(&dD, (size_t)M * N * 4)); CK(cudaMalloc(&dCyc, 8)); CK(cudaMalloc(&dStatus, 4));
  CK(cudaMemcpy(dA, a_dev_host.data(), a_dev_host.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, b_dev_host.data(), b_dev_host.size() * 2, cudaMemcpyHostToDevice));
  if (chain) { CK(cudaMalloc(&dB2, b2_img.size() * 2)); CK(cudaMemcpy(dB2, b2_img.data(), b2_img.size() * 2, cudaMemcpyHostToDevice)); }
  CK(cudaMemset(dD, 0xFF, (size_t)M * N * 4)); CK(cudaMemset(dCyc, 0, 8)); CK(cudaMemset(dStatus, 0, 4));
  Params p{dA, dB, dB2, dD, N, K, ts ? 0 : variant, ts ? variant : 0, reps, dCyc, dStatus};
  if (chain) {
    size_t smem_bytes = (size_t)N * K * 2 + (size_t)N * N * 2 + 64;
    CK(cudaFuncSetAttribute(chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    chain_kernel<<<1, 128, smem_bytes>>>(p, K);
  } else {
    size_t smem_bytes = (ts ? 0 : (size_t)128 * K * 2) + (size_t)Nh * K * 2 + 64;
    if (cg == 1) { if (ts) launch<1, true>(p, smem_bytes); else launch<1, false>(p, smem_bytes); }
    else         { if (ts) launch<2, true>(p, smem_bytes); else launch<2, false>(p, smem_bytes); }
  }
  cudaError_t e = cudaDeviceSynchronize();
  int status = 0; long long cyc = 0;
  if (e != cudaSuccess) { printf("%s: KERNEL ERROR %s\n", mode.c_str(), cudaGetErrorString(e)); return 3; }
  CK(cudaMemcpy(&status, dStatus, 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(&cyc, dCyc, 8, cudaMemcpyDeviceToHost));
  if (timing) {
    long long n_mma = (long long)reps * (K / 16);
    printf("time %s cg=%d N=%d K=%d reps=%d: %lld cycles total, %.1f cycles/MMA (model N/2=%d)\n",
           ts ? "ts" : "ss", cg, N, K, reps, cyc, (double)cyc / n_mma, N / 2);
    return 0;
  }
  std::vector<float> D((size_t)M * N);
  CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
  double max_err = 0, max_ref = 0;
  if (chain) {
    std::vector<float> D1((size_t)128 * N);
    for (int r = 0; r < 128; ++r) for (int n = 0; n < N; ++n) {
      double s = 0; for (int k = 0; k < K; ++k) s += (double)bf2f(A[(size_t)r * K + k]) * bf2f(B[(size_t)n * K + k]);
      D1[(size_t)r * N + n] = bf2f(f2bf((float)s));
    }
    for (int r = 0; r < 128; ++r) for (int n = 0; n < N; ++n) {
      double s = 0; for (int k = 0; k < N; ++k) s += (double)D1[(size_t)r * N + k] * bf2f(B2[(size_t)n * N + k]);
      max_err = fmax(max_err, fabs(s - D[(size_t)r * N + n])); max_ref = fmax(max_ref, fabs(s));
    }
  } else {
    for (int r = 0; r < M; ++r) for (int n = 0; n < N; ++n) {
      double s = 0; for (int k = 0; k < K; ++k) s += (double)bf2f(A[(size_t)r * K + k]) * bf2f(B[(size_t)n * K + k]);
      max_err = fmax(max_err, fabs(s - D[(size_t)r * N + n])); max_ref = fmax(max_ref, fabs(s));
    }
  }
  bool ok = max_err <= 1e-3 * fmax(1.0, max_ref) && std::isfinite(max_err);
  printf("%s cg=%d N=%d K=%d variant=%d: %s max_abs_err=%.4g max_ref=%.4g status=%d cycles=%lld\n",
         mode.c_str(), cg, N, K, variant, ok ? "PASS" : "FAIL", max_err, max_ref, status, cyc);
  return ok ? 0 : 1;
}
