// wide_fwd.cu -- layer-at-a-time PETS-ensemble forward for layers wider than the fused kernel's 254
// columns (ensemble_fwd.cu), e.g. the (512, 512) nets every shipped CMBPO config uses
// (examples/config/antsafe/cmbpo_uc.py:36) and the 200 -> 1024 width sweep of BASELINE.json.
//
// One launch computes ONE FC layer (mbpo/models/fc.py:105-114) for all members of a row slab:
//     out[e] = act(A[e] . W[e] + b[e])
// A and out are bf16 tile images in the "chunked canonical" layout of cmbpo_internal.cuh; the head layer
// writes the same tile-major fp32 scratch as the fused kernel, so everything downstream is shared.
//
// Shape: persistent CTA pairs (cluster of 2, tcgen05 cta_group::2, M = 256).  A work item is
// (member slot, 256-row tile, 256-column N chunk); its K loop streams 64-column chunks of A (this
// CTA's 128 rows) and of W^T (this CTA's N half) through a 6-stage shared-memory ring filled by 1-D
// bulk async copies (TMA engine).  Two 256-column fp32 TMEM accumulators: the epilogue of item i
// (bias + activation + bf16 pack + 16-byte stores straight into the next layer's operand image) runs
// under the MMAs of item i+1.  Warp roles: 0 = loader (+ relay of the peer CTA's arrivals to the
// leader's barrier), 1 = MMA issuer (leader CTA), 2 = TMEM alloc, 3 idle, 4..11 = epilogue (TMEM lane
// quadrant = warp % 4, column half = (warp - 4) / 4: two warps per SM sub-partition keep the XU busy).
#include "cmbpo_internal.cuh"
#include "sm100_ptx.cuh"

namespace cmbpo {
using namespace sm100;

constexpr int kWideThreads = 384;
constexpr int kWideStages = 6;
constexpr int kWideEpiWarps = 8;
constexpr uint32_t kWideStageBytes = 2u * kWideBlk;                    // A block + B block
constexpr uint32_t kWideBarOff = kWideStages * kWideStageBytes;        // barriers behind the ring
constexpr uint32_t kWideSmemBytes = kWideBarOff + 32u * 8u + 16u;

constexpr int WB_FULL = 0;         // [kWideStages]
constexpr int WB_EMPTY = 8;        // [kWideStages]
constexpr int WB_ACC_FULL = 16;    // [2]
constexpr int WB_ACC_EMPTY = 18;   // [2]

enum WideWaitCode : int { WWC_LOADER = 11, WWC_MMA_FULL = 12, WWC_MMA_ACC = 13, WWC_EPI_ACC = 14 };

// 1 / (1 + e^-x) to fp32 accuracy: ex2.approx (2 ulp) + rcp.approx (1 ulp)
__device__ __forceinline__ float sigmoid_precise(float x) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.4426950408889634f * x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
  return r;
}

template <int ACT, bool PRECISE>
__device__ __forceinline__ float wide_activate(float g) {
  if constexpr (PRECISE) {   // tf32 path (rtol 1e-3): tanh.approx's ~2^-11 relative error would eat the whole budget
    if constexpr (ACT == ACT_SWISH) return g * sigmoid_precise(g);              // fc.py:20, g = x (no pre-scaling)
    else if constexpr (ACT == ACT_RELU) return fmaxf(g, 0.f);
    else if constexpr (ACT == ACT_TANH) return fmaf(2.0f, sigmoid_precise(2.0f * g), -1.0f);
    else if constexpr (ACT == ACT_SIGMOID) return sigmoid_precise(g);
    else return g;
  }
  // g is the pre-activation, already halved for ACT_SWISH / ACT_SIGMOID (weights and bias pre-scaled)
  if constexpr (ACT == ACT_SWISH) return fmaf(g, tanh_approx(g), g);          // x*sigmoid(x), fc.py:20
  else if constexpr (ACT == ACT_RELU) return fmaxf(g, 0.f);
  else if constexpr (ACT == ACT_TANH) return tanh_approx(g);
  else if constexpr (ACT == ACT_SIGMOID) return fmaf(0.5f, tanh_approx(g), 0.5f);
  else return g;
}

// Work items of one cluster: i = cluster_id, cluster_id + n_clusters, ...; i -> (slot e, tile, n chunk)
// with the N chunk fastest, then tiles, then members (all clusters work on one member's weights at a time).
struct WideItems {
  long long i, step, total;
  int n_tiles, n_chunks;
  __device__ __forceinline__ WideItems(int cluster_id, int n_clusters, int E, int n_tiles_, int n_chunks_)
      : i(cluster_id), step(n_clusters), total((long long)E * n_tiles_ * n_chunks_), n_tiles(n_tiles_), n_chunks(n_chunks_) {}
  __device__ __forceinline__ bool valid() const { return i < total; }
  __device__ __forceinline__ void next() { i += step; }
  __device__ __forceinline__ int chunk() const { return (int)(i % n_chunks); }
  __device__ __forceinline__ int tile() const { return (int)((i / n_chunks) % n_tiles); }
  __device__ __forceinline__ int slot() const { return (int)(i / ((long long)n_chunks * n_tiles)); }
  __device__ __forceinline__ long long count() const { return i < total ? (total - i + step - 1) / step : 0; }
};

// TF32: the operand images hold fp32 words (tf32-rounded) in 16 KB blocks of 32 K columns -- the same byte geometry as
// the bf16 blocks of 64 (8 core matrices of 128 bytes along K, LBO 128, SBO 1024) -- so the loader, the ring and the
// descriptor stepping (4 MMAs per stage, 256 bytes apart) are unchanged; tcgen05.mma kind::tf32 consumes K = 8 per
// instruction at half the bf16 rate, and the epilogue evaluates the activation to fp32 accuracy.
template <int HACT, bool TF32>
__global__ void __launch_bounds__(kWideThreads, 1) wide_layer_kernel(const __grid_constant__ WideParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const WideLayerPlan& lp = p.lp;
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const uint32_t cta = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const long long n_rows_all = p.n_rows_dev ? (long long)(*p.n_rows_dev) : p.n_rows;
  long long live = n_rows_all - p.row0;
  if (live > p.slab_rows) live = p.slab_rows;
  if (live < 0) live = 0;
  const int n_tiles = (int)((live + 2 * kTileRows - 1) / (2 * kTileRows));
  const int KC = lp.k_chunks, NCH = lp.n_chunks;

  const uint32_t bar0 = smem_u32(smem + kWideBarOff);
  auto bar = [bar0](int i) -> uint32_t { return bar0 + 8u * (uint32_t)i; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kWideBarOff + 32u * 8u);

  if (tid == 0) {
    for (int s = 0; s < kWideStages; ++s) {
      mbar_init(bar(WB_FULL + s), cta == 0 ? 2 : 1);    // leader: own fill + relayed peer fill
      mbar_init(bar(WB_EMPTY + s), 1);                  // tcgen05.commit
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(bar(WB_ACC_FULL + a), 1);               // tcgen05.commit
      mbar_init(bar(WB_ACC_EMPTY + a), 2 * kWideEpiWarps);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc<2>(smem_u32(tmem_slot), 512);
    tmem_relinquish<2>();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == 0) {
    // ===== loader: fill n of the (item, k chunk) walk goes to ring stage n % kWideStages
    if (lane == 0) {
      WideItems it(cluster_id, n_clusters, p.E, n_tiles, NCH);
      const long long total = it.count() * KC;
      long long fill_n = 0, relay_n = 0;
      int f_stage = 0, r_stage = 0, f_j = 0; uint32_t f_use = 0, r_use = 0;
      uint32_t idle = 0;
      // per-item state, decoded once per item (the 64-bit divisions of the item walk are slow on one thread)
      const uint8_t* a_src = nullptr; const uint8_t* b_src = nullptr; uint32_t b_bytes = 0;
      while (fill_n < total || (cta == 1 && relay_n < total)) {
        bool progress = false;
        if (fill_n < total && (f_use == 0 || mbar_try_wait(bar(WB_EMPTY + f_stage), (f_use - 1) & 1))) {
          if (f_j == 0) {
            const int e = it.slot(), tile = it.tile(), c = it.chunk();
            const int chunk_n = min(kWideNC, lp.npad - c * kWideNC);
            b_bytes = (uint32_t)(chunk_n >> 1) * 128u;
            const int in_slot = p.per_slot_input ? wide_input_of(p, e) : 0;
            a_src = p.a_img + (size_t)in_slot * p.a_slot_stride + (size_t)(tile * 2 + (int)cta) * KC * kWideBlk;
            b_src = p.w_img + (((size_t)wide_member_of(p, e) * NCH + c) * 2 + cta) * KC * kWideBlk;
          }
          const uint32_t dst = smem_u32(smem + (uint32_t)f_stage * kWideStageBytes);
          mbar_arrive_expect_tx(bar(WB_FULL + f_stage), kWideBlk + b_bytes);
          bulk_g2s(dst, a_src, kWideBlk, bar(WB_FULL + f_stage));
          bulk_g2s(dst + kWideBlk, b_src, b_bytes, bar(WB_FULL + f_stage));
          a_src += kWideBlk; b_src += kWideBlk;
          if (++f_j == KC) { f_j = 0; it.next(); }
          if (++f_stage == kWideStages) { f_stage = 0; ++f_use; }
          ++fill_n; progress = true;
        }
        if (cta == 1 && relay_n < fill_n && mbar_try_wait(bar(WB_FULL + r_stage), r_use & 1)) {
          // light (default-semantics) arrive: the data it announces was written by the async proxy and is
          // read by the async proxy (tcgen05.mma); the .release.cluster form costs a MEMBAR.ALL.GPU + L1
          // invalidate per K chunk, which alone held this kernel at ~2000 cycles per stage
          mbar_arrive_cluster_light(bar(WB_FULL + r_stage), 0);
          if (++r_stage == kWideStages) { r_stage = 0; ++r_use; }
          ++relay_n; progress = true;
        }
        if (progress) { idle = 0; continue; }
        __nanosleep(32);
        if (++idle > (SM100_WAIT_LIMIT >> 2)) {
          if (p.status) { *reinterpret_cast<volatile int*>(p.status) = (WWC_LOADER | 0x40000000); __threadfence_system(); }
          __trap();
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA): whole warp walks the loop, one elected lane issues
    if (cta == 0) {
      int stage = 0; uint32_t use = 0, n_item = 0;
      for (WideItems it(cluster_id, n_clusters, p.E, n_tiles, NCH); it.valid(); it.next(), ++n_item) {
        const int acc = n_item & 1;
        const uint32_t acc_use = n_item >> 1;
        const int chunk_n = min(kWideNC, lp.npad - it.chunk() * kWideNC);
        const uint32_t idesc = make_idesc(256, chunk_n, TF32 ? kFmtTF32 : kFmtBF16);
        if (acc_use > 0) mbar_wait(bar(WB_ACC_EMPTY + acc), (acc_use - 1) & 1, p.status, WWC_MMA_ACC);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kWideNC;
        for (int j = 0; j < KC; ++j) {
          mbar_wait(bar(WB_FULL + stage), use & 1, p.status, WWC_MMA_FULL);
          tc_fence_after();
          const uint32_t a_base = smem_u32(smem + (uint32_t)stage * kWideStageBytes);
          const uint32_t b_base = a_base + kWideBlk;
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < kWideKC / 16; ++ks) {
              if constexpr (TF32)
                mma_tf32_ss<2>(d_tmem, make_smem_desc(a_base + ks * 256, 128, 1024),
                               make_smem_desc(b_base + ks * 256, 128, 1024), idesc, (j > 0 || ks > 0) ? 1u : 0u);
              else
                mma_bf16_ss<2>(d_tmem, make_smem_desc(a_base + ks * 256, 128, 1024),
                               make_smem_desc(b_base + ks * 256, 128, 1024), idesc, (j > 0 || ks > 0) ? 1u : 0u);
            }
            mma_commit_2sm(bar(WB_EMPTY + stage), 0x3);
            if (j == KC - 1) mma_commit_2sm(bar(WB_ACC_FULL + acc), 0x3);
          }
          __syncwarp();
          if (++stage == kWideStages) { stage = 0; ++use; }
        }
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue warps: TMEM lane quadrant q, column half `half` of the item's N chunk
    const int q = warp & 3, half = (warp - 4) >> 2;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const int row_in_tile = q * 32 + lane;
    const uint32_t row_off = (uint32_t)(row_in_tile >> 3) * 1024u + (uint32_t)(row_in_tile & 7) * 16u;
    const int kc_next = lp.npad / lp.kc;       // K chunks of the image this layer writes (hidden layers)
    uint32_t n_item = 0;
    for (WideItems it(cluster_id, n_clusters, p.E, n_tiles, NCH); it.valid(); it.next(), ++n_item) {
      const int acc = n_item & 1;
      const int e = it.slot(), tile = it.tile(), c = it.chunk();
      const int chunk_n = min(kWideNC, lp.npad - c * kWideNC);
      const int nch = chunk_n >> 4, h0 = (nch + 1) >> 1;
      const int cb = half ? h0 : 0, ce = half ? nch : h0;
      const float* bias = p.bias + (size_t)wide_member_of(p, e) * lp.npad + c * kWideNC;
      mbar_wait(bar(WB_ACC_FULL + acc), (n_item >> 1) & 1, p.status, WWC_EPI_ACC);
      tc_fence_after();
      const uint32_t d_addr = tmem_base + lane_base + acc * kWideNC;
      const long long blk = (long long)tile * 2 + cta;     // 128-row block inside the slab
      uint32_t ra[16], rb[16];
      if (cb < ce) tmem_ld16(d_addr + 16 * cb, ra);
      if (!p.is_head) {
        uint8_t* out_blk = p.out_img + (size_t)e * p.out_slot_stride + (size_t)blk * kc_next * kWideBlk + row_off;
        auto finish = [&](int c16, const uint32_t (&r)[16]) {
          const int n0 = c * kWideNC + 16 * c16;             // first output column of this piece
          float b[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(bias + 16 * c16) + i);
            b[4 * i] = v.x; b[4 * i + 1] = v.y; b[4 * i + 2] = v.z; b[4 * i + 3] = v.w;
          }
          if constexpr (TF32) {   // 16 fp32 words = four 16-byte core-matrix rows, 128 bytes apart, inside one 32-column chunk
            uint8_t* dst = out_blk + (size_t)(n0 >> 5) * kWideBlk + (size_t)((n0 & 31) >> 2) * 128;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              float4 v;
              v.x = round_tf32(wide_activate<HACT, true>(__uint_as_float(r[4 * i]) + b[4 * i]));
              v.y = round_tf32(wide_activate<HACT, true>(__uint_as_float(r[4 * i + 1]) + b[4 * i + 1]));
              v.z = round_tf32(wide_activate<HACT, true>(__uint_as_float(r[4 * i + 2]) + b[4 * i + 2]));
              v.w = round_tf32(wide_activate<HACT, true>(__uint_as_float(r[4 * i + 3]) + b[4 * i + 3]));
              *reinterpret_cast<float4*>(dst + 128 * i) = v;
            }
          } else {
            uint32_t o[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
              o[i] = pack_bf16x2(wide_activate<HACT, false>(__uint_as_float(r[2 * i]) + b[2 * i]),
                                 wide_activate<HACT, false>(__uint_as_float(r[2 * i + 1]) + b[2 * i + 1]));
            uint8_t* dst = out_blk + (size_t)(n0 >> 6) * kWideBlk + (size_t)((n0 & 63) >> 3) * 128;
            *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
            *reinterpret_cast<uint4*>(dst + 128) = make_uint4(o[4], o[5], o[6], o[7]);
          }
        };
        for (int c16 = cb; c16 < ce; c16 += 2) {
          tmem_wait_ld();
          if (c16 + 1 < ce) tmem_ld16(d_addr + 16 * (c16 + 1), rb);
          finish(c16, ra);
          if (c16 + 1 < ce) {
            tmem_wait_ld();
            if (c16 + 2 < ce) tmem_ld16(d_addr + 16 * (c16 + 2), ra);
            finish(c16 + 1, rb);
          }
        }
      } else {
        // head: (acc + b) * s + shift into the tile-major fp32 scratch (bnn.py:878-893)
        const long long gblk = p.head_blk0 + blk;
        const bool live_row = p.row0 + blk * kTileRows + row_in_tile < n_rows_all;
        float* out = p.heads + ((size_t)(gblk * p.E + e) * p.n_head) * kTileRows + row_in_tile;
        for (int c16 = cb; c16 < ce; ++c16) {
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 16; ++i) rb[i] = ra[i];
          if (c16 + 1 < ce) tmem_ld16(d_addr + 16 * (c16 + 1), ra);
          if (live_row) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int col = 16 * c16 + i;
              if (col < p.n_head) {
                const float sc = p.head_scale ? __ldg(p.head_scale + col) : 1.f;
                const float sh = p.head_shift ? __ldg(p.head_shift + col) : 0.f;
                out[(size_t)col * kTileRows] = fmaf(__uint_as_float(rb[i]) + __ldg(bias + col), sc, sh);
              }
            }
          }
        }
      }
      tmem_wait_ld();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster_light(bar(WB_ACC_EMPTY + acc), 0);
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync();
  if (warp == 2) tmem_dealloc<2>(tmem_base, 512);
}

// ---------------------------------------------------------------------------------------
// weight packing for the wide path: fp32 W [E][K][N], b [E][1][N] -> bf16 blocks
//   [e][n chunk c][cta half h][k chunk j] of kWideBlk bytes, element (nl, kk) of a block at
//   (nl/8)*1024 + (kk/8)*128 + (nl%8)*16 + (kk%8)*2, and the fp32 bias table [E][npad];
// both pre-scaled by `prescale` (0.5 for swish / sigmoid: the epilogue works on x/2).
// ---------------------------------------------------------------------------------------
__global__ void pack_wide_layer_kernel(const float* __restrict__ W, const float* __restrict__ b, uint8_t* img,
                                       float* bias_tab, int E, WideLayerPlan lp, int tf32) {
  const long long per_e = (long long)lp.npad * lp.kpad;
  const long long total = (long long)E * per_e;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int e = (int)(idx / per_e);
    const long long rem = idx - (long long)e * per_e;
    const int n = (int)(rem / lp.kpad), k = (int)(rem - (long long)n * lp.kpad);
    const int c = n / kWideNC, n_in = n - c * kWideNC;
    const int chunk_n = min(kWideNC, lp.npad - c * kWideNC), nh = chunk_n >> 1;
    const int h = n_in / nh, nl = n_in - h * nh;
    const int j = k / lp.kc, kk = k - j * lp.kc;
    float v = 0.f;
    if (lp.split3) {   // rows [W_hi ; W_hi ; W_lo] against the staged columns [x_hi | x_lo | x_hi]
      const int part = k / lp.k_real, ks = k - part * lp.k_real;
      if (n < lp.n_real && part < 3) {
        const float w = W[((size_t)e * lp.k_real + ks) * lp.n_real + n] * lp.prescale;
        const float hi = sm100::round_tf32(w);
        v = part < 2 ? hi : w - hi;
      }
    } else if (n < lp.n_real && k < lp.k_real) {
      v = W[((size_t)e * lp.k_real + k) * lp.n_real + n] * lp.prescale;
    }
    const size_t blk = ((((size_t)e * lp.n_chunks + c) * 2 + h) * lp.k_chunks + j) * kWideBlk;
    if (tf32) {   // fp32 words, 8x4 core matrices (tcgen05.mma kind::tf32)
      const size_t off = blk + (size_t)(nl >> 3) * 1024 + (size_t)(kk >> 2) * 128 + (size_t)(nl & 7) * 16 + (size_t)(kk & 3) * 4;
      *reinterpret_cast<float*>(img + off) = sm100::round_tf32(v);
    } else {
      const size_t off = blk + (size_t)(nl >> 3) * 1024 + (size_t)(kk >> 3) * 128 + (size_t)(nl & 7) * 16 + (size_t)(kk & 7) * 2;
      *reinterpret_cast<__nv_bfloat16*>(img + off) = __float2bfloat16_rn(v);
    }
    if (k == 0) bias_tab[(size_t)e * lp.npad + n] = (n < lp.n_real && b) ? b[(size_t)e * lp.n_real + n] * lp.prescale : 0.f;
  }
}

cudaError_t pack_wide_layer(const float* W, const float* b, uint8_t* img, float* bias_tab, int E,
                            const WideLayerPlan& lp, int tf32, cudaStream_t stream) {
  const long long total = (long long)E * lp.npad * lp.kpad;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 8192) blocks = 8192;
  pack_wide_layer_kernel<<<blocks, 256, 0, stream>>>(W, b, img, bias_tab, E, lp, tf32);
  return cudaGetLastError();
}

template <int HACT, bool TF32>
static cudaError_t launch_wide_variant2(const WideParams& p, int num_sms, cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(wide_layer_kernel<HACT, TF32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kWideSmemBytes);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  int clusters = num_sms / 2;
  if (!p.n_rows_dev) {
    const long long n_tiles = (p.slab_rows + 2 * kTileRows - 1) / (2 * kTileRows);
    const long long items = n_tiles * p.E * p.lp.n_chunks;
    if (items < clusters) clusters = (int)(items > 0 ? items : 1);
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(kWideThreads);
  cfg.dynamicSmemBytes = kWideSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, wide_layer_kernel<HACT, TF32>, p);
}
template <int HACT>
static cudaError_t launch_wide_variant(const WideParams& p, int num_sms, cudaStream_t stream) {
  return p.tf32 ? launch_wide_variant2<HACT, true>(p, num_sms, stream) : launch_wide_variant2<HACT, false>(p, num_sms, stream);
}

// act: activation of THIS layer (ACT_NONE for the head)
cudaError_t launch_wide_layer(const WideParams& p, int act, int num_sms, cudaStream_t stream) {
  switch (act) {
    case ACT_SWISH:   return launch_wide_variant<ACT_SWISH>(p, num_sms, stream);
    case ACT_RELU:    return launch_wide_variant<ACT_RELU>(p, num_sms, stream);
    case ACT_TANH:    return launch_wide_variant<ACT_TANH>(p, num_sms, stream);
    case ACT_SIGMOID: return launch_wide_variant<ACT_SIGMOID>(p, num_sms, stream);
    case ACT_NONE:    return launch_wide_variant<ACT_NONE>(p, num_sms, stream);
    default:          return cudaErrorInvalidValue;
  }
}

}  // namespace cmbpo
