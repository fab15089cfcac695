// ensemble_fwd.cu -- fused PETS-ensemble MLP forward for sm_100a.
//
// Replaces, for the rollout step, the TF graph the reference builds in
// BNN._compile_outputs (mbpo/models/bnn.py:852-903): scaler_in.transform (utils.py:161) ->
// (L+1) x FC.compute_output_tensor (fc.py:105-114) -> head split + inverse scalers
// (bnn.py:878-893, utils.py:172,192).  exp() of the log-variance is left to the consumer
// kernels (step_post.cu) which need the log-variance anyway.
//
// Shape of the kernel (see DESIGN.md "K1"):
//   * persistent CTA *pairs* (cluster of 2, tcgen05 cta_group::2).  A pair owns a 256-row
//     super-tile; each CTA holds its own 128 rows (TMEM lanes) and HALF of every layer's
//     weight matrix (N split), so one member's whole bf16 weight set (~150 KB per CTA at
//     4x200) is resident in shared memory and refilled layer by layer by 1-D bulk async
//     copies (TMA engine) as soon as the tensor pipe releases a slot.
//   * activations never leave the SM: layer l's fp32 accumulator (TMEM) is read by the
//     epilogue warps, bias is already in the GEMM (two extra K rows hold bias hi/lo against
//     two constant-one A columns), swish is evaluated as g + g*tanh(g) on g = x/2 (weights are
//     pre-scaled by 1/2, exact), and the bf16 result is written back IN PLACE over the
//     accumulator's own columns, where the next layer's tcgen05.mma reads it as its TMEM
//     A operand, chunk by chunk (mbarrier per 32 columns) while the epilogue is still running.
//   * warp roles: 0 = weight loader, 1 = MMA issuer (leader CTA), 2 = TMEM alloc + peer->leader
//     relay of weight arrivals, 3 = input stager (concat + scaler + bf16 + canonical smem
//     layout), 4..11 = epilogue (two warps per TMEM lane quadrant, alternating 32-column chunks).
#include "cmbpo_internal.cuh"
#include "sm100_ptx.cuh"

namespace cmbpo {
using namespace sm100;

constexpr int kFwdThreads = 384;   // warps 0-3: loader / MMA issuer / alloc+relay / input stager; 4-11: epilogue

struct SmemLayout {
  uint32_t off_a0, off_const, off_bar, off_tmem, total;
};
__host__ __device__ inline SmemLayout make_smem_layout(const ModelPlan& pl) {
  SmemLayout s;
  s.off_a0 = pl.w_smem_bytes;
  s.off_const = s.off_a0 + 2 * pl.a0_bytes;
  uint32_t nconst = 2u * pl.in_dim + 2u * pl.n_head;
  s.off_bar = (s.off_const + 4u * nconst + 15u) & ~15u;
  s.off_tmem = s.off_bar + 40u * 8u;
  s.total = s.off_tmem + 16u;
  return s;
}
uint32_t fwd_smem_bytes(const ModelPlan& pl) { return make_smem_layout(pl).total; }

// barrier indices inside the barrier block
constexpr int BAR_W_FULL = 0;            // [kMaxLayers]
constexpr int BAR_W_EMPTY = 8;           // [kMaxLayers]
constexpr int BAR_A0_FULL = 16;          // [2]
constexpr int BAR_A0_EMPTY = 18;         // [2]
constexpr int BAR_D_FULL = 20;           // [1]
constexpr int BAR_A_READY = 24;          // [8] one per 32-column chunk

enum WaitCode : int {
  WC_LOADER_EMPTY = 1, WC_RELAY_FULL = 2, WC_STAGER_EMPTY = 3, WC_MMA_WFULL = 4,
  WC_MMA_A0 = 5, WC_MMA_AREADY = 6, WC_EPI_DFULL = 7
};

// debug timeline: role r in [0,4) owns trace[r*2048 ..]; entries are (clock64, tag) pairs
constexpr int kTraceSlots = 1024;
struct Tracer {
  long long* base; int n;
  __device__ __forceinline__ void operator()(int tag) {
    if (base && n < kTraceSlots) { base[2 * n] = clock64(); base[2 * n + 1] = tag; ++n; }
  }
};

template <int ACT>
__device__ __forceinline__ float activate(float g) {
  // g is the pre-activation, already halved for ACT_SWISH / ACT_SIGMOID
  if constexpr (ACT == ACT_SWISH) {        // x*sigmoid(x) = g + g*tanh(g), g = x/2   (fc.py:20)
    return fmaf(g, tanh_approx(g), g);
  } else if constexpr (ACT == ACT_RELU) {
    return fmaxf(g, 0.f);
  } else if constexpr (ACT == ACT_TANH) {
    return tanh_approx(g);
  } else if constexpr (ACT == ACT_SIGMOID) {  // 0.5 + 0.5*tanh(x/2)
    return fmaf(0.5f, tanh_approx(g), 0.5f);
  } else {
    return g;
  }
}

// Convert one accumulator chunk (W fp32 columns starting at col0) to packed bf16 in registers.
template <int ACT, int W>
__device__ __forceinline__ void convert_chunk(const uint32_t (&r)[32], uint32_t (&o)[16], int col0,
                                              int ones_col) {
  if (ones_col >= col0 - 1 && ones_col < col0 + W) {
    // the chunk that holds the next layer's two bias-carrier columns (warp-uniform, once per layer)
#pragma unroll
    for (int i = 0; i < W / 2; ++i) {
      float a = activate<ACT>(__uint_as_float(r[2 * i]));
      float b = activate<ACT>(__uint_as_float(r[2 * i + 1]));
      if (col0 + 2 * i == ones_col || col0 + 2 * i == ones_col + 1) a = 1.0f;
      if (col0 + 2 * i + 1 == ones_col || col0 + 2 * i + 1 == ones_col + 1) b = 1.0f;
      o[i] = pack_bf16x2(a, b);
    }
  } else {
#pragma unroll
    for (int i = 0; i < W / 2; ++i)
      o[i] = pack_bf16x2(activate<ACT>(__uint_as_float(r[2 * i])), activate<ACT>(__uint_as_float(r[2 * i + 1])));
  }
}

__device__ __forceinline__ void signal_chunk(uint32_t bar, int lane) {
  tmem_wait_st();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive_cluster_light(bar, 0);
}

// Hidden-layer epilogue for one warp: accumulator -> activation -> bf16 -> same columns, signalling
// the MMA issuer per 32-column chunk.  Two warps share each TMEM lane quadrant (`half` = 0/1 takes the
// even / odd chunks) so that every SM sub-partition has two instruction streams: the MUFU.TANH ->
// FFMA dependency stalls of one are filled by the other, and the XU pipe (16 tanh/clk/SM -- the
// binding unit of this kernel) stays busy.  The TMEM load of the warp's next chunk is in flight while
// the current one is converted.
template <int ACT>
__device__ __forceinline__ void hidden_epilogue(uint32_t d_addr, int npad, int ones_col,
                                                uint32_t bar_a_ready0, int lane, int half, Tracer& tr) {
  const int nfull = npad >> 5;                       // 32-wide chunks
  const int nch = nfull + ((npad & 31) ? 1 : 0);     // + one 16-wide chunk at the end
  uint32_t ra[32], rb[32], o[16];
  auto load = [&](int c, uint32_t (&r)[32]) {
    if (c < nfull) tmem_ld32(d_addr + 32 * c, r); else tmem_ld16_lo(d_addr + 32 * c, r);
  };
  auto finish = [&](int c, uint32_t (&r)[32]) {
    if (c < nfull) { convert_chunk<ACT, 32>(r, o, 32 * c, ones_col); tmem_st16(d_addr + 16 * c, o); }
    else { convert_chunk<ACT, 16>(r, o, 32 * c, ones_col); tmem_st8_lo(d_addr + 16 * c, o); }
    signal_chunk(bar_a_ready0 + 8 * c, lane);
  };
  int c = half;
  if (c < nch) load(c, ra);
  for (; c < nch; c += 4) {
    tr(300);
    tmem_wait_ld();
    if (c + 2 < nch) load(c + 2, rb);
    finish(c, ra);
    tr(302);
    if (c + 2 < nch) {
      tmem_wait_ld();
      if (c + 4 < nch) load(c + 4, ra);
      finish(c + 2, rb);
    }
  }
}

// HACT = activation of every hidden layer (construct_model applies one `activation` to all of
// them, constructor.py:57-60); a kernel template so each variant gets its own register allocation.
template <int HACT>
__global__ void __launch_bounds__(kFwdThreads, 1) ensemble_fwd_kernel(const __grid_constant__ FwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const ModelPlan& pl = p.plan;
  const SmemLayout sl = make_smem_layout(pl);
  const int tid = threadIdx.x, lane = tid & 31;
  // warp index through a shuffle so the compiler knows it is warp-uniform: role branches stay
  // convergent and the MMA operands live in uniform registers (no per-instruction R2UR waterfall)
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const uint32_t cta = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const long long n_rows = p.n_rows_dev ? (long long)(*p.n_rows_dev) : p.n_rows;
  const int n_tiles = (int)((n_rows + 2 * kTileRows - 1) / (2 * kTileRows));
  const int E = pl.E, L = pl.L;

  float* s_const = reinterpret_cast<float*>(smem + sl.off_const);
  float* s_in_mu = s_const;
  float* s_in_inv = s_const + pl.in_dim;
  float* s_head_scale = s_const + 2 * pl.in_dim;
  float* s_head_shift = s_const + 2 * pl.in_dim + pl.n_head;
  const uint32_t bar0 = smem_u32(smem + sl.off_bar);
  auto bar = [bar0](int i) -> uint32_t { return bar0 + 8u * (uint32_t)i; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + sl.off_tmem);

  // ---------------------------------------------------------------- setup
  if (tid == 0) {
    for (int l = 0; l < kMaxLayers; ++l) {
      mbar_init(bar(BAR_W_FULL + l), cta == 0 ? 2 : 1);
      mbar_init(bar(BAR_W_EMPTY + l), 1);
      mbar_init(bar(BAR_A_READY + l), 8);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar(BAR_A0_FULL + b), 2);
      mbar_init(bar(BAR_A0_EMPTY + b), 1);
    }
    mbar_init(bar(BAR_D_FULL), 1);
    fence_mbar_init();
  }
  for (int i = tid; i < pl.in_dim; i += blockDim.x) {
    s_in_mu[i] = p.in_mu ? p.in_mu[i] : 0.f;
    s_in_inv[i] = p.in_inv ? p.in_inv[i] : 1.f;
  }
  for (int i = tid; i < pl.n_head; i += blockDim.x) {
    s_head_scale[i] = p.head_scale ? p.head_scale[i] : 1.f;
    s_head_shift[i] = p.head_shift ? p.head_shift[i] : 0.f;
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

  // ---------------------------------------------------------------- roles
  if (warp == 0) {
    // ===== weight loader: refill layer slot l with the next (tile, member) as soon as free
    if (lane == 0) {
      uint32_t unit = 0;
      for (int tile = cluster_id; tile < n_tiles; tile += n_clusters) {
        for (int e = 0; e < E; ++e, ++unit) {
          const int member = p.member_of_slot[e];
          const uint8_t* blk = p.wimg + (size_t)(member * 2 + (int)cta) * pl.blk_bytes;
          for (int l = 0; l < L; ++l) {
            const LayerPlan& lp = pl.layer[l];
            if (unit > 0) mbar_wait_relaxed(bar(BAR_W_EMPTY + l), (unit - 1) & 1, p.status, WC_LOADER_EMPTY);
            mbar_arrive_expect_tx(bar(BAR_W_FULL + l), lp.bytes);
            bulk_g2s(smem_u32(smem + lp.smem_off), blk + lp.img_off, lp.bytes, bar(BAR_W_FULL + l));
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA).  The whole warp walks the loop convergently (all lanes wait on the
    // barriers); only the tcgen05.mma / commit instructions are issued by one elected lane.
    if (cta == 0) {
      uint32_t unit = 0, n = 0, stage = 0, ar_bits = 0;
      Tracer tr{(p.trace && cluster_id == 0 && lane == 0) ? p.trace : nullptr, 0};
      for (int tile = cluster_id; tile < n_tiles; tile += n_clusters) {
        for (int e = 0; e < E; ++e, ++unit) {
          for (int l = 0; l < L; ++l, ++n) {
            tr(100 + l);
            const LayerPlan& lp = pl.layer[l];
            const uint32_t d_col = (n & 1) * kRegionCols;
            const uint32_t a_col = ((n & 1) ^ 1) * kRegionCols;
            const uint32_t idesc = make_idesc(256, lp.npad, kFmtBF16);
            const uint32_t b_base = smem_u32(smem + lp.smem_off);
            const uint32_t sbo = (uint32_t)(lp.kpad >> 3) * 128u;
            const int ksteps = lp.kpad >> 4;
            mbar_wait_cluster(bar(BAR_W_FULL + l), unit & 1, p.status, WC_MMA_WFULL);
            tr(110 + l);
            if (l == 0) {
              const uint32_t buf = stage & 1;
              if (p.per_slot_input || e == 0)
                mbar_wait_cluster(bar(BAR_A0_FULL + buf), (stage >> 1) & 1, p.status, WC_MMA_A0);
              tc_fence_after();
              const uint32_t a_base = smem_u32(smem + sl.off_a0 + buf * pl.a0_bytes);
              if (elect_one()) {
                for (int j = 0; j < ksteps; ++j)
                  mma_bf16_ss<2>(tmem_base + d_col, make_smem_desc(a_base + j * 256, 128, sbo),
                                 make_smem_desc(b_base + j * 256, 128, sbo), idesc, j > 0);
              }
              __syncwarp();
              if (p.per_slot_input || e == E - 1) {
                if (elect_one()) mma_commit_2sm(bar(BAR_A0_EMPTY + buf), 0x3);
                ++stage;
              }
            } else {
              const int nch = (lp.kpad + 31) >> 5;
              for (int c = 0; c < nch; ++c) {
                mbar_wait_cluster(bar(BAR_A_READY + c), (ar_bits >> c) & 1, p.status, WC_MMA_AREADY);
                ar_bits ^= (1u << c);
                tc_fence_after();
                tr(120 + c);
                const int j1 = min(2 * c + 2, ksteps);
                if (elect_one()) {
                  for (int j = 2 * c; j < j1; ++j)
                    mma_bf16_ts<2>(tmem_base + d_col, tmem_base + a_col + 8 * j,
                                   make_smem_desc(b_base + j * 256, 128, sbo), idesc, j > 0);
                }
                __syncwarp();
              }
            }
            if (elect_one()) {
              mma_commit_2sm(bar(BAR_W_EMPTY + l), 0x3);
              mma_commit_2sm(bar(BAR_D_FULL), 0x3);
            }
            __syncwarp();
            tr(130 + l);
          }
        }
      }
    }
  } else if (warp == 2) {
    // ===== relay (peer CTA): forward "my half of layer l has landed" to the leader's barrier
    if (cta == 1 && lane == 0) {
      uint32_t unit = 0;
      for (int tile = cluster_id; tile < n_tiles; tile += n_clusters)
        for (int e = 0; e < E; ++e, ++unit)
          for (int l = 0; l < L; ++l) {
            mbar_wait_relaxed(bar(BAR_W_FULL + l), unit & 1, p.status, WC_RELAY_FULL);
            mbar_arrive_cluster(bar(BAR_W_FULL + l), 0);
          }
    }
  } else if (warp == 3) {
    // ===== input stager: concat(segments) -> (x-mu)*inv -> bf16 -> canonical K-major smem tile
    const LayerPlan& l0 = pl.layer[0];
    const uint32_t sbo = (uint32_t)(l0.kpad >> 3) * 128u;
    uint32_t stage = 0;
    for (int tile = cluster_id; tile < n_tiles; tile += n_clusters) {
      const int n_stage = p.per_slot_input ? E : 1;
      for (int e = 0; e < n_stage; ++e, ++stage) {
        const uint32_t buf = stage & 1;
        if (stage >= 2) mbar_wait_relaxed(bar(BAR_A0_EMPTY + buf), ((stage >> 1) + 1) & 1, p.status, WC_STAGER_EMPTY);
        uint8_t* a0 = smem + sl.off_a0 + buf * pl.a0_bytes;
        const int in_slot = p.per_slot_input ? p.input_of_slot[e] : 0;
        for (int i = 0; i < 4; ++i) {
          const int r = lane + 32 * i;
          const long long g = (long long)tile * (2 * kTileRows) + (long long)cta * kTileRows + r;
          uint8_t* rowp = a0 + (uint32_t)(r >> 3) * sbo + (uint32_t)(r & 7) * 16u;
          const bool live = g < n_rows;
          int k = 0;
          for (int s = 0; s < kMaxSeg; ++s) {
            const int w = p.seg_width[s];
            if (w == 0) continue;
            const float* src = p.seg_ptr[s] + (size_t)in_slot * p.seg_slot_stride[s] + (size_t)g * w;
            for (int kk = 0; kk < w; ++kk, ++k) {
              float x = live ? __ldg(src + kk) : 0.f;
              float xn = (x - s_in_mu[k]) * s_in_inv[k];
              *reinterpret_cast<__nv_bfloat16*>(rowp + (uint32_t)(k >> 3) * 128u + (uint32_t)(k & 7) * 2u) =
                  __float2bfloat16_rn(live ? xn : 0.f);
            }
          }
          for (; k < l0.kpad; ++k) {  // bias carrier columns, then zero padding
            float v = (k < pl.in_dim + 2) ? 1.0f : 0.0f;
            *reinterpret_cast<__nv_bfloat16*>(rowp + (uint32_t)(k >> 3) * 128u + (uint32_t)(k & 7) * 2u) =
                __float2bfloat16_rn(v);
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(bar(BAR_A0_FULL + buf), 0);
      }
    }
  } else {
    // ===== epilogue warps 4..11: TMEM lane quadrant q (the hardware ties a warp to lanes 32*(warp%4)..),
    // column half `half`
    const int q = warp & 3;
    const int half = (warp - 4) >> 2;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const int row_in_tile = q * 32 + lane;
    uint32_t n = 0;
    Tracer tr{(p.trace && cluster_id == 0 && cta == 0 && warp == 4 && lane == 0) ? p.trace + 2 * kTraceSlots : nullptr, 0};
    for (int tile = cluster_id; tile < n_tiles; tile += n_clusters) {
      const long long grow = (long long)tile * (2 * kTileRows) + (long long)cta * kTileRows + row_in_tile;
      for (int e = 0; e < E; ++e) {
        for (int l = 0; l < L; ++l, ++n) {
          const LayerPlan& lp = pl.layer[l];
          const uint32_t d_addr = tmem_base + lane_base + (n & 1) * kRegionCols;
          tr(200 + l);
          mbar_wait(bar(BAR_D_FULL), n & 1, p.status, WC_EPI_DFULL);
          tc_fence_after();
          tr(210 + l);
          if (l < L - 1) {
            const int ones_col = lp.n_real;  // next layer's bias-carrier columns
            const uint32_t ar0 = bar(BAR_A_READY);
            hidden_epilogue<HACT>(d_addr, lp.npad, ones_col, ar0, lane, half, tr);
            tr(220 + l);
          } else {
            // head: inverse output scaler on the mean, 2*log(s) shift on the log-variance
            // (bnn.py:878-893); written as [slot][column][row] so the store is coalesced.
            float* out = p.heads + (size_t)e * pl.n_head * p.row_stride + grow;
            for (int c0 = 16 * half; c0 < lp.npad; c0 += 32) {
              uint32_t r[16];
              tmem_ld16(d_addr + c0, r);
              tmem_wait_ld();
              if (grow < n_rows) {
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                  const int j = c0 + i;
                  if (j < pl.n_head)
                    out[(size_t)j * p.row_stride] = fmaf(__uint_as_float(r[i]), s_head_scale[j], s_head_shift[j]);
                }
              }
            }
          }
        }
      }
    }
  }

  // ---------------------------------------------------------------- teardown
  tc_fence_before();
  __syncthreads();
  cluster_sync();
  if (warp == 2) tmem_dealloc<2>(tmem_base, 512);
}

// ---------------------------------------------------------------------------------------
// weight packing: fp32 [E][K][N] (+ bias [E][1][N]) -> bf16 canonical K-major image per
// (member, N-half):  element (n, k) of the half at
//     (n/8)*(kpad/8)*128 + (k/8)*128 + (n%8)*16 + (k%8)*2
// rows k_real, k_real+1 of K carry bias hi / lo; everything else in the padding is zero.
// ---------------------------------------------------------------------------------------
__global__ void pack_layer_kernel(const float* __restrict__ W, const float* __restrict__ b, uint8_t* img,
                                  int E, LayerPlan lp, uint32_t blk_bytes) {
  const int nhalf = lp.npad >> 1;
  const long long per_blk = (long long)nhalf * lp.kpad;
  const long long total = (long long)E * 2 * per_blk;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(idx % lp.kpad);
    const int nl = (int)((idx / lp.kpad) % nhalf);
    const int h = (int)((idx / per_blk) & 1);
    const int e = (int)(idx / (2 * per_blk));
    const int n = h * nhalf + nl;
    float v = 0.f;
    if (n < lp.n_real) {
      if (k < lp.k_real) {
        v = W[((size_t)e * lp.k_real + k) * lp.n_real + n] * lp.prescale;
      } else if (k == lp.k_real || k == lp.k_real + 1) {
        const float bias = (b ? b[(size_t)e * lp.n_real + n] : 0.f) * lp.prescale;
        const float hi = __bfloat162float(__float2bfloat16_rn(bias));
        v = (k == lp.k_real) ? hi : (bias - hi);
      }
    }
    const size_t off = (size_t)(e * 2 + h) * blk_bytes + lp.img_off +
                       (size_t)(nl >> 3) * ((size_t)(lp.kpad >> 3) * 128) + (size_t)(k >> 3) * 128 +
                       (size_t)(nl & 7) * 16 + (size_t)(k & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(img + off) = __float2bfloat16_rn(v);
  }
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
cudaError_t pack_layer(const float* W, const float* b, uint8_t* img, int E, const LayerPlan& lp,
                       uint32_t blk_bytes, cudaStream_t stream) {
  const long long total = (long long)E * lp.npad * lp.kpad;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 4096) blocks = 4096;
  pack_layer_kernel<<<blocks, 256, 0, stream>>>(W, b, img, E, lp, blk_bytes);
  return cudaGetLastError();
}

template <int HACT>
static cudaError_t launch_fwd_variant(const FwdParams& p, int num_sms, cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(ensemble_fwd_kernel<HACT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         227 * 1024);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const long long n_tiles = (p.n_rows + 2 * kTileRows - 1) / (2 * kTileRows);
  int clusters = num_sms / 2;
  if (!p.n_rows_dev && n_tiles < clusters) clusters = (int)(n_tiles > 0 ? n_tiles : 1);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(kFwdThreads);
  cfg.dynamicSmemBytes = p.plan.smem_total;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, ensemble_fwd_kernel<HACT>, p);
}

// hidden_act: the (single) activation of the hidden layers
cudaError_t launch_ensemble_fwd(const FwdParams& p, int hidden_act, int num_sms, cudaStream_t stream) {
  switch (hidden_act) {
    case ACT_SWISH:   return launch_fwd_variant<ACT_SWISH>(p, num_sms, stream);
    case ACT_RELU:    return launch_fwd_variant<ACT_RELU>(p, num_sms, stream);
    case ACT_TANH:    return launch_fwd_variant<ACT_TANH>(p, num_sms, stream);
    case ACT_SIGMOID: return launch_fwd_variant<ACT_SIGMOID>(p, num_sms, stream);
    case ACT_NONE:    return launch_fwd_variant<ACT_NONE>(p, num_sms, stream);
    default:          return cudaErrorInvalidValue;
  }
}

}  // namespace cmbpo
