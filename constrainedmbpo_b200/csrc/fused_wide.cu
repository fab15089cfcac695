// fused_wide.cu -- whole-stack PETS-ensemble forward for nets whose hidden layers are too wide for the two-stream
// fused kernel (ensemble_fwd.cu, <= 254 columns) but still fit ONE shared-memory activation tile: widths up to 512
// columns in bf16 -- the (512, 512) nets every shipped CMBPO config uses (examples/config/antsafe/cmbpo_uc.py:36,
// mbpo/models/constructor.py:9-67) -- and up to 256 columns with tf32 operands.
//
// Why: the layer-at-a-time kernel (wide_fwd.cu) streams BOTH operands of every (tile, N chunk) item through shared
// memory (64 B per clock and SM at full tensor rate, above what L2 delivers) and round-trips the bf16 activations of
// E members through HBM between layers.  Here the activations of a 256-row tile never leave the SM pair:
//
//   * A: one activation tile per CTA (128 rows x up to 512 columns, "chunked canonical" 16 KB blocks of
//     cmbpo_internal.cuh), written by the epilogue of layer l and read by the MMAs of layer l + 1; the staged input
//     tile of layer 0 (K0's image) has a buffer of its own and is reused by all members of the tile.
//   * B: the same packed weight image as wide_fwd.cu, streamed in 16 KB blocks (this CTA's N half x 64 k) through a
//     ring of bulk async copies -- the only operand traffic left: 32 B per clock and SM.
//   * D: the layer's fp32 accumulator in TMEM, up to two 256-column N chunks (columns 0..511); the head uses chunk 0.
//
// Schedule of one (tile, member) instance, tcgen05 cta_group::2, M = 256 rows per CTA pair:
//   MMAs of layer l (N chunk 0, then 1; K blocks in order) -> commit LAYER_DONE -> the 16 epilogue warps convert the
//   accumulator in block groups of 128 columns (two K blocks of the next layer): bias + activation + bf16 pack + 16-byte
//   shared-memory stores, then fence.proxy.async and an arrival on A_RDY[group].  The issuer starts N chunk 0 of layer
//   l + 1 as soon as the conversion has left the chunk's accumulator columns and then follows it group by group.  In a
//   two-chunk layer the conversion of chunk 0 does not wait for the layer: it starts on ACC0_FULL and overwrites block
//   group g of the activation tile once chunk 1's MMAs have read it (A_FREE[g]), so half of every conversion runs under
//   MMAs of the same layer and the other half under the first chunk of the next one.
//   The accumulation order per output element (K blocks in order, 4 k-steps each) and the epilogue arithmetic are
//   those of wide_fwd.cu: the two kernels agree bit for bit (tests/test_gpu_wide.py).
//
// Warp roles: 0 = weight loader (+ relay of the peer CTA's arrivals to the leader's barrier), 1 = MMA issuer (leader
// CTA), 2 = TMEM alloc, 3 = input-tile loader, 4..19 = epilogue (TMEM lane quadrant = warp % 4, column quarter of every
// block group = (warp - 4) / 4).  Work split: the flat list of (tile, member) instances, member fastest, is cut into one
// contiguous range per CTA pair.  Nets whose layers are all one N chunk run two row tiles per instance (see NS below).
// Measurements and what bounds the kernel: profiles/r02_fused_wide_trace_notes.txt.
#include "cmbpo_internal.cuh"
#include "sm100_ptx.cuh"

#ifdef FW_EXP_TESTPOLL   // timing experiment: non-blocking probes (mbarrier.test_wait) in the loader / relay loops -- no difference measured
#define FW_POLL mbar_test_wait
#else
#define FW_POLL mbar_try_wait
#endif
#ifdef FW_EXP_PLAINWAIT   // timing experiment: spin-waiting epilogue warps
#define FW_EPI_WAIT mbar_wait
#else
#define FW_EPI_WAIT mbar_wait_parked
#endif

namespace cmbpo {
using namespace sm100;

constexpr int kFwThreads = 640;
constexpr int kFwEpiWarps = 16;

constexpr int FB_W_FULL = 0;         // [kFusedWideMaxStages]
constexpr int FB_W_EMPTY = 8;        // [kFusedWideMaxStages]
constexpr int FB_A_RDY = 16;         // [kFusedWideMaxABlocks / kFwGroup] blocks of group g of the activation tile written (and every accumulator column below their end read)
#ifndef FW_GROUP
#define FW_GROUP 2
#endif
constexpr int kFwGroup = FW_GROUP;          // K blocks per A_RDY signal: one fence.proxy.async (300-400 cycles) per 128 bf16 columns and warp
constexpr int FB_A0_FULL = 24;
constexpr int FB_A0_EMPTY = 25;
constexpr int FB_LAYER_DONE = 26;    // all MMAs of a layer complete (accumulator full, A tile no longer read)
constexpr int FB_HEAD_DRAINED = 27;  // the head accumulator is in registers
constexpr int FB_ACC0_FULL = 28;     // two-chunk layers: N chunk 0 complete (its conversion starts under the MMAs of chunk 1)
constexpr int FB_A_FREE = 29;        // [kFwEarlyGroups] two-chunk layers: chunk 1's MMAs have read block group g of the activation tile
constexpr int kFwEarlyGroups = (kWideNC / kWideKC) / kFwGroup;    // block groups under N chunk 0 (256 bf16 columns)
static_assert(kFwGroup == 2 || kFwGroup == 4, "block groups of 2 or 4 K blocks");
constexpr int FB_S1 = 31;            // second row-tile stream: A0_FULL, A0_EMPTY, LAYER_DONE, HEAD_DRAINED at FB_S1 + 0 .. 3; its A_RDY at FB_A_RDY + 4 + g
constexpr int kFwBars = 35;
static_assert(kFwBars * 8 + 16 <= 512, "barrier area");
constexpr uint32_t kFwBarBytes = 512;   // barriers + TMEM slot, padded so the bias words behind them are 16-byte aligned
constexpr int kFwBiasFloats = 512;   // widest hidden layer (8 blocks x 64 bf16 columns)

// Debug timeline (-DCMBPO_TRACE, `make trace`): cluster 0's MMA issuer (role 0) and two epilogue warps of the leader CTA
// (role 1: quadrant 0 part 0, role 2: quadrant 0 part 3) append (clock64, tag) pairs; tools/fw_trace_report.py prints them.
constexpr int kFwTraceSlots = 2048;
#ifdef CMBPO_TRACE
#define FTR_DECL(role, cond) long long* tr_base = ((cond) && p.trace) ? p.trace + (role) * 2 * kFwTraceSlots : nullptr; int tr_n = 0
#define FTR(tag) do { if (tr_base && tr_n < kFwTraceSlots) { tr_base[2 * tr_n] = clock64(); tr_base[2 * tr_n + 1] = (tag); ++tr_n; } } while (0)
#else
#define FTR_DECL(role, cond)
#define FTR(tag)
#endif

enum FusedWideWaitCode : int { FWC_LOADER = 21, FWC_A0 = 22, FWC_MMA_W = 23, FWC_MMA_A = 24, FWC_MMA_A0 = 25, FWC_MMA_HEAD = 26, FWC_EPI = 27 };

__device__ __forceinline__ float fw_sigmoid_precise(float x) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.4426950408889634f * x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
  return r;
}
// same arithmetic as wide_fwd.cu's wide_activate (bit-identical results are tested)
template <int ACT, bool PRECISE>
__device__ __forceinline__ float fw_activate(float g) {
  if constexpr (PRECISE) {
    if constexpr (ACT == ACT_SWISH) return g * fw_sigmoid_precise(g);
    else if constexpr (ACT == ACT_RELU) return fmaxf(g, 0.f);
    else if constexpr (ACT == ACT_TANH) return fmaf(2.0f, fw_sigmoid_precise(2.0f * g), -1.0f);
    else if constexpr (ACT == ACT_SIGMOID) return fw_sigmoid_precise(g);
    else return g;
  }
#ifdef FW_EXP_NOTANH   // timing experiment: no MUFU work (results are wrong)
  return fmaf(g, g, g);
#endif
  if constexpr (ACT == ACT_SWISH) return fmaf(g, tanh_approx(g), g);
  else if constexpr (ACT == ACT_RELU) return fmaxf(g, 0.f);
  else if constexpr (ACT == ACT_TANH) return tanh_approx(g);
  else if constexpr (ACT == ACT_SIGMOID) return fmaf(0.5f, tanh_approx(g), 0.5f);
  else return g;
}

__device__ __forceinline__ int fw_member_of(const FusedWideParams& p, int j) {
  return p.perm_dev ? perm_at(p.perm_dev, j, p.E) : p.member_of_slot[j];
}
__device__ __forceinline__ int fw_input_of(const FusedWideParams& p, int j) {
  return p.perm_dev ? perm_at(p.perm_dev, perm_at(p.perm_dev, j, p.E), p.E) : p.input_of_slot[j];
}

// the instances [begin, end) of one CTA pair: i -> (tile = i / E, slot e = i % E)
struct FwInst {
  long long i, begin, end, tile;
  int e, E, per_slot;
  __device__ __forceinline__ FwInst(long long b, long long en, int E_, int per_slot_)
      : i(b), begin(b), end(en), tile(b / E_), e((int)(b % E_)), E(E_), per_slot(per_slot_) {}
  __device__ __forceinline__ bool valid() const { return i < end; }
  __device__ __forceinline__ void next() { ++i; if (++e == E) { e = 0; ++tile; } }
  __device__ __forceinline__ bool new_input() const { return per_slot || i == begin || e == 0; }       // the input tile changes
  __device__ __forceinline__ bool last_use() const { return per_slot || i == end - 1 || e == E - 1; }  // last instance on this input tile
};

template <int HACT, bool TF32>
__global__ void __launch_bounds__(kFwThreads, 1) fused_wide_kernel(const __grid_constant__ FusedWideParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const uint32_t cta = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const long long n_rows = p.n_rows_dev ? (long long)(*p.n_rows_dev) : p.n_rows;
  const long long n_tiles = (n_rows + 2 * kTileRows - 1) / (2 * kTileRows);
  const int E = p.E, L = p.L, S = p.stages;
  // Row-tile streams.  Nets whose layers are all ONE N chunk (<= 256 columns) leave half of TMEM and of the activation
  // area free: they run TWO row tiles per instance (unit = tile pair, stream s = tile 2 unit + s), each with its own
  // accumulator, activation tile, input tile and barriers, through the same weight ring -- the conversion of one stream's
  // layer runs under the other stream's MMAs (a one-chunk layer has nothing of its own to hide it behind).
  const int NS = p.streams;
  const long long n_units = (NS == 2) ? (n_tiles + 1) >> 1 : n_tiles;
  auto tile_of = [&](long long unit, int s) -> long long { return NS == 2 ? 2 * unit + s : unit; };
  auto streams_of = [&](long long unit) -> int { return (NS == 2 && 2 * unit + 1 < n_tiles) ? 2 : 1; };   // the last pair may be a lone tile
  const long long total = n_units * E;
  const long long i_begin = total * cluster_id / n_clusters, i_end = total * (cluster_id + 1) / n_clusters;

  const uint32_t a_bytes = (uint32_t)p.a_blocks * kWideBlk, a0_bytes = (uint32_t)p.a0_blocks * kWideBlk;
  const uint32_t off_a0 = (uint32_t)NS * a_bytes;
  const uint32_t off_ring = off_a0 + (uint32_t)NS * a0_bytes;
  const uint32_t off_bar = off_ring + (uint32_t)S * kWideBlk;
  const uint32_t bar0 = smem_u32(smem + off_bar);
  auto bar = [bar0](int i) -> uint32_t { return bar0 + 8u * (uint32_t)i; };
  auto bar_a_rdy = [&](int s, int g) -> uint32_t { return bar(FB_A_RDY + 4 * s + g); };
  auto bar_a0_full = [&](int s) -> uint32_t { return bar(s ? FB_S1 + 0 : FB_A0_FULL); };
  auto bar_a0_empty = [&](int s) -> uint32_t { return bar(s ? FB_S1 + 1 : FB_A0_EMPTY); };
  auto bar_layer_done = [&](int s) -> uint32_t { return bar(s ? FB_S1 + 2 : FB_LAYER_DONE); };
  auto bar_head_drained = [&](int s) -> uint32_t { return bar(s ? FB_S1 + 3 : FB_HEAD_DRAINED); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + off_bar + kFwBars * 8u);
  float* bias_s = reinterpret_cast<float*>(smem + off_bar + kFwBarBytes);   // [kFwBiasFloats] bias of the pass being converted

  if (tid == 0) {
    for (int s = 0; s < kFusedWideMaxStages; ++s) {
      mbar_init(bar(FB_W_FULL + s), cta == 0 ? 2 : 1);    // leader: own fill + relayed peer fill
      mbar_init(bar(FB_W_EMPTY + s), 1);                  // tcgen05.commit
    }
    for (int j = 0; j < kFusedWideMaxABlocks; ++j) mbar_init(bar(FB_A_RDY + j), 2 * kFwEpiWarps);
    for (int st = 0; st < 2; ++st) {
      mbar_init(bar_a0_full(st), cta == 0 ? 2 : 1);
      mbar_init(bar_a0_empty(st), 1);                     // tcgen05.commit
      mbar_init(bar_layer_done(st), 1);                   // tcgen05.commit
      mbar_init(bar_head_drained(st), 2 * kFwEpiWarps);
    }
    mbar_init(bar(FB_ACC0_FULL), 1);                      // tcgen05.commit
    for (int g = 0; g < kFwEarlyGroups; ++g) mbar_init(bar(FB_A_FREE + g), 1);   // tcgen05.commit
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
    // ===== weight loader: fill n of the (instance, layer, N chunk, K block) walk goes to ring stage n % S
    if (lane == 0) {
      long long per_inst = 0;
      for (int l = 0; l < L; ++l) per_inst += (long long)p.lp[l].n_chunks * p.lp[l].k_chunks;
      long long n_fill = 0;            // every stream of an instance walks the layer's blocks on its own
      for (FwInst ic(i_begin, i_end, E, p.per_slot_input); ic.valid(); ic.next()) n_fill += per_inst * streams_of(ic.tile);
      long long fill_n = 0, relay_n = 0;
      int f_stage = 0, r_stage = 0; uint32_t f_use = 0, r_use = 0, idle = 0;
      FwInst it(i_begin, i_end, E, p.per_slot_input);
      int l = 0, st = 0, c = 0, j = 0;
      const uint8_t* b_src = nullptr; uint32_t b_bytes = 0;
      while (fill_n < n_fill || (cta == 1 && relay_n < n_fill)) {
        bool progress = false;
        if (fill_n < n_fill && (f_use == 0 || FW_POLL(bar(FB_W_EMPTY + f_stage), (f_use - 1) & 1))) {
          const WideLayerPlan& lp = p.lp[l];
          if (j == 0) {
            const int chunk_n = min(kWideNC, lp.npad - c * kWideNC);
            b_bytes = (uint32_t)(chunk_n >> 1) * 128u;
#ifdef FW_EXP_HALFW   // timing experiment: half of every weight block (results are wrong)
            b_bytes >>= 1;
#endif
            b_src = p.w_img + lp.w_off + ((((size_t)fw_member_of(p, it.e) * lp.n_chunks + c) * 2 + cta) * lp.k_chunks) * kWideBlk;
          }
          mbar_arrive_expect_tx(bar(FB_W_FULL + f_stage), b_bytes);
          bulk_g2s(smem_u32(smem + off_ring + (uint32_t)f_stage * kWideBlk), b_src, b_bytes, bar(FB_W_FULL + f_stage));
          b_src += kWideBlk;
          if (++j == lp.k_chunks) { j = 0; if (++c == lp.n_chunks) { c = 0; if (++st == streams_of(it.tile)) { st = 0; if (++l == L) { l = 0; it.next(); } } } }
          if (++f_stage == S) { f_stage = 0; ++f_use; }
          ++fill_n; progress = true;
        }
        if (cta == 1 && relay_n < fill_n && FW_POLL(bar(FB_W_FULL + r_stage), r_use & 1)) {
          mbar_arrive_cluster_light(bar(FB_W_FULL + r_stage), 0);   // async-proxy data consumed by the async proxy: light arrive
          if (++r_stage == S) { r_stage = 0; ++r_use; }
          ++relay_n; progress = true;
        }
        if (progress) { idle = 0; continue; }
        __nanosleep(32);
        if (++idle > (SM100_WAIT_LIMIT >> 2)) {
          if (p.status) { *reinterpret_cast<volatile int*>(p.status) = (FWC_LOADER | 0x40000000); __threadfence_system(); }
          __trap();
        }
      }
    }
  } else if (warp == 3) {
    // ===== input-tile loader: K0's staged layer-0 operand blocks of this CTA's 128 rows, one fill per input tile
    if (lane == 0) {
      long long n_fill[2] = {0, 0}, fill_n[2] = {0, 0}, relay_n[2] = {0, 0};
      for (FwInst ic(i_begin, i_end, E, p.per_slot_input); ic.valid(); ic.next())
        if (ic.new_input()) { ++n_fill[0]; if (streams_of(ic.tile) == 2) ++n_fill[1]; }
      uint32_t idle = 0;
      FwInst it0(i_begin, i_end, E, p.per_slot_input), it1 = it0;       // next instance to stage, per stream
      while (fill_n[0] < n_fill[0] || fill_n[1] < n_fill[1] || (cta == 1 && (relay_n[0] < n_fill[0] || relay_n[1] < n_fill[1]))) {
        bool progress = false;
#pragma unroll
        for (int st = 0; st < 2; ++st) {
          FwInst& it = st ? it1 : it0;
          if (fill_n[st] < n_fill[st] && (fill_n[st] == 0 || FW_POLL(bar_a0_empty(st), (uint32_t)(fill_n[st] - 1) & 1))) {
            while (!(it.new_input() && st < streams_of(it.tile))) it.next();
            const int in_slot = p.per_slot_input ? fw_input_of(p, it.e) : 0;
            const uint8_t* src = p.a0_img + (size_t)in_slot * p.a0_slot_stride + (size_t)(tile_of(it.tile, st) * 2 + cta) * a0_bytes;
            mbar_arrive_expect_tx(bar_a0_full(st), a0_bytes);
            bulk_g2s(smem_u32(smem + off_a0 + st * a0_bytes), src, a0_bytes, bar_a0_full(st));
            it.next();
            ++fill_n[st]; progress = true;
          }
          if (cta == 1 && relay_n[st] < fill_n[st] && FW_POLL(bar_a0_full(st), (uint32_t)relay_n[st] & 1)) {
            mbar_arrive_cluster_light(bar_a0_full(st), 0);
            ++relay_n[st]; progress = true;
          }
        }
        if (progress) { idle = 0; continue; }
        __nanosleep(64);
        if (++idle > (SM100_WAIT_LIMIT >> 2)) {
          if (p.status) { *reinterpret_cast<volatile int*>(p.status) = (FWC_A0 | 0x40000000); __threadfence_system(); }
          __trap();
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA): the whole warp walks the loop, one lane issues.  A lone warp runs its dependent
    // instructions 4-6 cycles apart, so everything per K block that is not a wait, an MMA or a commit is kept out of the
    // inner loop (the first version spent ~650 cycles per 512-cycle K block and bounded the kernel: trace in
    // profiles/r02_fused_wide_trace_notes.txt): descriptors are running 32-bit words, the A_RDY waits sit at block-group
    // boundaries and the rare commits behind the loops.
    if (cta == 0) {
      FTR_DECL(0, cluster_id == 0 && lane == 0);
      const bool leader = elect_one();
      uint32_t stage = 0, use = 0;
      uint32_t a_par[2] = {0, 0};  // per stream, bit g: parity of the completion of A_RDY[g] the next consuming layer waits for
      uint32_t a0_n[2] = {0, 0}, hd_n[2] = {0, 0};     // input tiles / heads of the stream so far
      // operand descriptors as (lo, hi) words: the 14-bit address field of lo counts 16-byte units, so a K block is
      // +kWideBlk/16 and a k-step +16; hi (SBO, version) never changes
      const uint64_t d0 = make_smem_desc(0, 128, 1024);
      const uint32_t desc_hi = (uint32_t)(d0 >> 32), desc_lo0 = (uint32_t)d0;
      const uint32_t a_lo_tile = desc_lo0 | (smem_u32(smem) >> 4), a0_lo_tile = desc_lo0 | (smem_u32(smem + off_a0) >> 4);
      const uint32_t b_lo_ring = desc_lo0 | (smem_u32(smem + off_ring) >> 4);
      const uint32_t w_full0 = bar(FB_W_FULL), w_empty0 = bar(FB_W_EMPTY);
      for (FwInst it(i_begin, i_end, E, p.per_slot_input); it.valid(); it.next()) {
        const int ns_inst = streams_of(it.tile);
        for (int l = 0; l < L; ++l) {
          const WideLayerPlan& lp = p.lp[l];
          const int KC = lp.k_chunks, NCH = lp.n_chunks;
          const int n_kgrp = (KC + kFwGroup - 1) / kFwGroup;     // block groups of this layer's K axis
#pragma unroll
         for (int st = 0; st < 2; ++st) {                         // streams alternate layer by layer: X.l, Y.l, X.l+1, ...
          if (st >= ns_inst) break;
          uint32_t seen = 0;
          auto wait_group = [&](int g) {     // group g of the stream's activation tile written by the previous layer's conversion
            if (!((seen >> g) & 1u)) {
              mbar_wait(bar_a_rdy(st, g), (a_par[st] >> g) & 1u, p.status, FWC_MMA_A);
              seen |= 1u << g;
            }
          };
          for (int c = 0; c < NCH; ++c) {
            const int chunk_n = min(kWideNC, lp.npad - c * kWideNC);
            const uint32_t idesc = make_idesc(256, chunk_n, TF32 ? kFmtTF32 : kFmtBF16);
            const uint32_t d_tmem = tmem_base + (uint32_t)((st + c) * kWideNC);     // (two streams: one chunk each)
            FTR(10000 + l * 1000 + c * 100 + st * 10);
            if (l == 0) {
              if (c == 0) {
                if (it.new_input()) { mbar_wait(bar_a0_full(st), a0_n[st] & 1, p.status, FWC_MMA_A0); ++a0_n[st]; }
                // the stream's previous head sits in its accumulator columns 0..; everything above was drained by its last
                // hidden conversion, which completed before the head MMAs were issued
                if (hd_n[st] > 0) mbar_wait(bar_head_drained(st), (hd_n[st] - 1) & 1, p.status, FWC_MMA_HEAD);
              }
            } else {
              // accumulator columns [256 c, 256 c + chunk_n) must have been read by the conversion of layer l - 1, which
              // walks the columns in order: wait for the group whose last block ends at or beyond them
              int drain = (c * kWideNC + chunk_n + lp.kc - 1) / lp.kc;
              if (drain > KC) drain = KC;
              wait_group((drain - 1) / kFwGroup);
            }
            uint32_t a_lo = (l == 0) ? a0_lo_tile + (uint32_t)st * (a0_bytes >> 4) : a_lo_tile + (uint32_t)st * (a_bytes >> 4);
            uint32_t acc = 0;
            for (int g = 0; g < n_kgrp; ++g) {
              if (l > 0) wait_group(g);
              FTR(15000 + l * 1000 + c * 100 + g);
              const int j_end = min((g + 1) * kFwGroup, KC);
              for (int j = g * kFwGroup; j < j_end; ++j) {
                mbar_wait(w_full0 + 8u * stage, use & 1, p.status, FWC_MMA_W);
                tc_fence_after();
                if (leader) {
                  const uint32_t b_lo = b_lo_ring + stage * (kWideBlk >> 4);
                  mma_ss_lohi<2, TF32>(d_tmem, a_lo, b_lo, desc_hi, idesc, acc);
                  mma_ss_lohi<2, TF32>(d_tmem, a_lo + 16, b_lo + 16, desc_hi, idesc, 1u);
                  mma_ss_lohi<2, TF32>(d_tmem, a_lo + 32, b_lo + 32, desc_hi, idesc, 1u);
                  mma_ss_lohi<2, TF32>(d_tmem, a_lo + 48, b_lo + 48, desc_hi, idesc, 1u);
                  mma_commit_2sm(w_empty0 + 8u * stage, 0x3);
                }
                acc = 1u;
                a_lo += kWideBlk >> 4;
                if (++stage == (uint32_t)S) { stage = 0; ++use; }
              }
              // two-chunk layers: chunk 1 has read block group g of the activation tile: the conversion of chunk 0 (running
              // under these MMAs) may overwrite it
              if (NCH == 2 && c == 1 && g < kFwEarlyGroups && leader) mma_commit_2sm(bar(FB_A_FREE + g), 0x3);
            }
            FTR(20000 + l * 1000 + c * 100);
            if (leader) {
              if (NCH == 2 && c == 0) mma_commit_2sm(bar(FB_ACC0_FULL), 0x3);    // chunk 0 complete: its conversion starts now
              if (NCH == 2 && c == 1)                                             // groups this layer's K axis does not reach
                for (int g = n_kgrp; g < kFwEarlyGroups; ++g) mma_commit_2sm(bar(FB_A_FREE + g), 0x3);
              if (c == NCH - 1) {
                mma_commit_2sm(bar_layer_done(st), 0x3);
                if (l == 0 && it.last_use()) mma_commit_2sm(bar_a0_empty(st), 0x3);
              }
            }
            __syncwarp();
          }
          if (l > 0) a_par[st] ^= (1u << n_kgrp) - 1u;     // one completion of every group's A_RDY consumed
         }
        }
#pragma unroll
        for (int st = 0; st < 2; ++st) if (st < ns_inst) ++hd_n[st];
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue warps 4..19: TMEM lane quadrant q (the hardware ties a warp to lanes 32 * (warp % 4) ..), column part
    // `part` of every block group: four warps per SM sub-partition hide each other's TMEM / bias-load / fence latencies
    const int q = warp & 3, part = (warp - 4) >> 2;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const int row_in_tile = q * 32 + lane;
    const uint32_t row_off = (uint32_t)(row_in_tile >> 3) * 1024u + (uint32_t)(row_in_tile & 7) * 16u;
    const uint32_t d_addr0 = tmem_base + lane_base;
    uint8_t* a_row0 = smem + row_off;
    uint32_t ld_n[2] = {0, 0}, two_n = 0;     // completions of LAYER_DONE (per stream) / of the two-chunk barriers consumed
    FTR_DECL(part == 0 ? 1 : 2, cluster_id == 0 && cta == 0 && q == 0 && (part == 0 || part == 3) && lane == 0);
    for (FwInst it(i_begin, i_end, E, p.per_slot_input); it.valid(); it.next()) {
      const int e = it.e;
      const int member = fw_member_of(p, e);
      const int ns_inst = streams_of(it.tile);
      for (int l = 0; l < L; ++l) {
        const WideLayerPlan& lp = p.lp[l];
        const float* bias = p.bias + lp.b_off + (size_t)member * lp.npad;
        uint32_t ra[16], rb[16];
        if (l < L - 1) {
          // A block group = kFwGroup blocks = GC columns; this warp converts columns [part * GC/4, (part + 1) * GC/4) of every
          // group in PPG 16-column pieces: piece t -> group t / PPG.  Groups beyond npad (a partial last group) have no
          // columns for the higher parts: those warps only arrive.
          constexpr int KCOLS = TF32 ? kWideKC / 2 : kWideKC;   // columns per 16 KB block
          constexpr int GC = kFwGroup * KCOLS;                 // 128 (bf16) / 64 (tf32)
          constexpr int PPG = GC / 64;                         // 2 / 1
          const int n_grp = (lp.npad + GC - 1) / GC;
          auto col_of = [&](int t) { return (t / PPG) * GC + part * (GC / 4) + (t % PPG) * 16; };
          int n_t = n_grp * PPG;                               // this warp's pieces with columns below npad (a prefix)
          while (n_t > 0 && col_of(n_t - 1) >= lp.npad) --n_t;
          // The bias of this warp's columns goes through shared memory: one 16-byte global load per lane and pass, issued
          // while the MMAs still run, instead of four L2-latency loads in front of every piece (a new member every instance
          // makes them cold misses, on all four warps of a sub-partition at once: ~270 cycles per piece in the ncu samples).
          // The four quadrant warps of a part write the same values to the same words; a warp reads them after its own
          // store, and the named barrier keeps the words of the previous pass alive until all four warps are done with it.
          {
            constexpr int FPG = GC / 4;                        // floats per group and warp: 32 (bf16) / 16 (tf32)
            const int g = lane / (FPG / 4), c4 = lane % (FPG / 4);
            const int col = g * GC + part * FPG + c4 * 4;
            const bool act = g < n_grp && col < lp.npad;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (act) v = __ldg(reinterpret_cast<const float4*>(bias + col));
            asm volatile("bar.sync %0, 128;" ::"r"(1 + part) : "memory");
            if (act) *reinterpret_cast<float4*>(bias_s + col) = v;
            __syncwarp();
          }
#pragma unroll
         for (int st = 0; st < 2; ++st) {                       // the streams' conversions alternate: X under Y's MMAs, Y under X's next layer
          if (st >= ns_inst) break;
          const uint32_t d_addr = d_addr0 + (uint32_t)(st * kWideNC);
          uint8_t* a_row = a_row0 + (size_t)st * a_bytes;
          // two-chunk layers: the columns under N chunk 0 (groups 0 .. kFwEarlyGroups) are converted while chunk 1's MMAs run
          const bool two = lp.n_chunks == 2;
          const int n_early = two ? min(n_t, kFwEarlyGroups * PPG) : 0;
          auto arrive_group = [&](int g, bool wrote) {
#ifndef FW_EXP_NOFENCE   // timing experiment: without the proxy fence (racy)
            if (wrote) fence_proxy_async_smem();         // generic-proxy smem writes -> visible to tcgen05.mma
#endif
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster_light(bar_a_rdy(st, g), 0);
          };
          auto finish = [&](int t, const uint32_t (&r)[16], const float (&b)[16]) {
            const int n0 = col_of(t);
            uint32_t o[TF32 ? 16 : 8];
            if constexpr (TF32) {
#pragma unroll
              for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(round_tf32(fw_activate<HACT, true>(__uint_as_float(r[i]) + b[i])));
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i)
                o[i] = pack_bf16x2(fw_activate<HACT, false>(__uint_as_float(r[2 * i]) + b[2 * i]),
                                   fw_activate<HACT, false>(__uint_as_float(r[2 * i + 1]) + b[2 * i + 1]));
            }
            // early groups: chunk 1's MMAs must have read the group's blocks before they are overwritten
            if (t < n_early && t % PPG == 0) { FW_EPI_WAIT(bar(FB_A_FREE + t / PPG), two_n & 1, p.status, FWC_EPI); FTR(35000 + l * 10 + t / PPG); }
#ifdef FW_EXP_NOSTS   // timing experiment: no activation stores (results are wrong)
            if (n0 < 0)
#endif
            if constexpr (TF32) {
              uint8_t* dst = a_row + (size_t)(n0 >> 5) * kWideBlk + (size_t)((n0 & 31) >> 2) * 128;
#pragma unroll
              for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(dst + 128 * i) = make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
            } else {
              uint8_t* dst = a_row + (size_t)(n0 >> 6) * kWideBlk + (size_t)((n0 & 63) >> 3) * 128;
              *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
              *reinterpret_cast<uint4*>(dst + 128) = make_uint4(o[4], o[5], o[6], o[7]);
            }
            if ((t + 1) % PPG == 0) { arrive_group(t / PPG, true); FTR(36000 + l * 10 + t / PPG); }   // this warp's share of the group is written
          };
          auto step = [&](int t, int t_end, const uint32_t (&r)[16], uint32_t (&r_next)[16]) {
            float b[16];                                       // warp-uniform shared-memory addresses (broadcast reads)
            const float4* src = reinterpret_cast<const float4*>(bias_s + col_of(t));
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float4 v = src[i];
              b[4 * i] = v.x; b[4 * i + 1] = v.y; b[4 * i + 2] = v.z; b[4 * i + 3] = v.w;
            }
            tmem_wait_ld();
            if (t + 1 < t_end) tmem_ld16(d_addr + col_of(t + 1), r_next);
            finish(t, r, b);
          };
          auto run = [&](int t0, int t1) {                     // pieces [t0, t1), TMEM loads one piece ahead
#ifdef FW_EXP_NOEPI   // timing experiment: arrivals only (results are wrong)
            for (int t = t0; t < t1; ++t) if ((t + 1) % PPG == 0) arrive_group(t / PPG, false);
            return;
#endif
            if (t0 < t1) tmem_ld16(d_addr + col_of(t0), ra);
            for (int t = t0; t < t1; t += 2) {
              step(t, t1, ra, rb);
              if (t + 1 < t1) step(t + 1, t1, rb, ra);
            }
          };
          FTR(30000 + l);
          if (two) {
            FW_EPI_WAIT(bar(FB_ACC0_FULL), two_n & 1, p.status, FWC_EPI);
            tc_fence_after();
            FTR(31000 + l);
            run(0, n_early);
            FTR(32000 + l);
          }
          FW_EPI_WAIT(bar_layer_done(st), ld_n[st] & 1, p.status, FWC_EPI);
          ++ld_n[st];
          tc_fence_after();
          FTR(33000 + l);
          run(n_early, n_t);
          FTR(34000 + l);
          if (two) ++two_n;
          for (int g = n_t / PPG; g < n_grp; ++g) arrive_group(g, false);   // groups this warp has no columns in
         }
        } else {
          // head: (acc + b) * s + shift into the tile-major fp32 scratch (bnn.py:878-893); the accumulator is released
          // (HEAD_DRAINED) as soon as the warp's columns are in registers, before the global stores
#pragma unroll
         for (int st = 0; st < 2; ++st) {
          if (st >= ns_inst) break;
          const uint32_t d_addr = d_addr0 + (uint32_t)(st * kWideNC);
          const long long blk = tile_of(it.tile, st) * 2 + cta;           // 128-row block of the batch
          FTR(30000 + l);
          FW_EPI_WAIT(bar_layer_done(st), ld_n[st] & 1, p.status, FWC_EPI);
          ++ld_n[st];
          tc_fence_after();
          FTR(33000 + l);
          const int nch = lp.npad >> 4;
          const int c_begin = 16 * ((part * nch) >> 2), c_end = 16 * (((part + 1) * nch) >> 2);
          const bool live_row = blk * kTileRows + row_in_tile < n_rows;
          float* out = p.heads + ((size_t)(blk * E + e) * p.n_head) * kTileRows + row_in_tile;
          auto drained = [&]() {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster_light(bar_head_drained(st), 0);
          };
          auto store16 = [&](int c0, const uint32_t (&r)[16]) {
            if (live_row) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int col = c0 + i;
                if (col < p.n_head) {
                  const float sc = p.head_scale ? __ldg(p.head_scale + col) : 1.f;
                  const float sh = p.head_shift ? __ldg(p.head_shift + col) : 0.f;
                  out[(size_t)col * kTileRows] = fmaf(__uint_as_float(r[i]) + __ldg(bias + col), sc, sh);
                }
              }
            }
          };
          if (c_begin >= c_end) drained();
          for (int c0 = c_begin; c0 < c_end; c0 += 32) {
            const bool two = c0 + 16 < c_end;
            tmem_ld16(d_addr + c0, ra);
            if (two) tmem_ld16(d_addr + c0 + 16, rb);
            tmem_wait_ld();
            if (c0 + 32 >= c_end) { drained(); FTR(34000 + l); }
            store16(c0, ra);
            if (two) store16(c0 + 16, rb);
          }
          FTR(37000 + l);
         }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync();
  if (warp == 2) tmem_dealloc<2>(tmem_base, 512);
}

size_t fused_wide_smem_bytes(const FusedWideParams& p) {
  return (size_t)(p.streams * (p.a_blocks + p.a0_blocks) + p.stages) * kWideBlk + kFwBarBytes + kFwBiasFloats * sizeof(float);
}

template <int HACT, bool TF32>
static cudaError_t launch_fused_wide_variant2(const FusedWideParams& p, int num_sms, cudaStream_t stream) {
  const size_t smem = fused_wide_smem_bytes(p);
  static size_t smem_set = 0;
  if (smem > smem_set) {
    cudaError_t e = cudaFuncSetAttribute(fused_wide_kernel<HACT, TF32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    smem_set = smem;
  }
  int clusters = num_sms / 2;
  if (!p.n_rows_dev) {
    const long long n_tiles = (p.n_rows + 2 * kTileRows - 1) / (2 * kTileRows);
    const long long inst = (p.streams == 2 ? (n_tiles + 1) / 2 : n_tiles) * p.E;
    if (inst < clusters) clusters = (int)(inst > 0 ? inst : 1);
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(kFwThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, fused_wide_kernel<HACT, TF32>, p);
}
template <int HACT>
static cudaError_t launch_fused_wide_variant(const FusedWideParams& p, int num_sms, cudaStream_t stream) {
  return p.tf32 ? launch_fused_wide_variant2<HACT, true>(p, num_sms, stream) : launch_fused_wide_variant2<HACT, false>(p, num_sms, stream);
}

// act: the activation every hidden layer shares (constructor.py:57-60)
cudaError_t launch_fused_wide(const FusedWideParams& p, int act, int num_sms, cudaStream_t stream) {
  switch (act) {
    case ACT_SWISH:   return launch_fused_wide_variant<ACT_SWISH>(p, num_sms, stream);
    case ACT_RELU:    return launch_fused_wide_variant<ACT_RELU>(p, num_sms, stream);
    case ACT_TANH:    return launch_fused_wide_variant<ACT_TANH>(p, num_sms, stream);
    case ACT_SIGMOID: return launch_fused_wide_variant<ACT_SIGMOID>(p, num_sms, stream);
    case ACT_NONE:    return launch_fused_wide_variant<ACT_NONE>(p, num_sms, stream);
    default:          return cudaErrorInvalidValue;
  }
}

}  // namespace cmbpo
