// rollout_ops.cu -- device-resident bookkeeping of the batched model rollout: what
// ModelSampler.sample does around env.step (softlearning/samplers/model_sampler.py:311-332,
// :386-474) and what ModelBuffer.store_multiple writes (softlearning/replay_pools/
// modelbuffer.py:130-173), restricted to the quantities the hot path owns.
//
// The reference compacts ~25 host arrays with boolean masks every step; here one call keeps
// everything on the device: decide (DKL-accumulate cutoff, max_samples early termination) ->
// stable rank -> pool write at column `ptr` -> length / terminal masks -> stable compaction of
// the surviving rows.  Row order is preserved exactly (boolean-mask semantics).
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
#include "../../include/cmbpo_b200.h"

namespace cmbpo {

constexpr int kRB = 256;  // rows per block

// device-resident rollout state = cmbpo_rollout_state of the public header
struct RolloutState {
  int n_alive;
  int ptr;
  long long total_samples;
  int n_stored_last;
  int finished;
  int tmp_total_certain, tmp_n_early, tmp_n_keep, tmp_do_store;
  unsigned int ticket_a, ticket_b;   // block counters of the two fused scan epilogues (zero between launches)
};
static_assert(sizeof(RolloutState) == sizeof(cmbpo_rollout_state), "state layout");

__device__ __forceinline__ int block_exclusive_scan(int flag, int* total) {
  // 256 threads; returns exclusive prefix of `flag` inside the block
  __shared__ int warp_sums[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned ballot = __ballot_sync(0xffffffffu, flag);
  const int intra = __popc(ballot & ((1u << lane) - 1u));
  if (lane == 0) warp_sums[warp] = __popc(ballot);
  __syncthreads();
  int base = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) { if (w < warp) base += warp_sums[w]; tot += warp_sums[w]; }
  __syncthreads();
  if (total) *total = tot;
  return base + intra;
}

// exclusive scan of arbitrary non-negative ints over the block's 256 threads
__device__ __forceinline__ int block_exclusive_scan_int(int v, int* total) {
  __shared__ int ws[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
  if (lane == 31) ws[warp] = incl;
  __syncthreads();
  int base = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) { if (w < warp) base += ws[w]; tot += ws[w]; }
  __syncthreads();
  if (total) *total = tot;
  return base + incl - v;
}

// "last block done" epilogue shared by k_decide and k_store: every block publishes its count and takes a ticket;
// the block that takes the last one scans the per-block counts in place (exclusive) and returns true with the
// total in *total_out (valid on every thread of that block).  Replaces two single-block scan launches per step.
__device__ __forceinline__ bool last_block_scan(int* blk_cnt, int my_count, int n_blocks, unsigned int* ticket, int* total_out) {
  __shared__ bool s_last;
  __shared__ int s_carry;
  if (threadIdx.x == 0) blk_cnt[blockIdx.x] = my_count;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) { s_last = atomicAdd(ticket, 1u) == gridDim.x - 1; s_carry = 0; }
  __syncthreads();
  if (!s_last) return false;
  __threadfence();
  for (int base = 0; base < n_blocks; base += kRB) {
    const int i = base + threadIdx.x;
    const int v = (i < n_blocks) ? __ldcg(blk_cnt + i) : 0;
    int tot;
    const int ex = block_exclusive_scan_int(v, &tot);
    if (i < n_blocks) blk_cnt[i] = s_carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) s_carry += tot;
    __syncthreads();
  }
  *total_out = s_carry;
  if (threadIdx.x == 0) *ticket = 0;   // ready for the next launch
  return true;
}

// idx / w for 0 <= idx < 2^20 and small w through a float reciprocal (exact: the quotient's fractional part stays
// at least 0.5/w away from an integer, far above the float rounding error); replaces a ~20-instruction integer division
// in the payload copy loops
__device__ __forceinline__ int fast_div(int idx, float inv_w) { return (int)(((float)idx + 0.5f) * inv_w); }


// Row payload scatter shared by k_store / k_store_payload / k_compact_dev: element idx of the block's contiguous source
// rows goes to dst[dst_row(r) * W + d] when dst_row(r) >= 0.  The loads of kPayU iterations are issued before the first
// store (explicit register batch): source and destination are plain pointers, so without the batch the compiler keeps
// every load behind the previous iteration's store and the loop becomes a chain of serial DRAM round trips
// (k_store: 178 us for 330 MB at 1 M rows, ~30 long-scoreboard stalls per issue, profiles/r02_small_kernels_ncu_summary.txt).
constexpr int kPayU = 8;
template <int NSRC, typename RowFn>
__device__ __forceinline__ void scatter_rows(const float* const (&src)[NSRC], float* const (&dst)[NSRC], int n_elems, int W,
                                             float inv_w, RowFn dst_row) {
  for (int base = threadIdx.x; base < n_elems; base += kRB * kPayU) {
    float v[NSRC][kPayU];
    long long o[kPayU];
#pragma unroll
    for (int u = 0; u < kPayU; ++u) {
      const int idx = base + u * kRB;
      o[u] = -1;
      if (idx < n_elems) {
        const int r = fast_div(idx, inv_w), d = idx - r * W;
        const long long row = dst_row(r);
        if (row >= 0) {
          o[u] = row * W + d;
#pragma unroll
          for (int k = 0; k < NSRC; ++k) v[k][u] = __ldg(src[k] + idx);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kPayU; ++u)
      if (o[u] >= 0) {
#pragma unroll
        for (int k = 0; k < NSRC; ++k) dst[k][o[u]] = v[k][u];
      }
  }
}

// A: too_uncertain flags (model_sampler.py:311-315) -> per-block count of still-certain rows
__global__ void __launch_bounds__(kRB) k_decide(cmbpo_rollout_io io, uint8_t* certain, int* blk_cnt, int* snap, int* hist) {
  RolloutState* st = static_cast<RolloutState*>(io.state_dev);
  const int n = st->n_alive;
  const int i = blockIdx.x * kRB + threadIdx.x;
  int c = 0;
  if (i < n) {
    bool unc = false;
    if (io.mode == 1) {  // 'uncertainty': next_dkl = accum[alive] + dkl_path >= dkl_lim
      const int pid = static_cast<const int*>(io.path_id_dev)[i];
      // float64 accumulator + float32 increment, compared with a Python float (model_sampler.py:311-312)
      const double nd = io.dkl_accum_dev[pid] + (double)static_cast<const float*>(io.dkl_path_dev)[i];
      unc = nd >= io.dkl_lim;
    }   // mode 2 ('iv_gae'): no uncertainty cutoff in the sampler (model_sampler.py:315)
    c = unc ? 0 : 1;
    certain[i] = (uint8_t)c;
  }
  int tot;
  block_exclusive_scan(c, &tot);
  int total;
  if (!last_block_scan(blk_cnt, tot, (n + kRB - 1) / kRB, &st->ticket_a, &total)) return;
  if (threadIdx.x == 0) {
    *snap = n;                       // the pre-advance alive count, needed by the compaction
    if (hist) *hist = n;
    // model_sampler.py:327-332: n = total_samples + alive - too_uncertain - max_samples
    long long n_early = 0;
    if (io.max_samples > 0 && io.mode != 2) {   // :327 `max_samples and not rollout_mode=='iv_gae'`
      n_early = st->total_samples + (long long)total - io.max_samples;
      if (n_early < 0) n_early = 0;
      if (n_early > total) n_early = total;
    }
    st->tmp_total_certain = total;
    st->tmp_n_early = (int)n_early;
    const int n_remaining = total - (int)n_early;
    st->tmp_do_store = n_remaining > 0 ? 1 : 0;   // model_sampler.py:339-342: nothing alive -> return
    st->n_stored_last = n_remaining;
  }
}

// C: stable rank among certain rows -> remaining mask -> pool write (modelbuffer.py:130-173),
//    dkl accumulate (:415), length mask (:437, quirk Q8) and terminal mask (:453-456) -> keep flag
__global__ void __launch_bounds__(kRB) k_store(cmbpo_rollout_io io, uint8_t* certain, const int* blk_off,
                                                uint8_t* keep, int* blk_keep_cnt) {
  RolloutState* st = static_cast<RolloutState*>(io.state_dev);
  const int n = st->n_alive;
  const int n_early = st->tmp_n_early, do_store = st->tmp_do_store;   // read before any block can take the last ticket
  const int i = blockIdx.x * kRB + threadIdx.x;
  const int c = (i < n) ? certain[i] : 0;
  const int rank = blk_off[blockIdx.x] + block_exclusive_scan(c, nullptr);
  const bool remaining = c && rank >= n_early && do_store;
  const int ptr = st->ptr;
  const int T = io.max_path_length;
  const bool path_end = (ptr + 1) >= (T - 1);  // path_length after increment >= max_path_length-1
  __shared__ int s_pid[kRB];
  __shared__ uint8_t s_rem[kRB];
  // 'iv_gae' (mode 2): every per-transition array carries a leading member axis -- step arrays
  // [E][n_paths (capacity)][.], pools [E][n_paths][T][.] (modelbuffer.py:48-51,:134-150)
  const int E = (io.mode == 2 && io.n_members > 1) ? io.n_members : 1;
  const size_t cap = (size_t)io.n_paths;
  int pid = 0;
  bool term = false;   // 2-D: the row's terminal; iv_gae: np.any(terminal, axis=0) (model_sampler.py:445)
  if (i < n) {
    pid = static_cast<const int*>(io.path_id_dev)[i];
    for (int e = 0; e < E; ++e) term = term || static_cast<const uint8_t*>(io.term_dev)[e * cap + i] != 0;
  }
  s_pid[threadIdx.x] = pid;
  s_rem[threadIdx.x] = remaining ? 1 : 0;
  if (remaining) {
    const size_t slot = (size_t)ptr * cap + pid;   // time-major storage [T][n_paths]: one step's stores are contiguous
    for (int e = 0; e < E; ++e) {
      const size_t se = e * cap * T + slot, ie = e * cap + i;
      static_cast<float*>(io.pool_rew_dev)[se] = static_cast<const float*>(io.rew_dev)[ie];
      static_cast<float*>(io.pool_cost_dev)[se] = io.cost_dev ? static_cast<const float*>(io.cost_dev)[ie] : 0.f;
      static_cast<uint8_t*>(io.pool_term_dev)[se] = static_cast<const uint8_t*>(io.term_dev)[ie] ? 1 : 0;
    }
    if (io.pool_val_dev) {   // val_buf / cval_buf[alive, ptr] = V(s_t), VC(s_t) (modelbuffer.py:141-146)
      for (int e = 0; e < E; ++e) {
        io.pool_val_dev[e * cap * T + slot] = io.val_cur_dev[e * cap + i];
        io.pool_cval_dev[e * cap * T + slot] = io.cval_cur_dev[e * cap + i];
      }
    }
    static_cast<float*>(io.pool_dyn_error_dev)[slot] =
        io.ep_var_mean_dev ? static_cast<const float*>(io.ep_var_mean_dev)[i] : 0.f;
    static_cast<uint8_t*>(io.pool_populated_dev)[slot] = 1;
    io.dkl_accum_dev[pid] += (double)static_cast<const float*>(io.dkl_path_dev)[i];
  }
  const int k = (remaining && !term && !path_end) ? 1 : 0;
  if (i < n) { keep[i] = (uint8_t)k; certain[i] = remaining ? 1 : 0; }   // certain[] now = 'stored' flags (k_store_payload)
  if (io.finish_ptr_dev && i < n && !k) {
    // this path ends now: what ModelSampler._finish_paths hands to finish_path_multiple (model_sampler.py:497-548)
    //  - cut before the store (too uncertain / max_samples): column ptr, bootstrap V(s_t), VC(s_t)   (:336, append both)
    //  - length reached after the store: column ptr+1, bootstrap V(s_t+1), VC(s_t+1)                 (:438, append both)
    //  - environment terminal: column ptr+1, value bootstrap 0, cost bootstrap VC(s_t+1)             (:456, append_vals=False)
    io.finish_ptr_dev[pid] = remaining ? ptr + 1 : ptr;
    for (int e = 0; e < E; ++e) {
      float lv, lc;
      if (!remaining) { lv = io.val_cur_dev[e * cap + i]; lc = io.cval_cur_dev[e * cap + i]; }
      else { lv = path_end ? io.val_next_dev[e * cap + i] : 0.f; lc = io.cval_next_dev[e * cap + i]; }
      io.last_val_dev[e * cap + pid] = lv;
      io.last_cval_dev[e * cap + pid] = lc;
    }
  }
  int tot;
  block_exclusive_scan(k, &tot);   // includes the __syncthreads that publishes s_pid / s_rem
  // row payloads: flat element loop over the block's 256 rows (source rows are contiguous, so loads are
  // fully coalesced; every thread has several independent load->store chains in flight -- a
  // warp-per-row loop here was a 32-deep serial latency chain, 35 us even for a handful of rows)
  const int OD = io.obs_dim, AD = io.act_dim;
  const float inv_od = 1.0f / (float)OD, inv_ad = AD > 0 ? 1.0f / (float)AD : 0.f;
  const int row_base = blockIdx.x * kRB;
  const int n_here = min(kRB, n - row_base);
  for (int e = 0; e < (E == 1 ? 1 : 0); ++e) {   // member-leading pools: payload moved by k_store_payload, one grid row per member
    const float* so = static_cast<const float*>(io.cur_obs_dev) + (e * cap + row_base) * OD;
    const float* sn = static_cast<const float*>(io.next_obs_dev) + (e * cap + row_base) * OD;
    float* po = static_cast<float*>(io.pool_obs_dev) + e * cap * T * OD;
    float* pn = static_cast<float*>(io.pool_next_obs_dev) + e * cap * T * OD;
    auto pool_row = [&](int r) -> long long { return s_rem[r] ? (long long)ptr * (long long)cap + s_pid[r] : -1; };
    {
      const float* const src2[2] = {so, sn};
      float* const dst2[2] = {po, pn};
      scatter_rows<2>(src2, dst2, n_here * OD, OD, inv_od, pool_row);
    }
    if (AD > 0) {
      const float* const src1[1] = {static_cast<const float*>(io.act_dev) + (e * cap + row_base) * AD};
      float* const dst1[1] = {static_cast<float*>(io.pool_act_dev) + e * cap * T * AD};
      scatter_rows<1>(src1, dst1, n_here * AD, AD, inv_ad, pool_row);
    }
  }
  // fused scan epilogue (was a separate single-block launch): the last block to finish its stores scans the
  // keep counts and advances the rollout state (every block read the state at its start)
  int total;
  if (!last_block_scan(blk_keep_cnt, tot, (n + kRB - 1) / kRB, &st->ticket_b, &total)) return;
  if (threadIdx.x == 0) {
    st->tmp_n_keep = total;
    if (do_store) {
      st->total_samples += st->n_stored_last;      // model_sampler.py:386
      st->ptr += 1;                                // modelbuffer.py:173
    }
    st->n_alive = do_store ? total : 0;
    st->finished = (st->n_alive == 0) ? 1 : 0;
  }
}

// 'iv_gae' payload copy: grid (row blocks, members), so the E member blocks of obs / next_obs / act move in parallel
// instead of in a per-block loop (k_store at 125k rows x 7 members: 520 us -> see profiles/).  Runs after k_store:
// `stored` are the flags it published, the column is the pre-increment ptr.
__global__ void __launch_bounds__(kRB) k_store_payload(cmbpo_rollout_io io, const uint8_t* stored, const int* n_before_dev) {
  const RolloutState* st = static_cast<const RolloutState*>(io.state_dev);
  if (!st->tmp_do_store) return;
  const int n = *n_before_dev;
  const int ptr = st->ptr - 1;                     // k_store's epilogue already advanced it
  const int T = io.max_path_length, OD = io.obs_dim, AD = io.act_dim;
  const size_t cap = (size_t)io.n_paths;
  const int e = blockIdx.y;
  const int row_base = blockIdx.x * kRB;
  const int n_here = min(kRB, n - row_base);
  if (n_here <= 0) return;
  __shared__ int s_pid[kRB];
  __shared__ uint8_t s_rem[kRB];
  const int i = row_base + threadIdx.x;
  s_pid[threadIdx.x] = (i < n) ? static_cast<const int*>(io.path_id_dev)[i] : 0;
  s_rem[threadIdx.x] = (i < n) ? stored[i] : 0;
  __syncthreads();
  const float inv_od = 1.0f / (float)OD, inv_ad = AD > 0 ? 1.0f / (float)AD : 0.f;
  const float* so = static_cast<const float*>(io.cur_obs_dev) + (e * cap + row_base) * OD;
  const float* sn = static_cast<const float*>(io.next_obs_dev) + (e * cap + row_base) * OD;
  float* po = static_cast<float*>(io.pool_obs_dev) + e * cap * T * OD;
  float* pn = static_cast<float*>(io.pool_next_obs_dev) + e * cap * T * OD;
  auto pool_row = [&](int r) -> long long { return s_rem[r] ? (long long)ptr * (long long)cap + s_pid[r] : -1; };
  {
    const float* const src2[2] = {so, sn};
    float* const dst2[2] = {po, pn};
    scatter_rows<2>(src2, dst2, n_here * OD, OD, inv_od, pool_row);
  }
  if (AD > 0) {
    const float* const src1[1] = {static_cast<const float*>(io.act_dev) + (e * cap + row_base) * AD};
    float* const dst1[1] = {static_cast<float*>(io.pool_act_dev) + e * cap * T * AD};
    scatter_rows<1>(src1, dst1, n_here * AD, AD, inv_ad, pool_row);
  }
}

// synthetic benchmark policy a = tanh(obs @ P); not part of the reference.  One thread per row: the
// block's 256 obs rows come in through a coalesced flat copy into shared memory (row stride OD + 1: conflict-free row
// reads); P sits in shared memory with its rows padded to a multiple of 4 columns and is read as 16-byte broadcasts, up to
// 24 action columns accumulated in registers per pass (one pass for every environment of the spec), and the actions leave
// through shared memory as a coalesced flat copy.  (Version 1 computed one output element per thread from a 64-row tile:
// 59 M warp instructions per 1 M rows; version 2 read P with scalar broadcasts, 9 shared loads per 8 FMAs, and was bound
// by the shared-memory pipe -- 198 us for 875 k Humanoid rows; this one is bound by its traffic.)
constexpr int kPolRows = 256;
template <int NG>
__device__ __forceinline__ void stub_policy_row(const float* __restrict__ xo, const float* __restrict__ sP, int AP, int OD,
                                                int a0, int AD, float* __restrict__ out) {
  float acc[NG * 4];
#pragma unroll
  for (int u = 0; u < NG * 4; ++u) acc[u] = 0.f;
  for (int d = 0; d < OD; ++d) {
    const float x = xo[d];
    const float4* pr = reinterpret_cast<const float4*>(sP + d * AP + a0);
#pragma unroll
    for (int g = 0; g < NG; ++g) {
      const float4 w = pr[g];
      acc[4 * g + 0] = fmaf(x, w.x, acc[4 * g + 0]);
      acc[4 * g + 1] = fmaf(x, w.y, acc[4 * g + 1]);
      acc[4 * g + 2] = fmaf(x, w.z, acc[4 * g + 2]);
      acc[4 * g + 3] = fmaf(x, w.w, acc[4 * g + 3]);
    }
  }
#pragma unroll
  for (int u = 0; u < NG * 4; ++u)
    if (a0 + u < AD) out[a0 + u] = tanhf(acc[u]);
}

__global__ void __launch_bounds__(kPolRows) k_stub_policy(const float* __restrict__ obs, const float* __restrict__ P,
                                                          float* __restrict__ act, const int* n_rows_dev,
                                                          long long n_rows, int OD, int AD, int n_slots) {
  extern __shared__ __align__(16) float sm[];
  const int AP = (AD + 3) & ~3;
  float* sP = sm;                         // [OD][AP], columns AD..AP-1 zero
  float* sO = sm + OD * AP;               // [kPolRows][OD + 1]
  float* sA = sO + kPolRows * (OD + 1);   // [kPolRows][AD]
  const float inv_ap = 1.0f / (float)AP;
  for (int i = threadIdx.x; i < OD * AP; i += blockDim.x) {
    const int d = fast_div(i, inv_ap), a = i - d * AP;
    sP[i] = a < AD ? P[d * AD + a] : 0.f;
  }
  const long long n = n_rows_dev ? (long long)(*n_rows_dev) : n_rows;
  const long long tiles_per_slot = (n + kPolRows - 1) / kPolRows;
  const long long n_tiles = tiles_per_slot * n_slots;   // slot e of obs / act starts at row e * n_rows (capacity)
  const float inv_od = 1.0f / (float)OD;
  const bool dense = (OD & 1) && (reinterpret_cast<uintptr_t>(obs) & 15) == 0 && (((size_t)n_rows * OD) & 3) == 0 &&
                     ((OD * AP) & 3) == 0;
  const int ld_o = dense ? OD : OD + 1;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long slot = tile / tiles_per_slot, r0 = (tile - slot * tiles_per_slot) * kPolRows;
    const int rows_here = (int)min((long long)kPolRows, n - r0);
    const size_t row_base = (size_t)(slot * n_rows + r0);
    __syncthreads();
    if (dense) {       // odd OD: rows at stride OD are conflict-free as they are, the tile is one flat 16-byte cp.async copy
      const float* src = obs + row_base * OD;
      const int n_el = rows_here * OD, n4 = n_el >> 2;
      for (int q = threadIdx.x; q < n4; q += kPolRows)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(sO + 4 * q)), "l"(src + 4 * q) : "memory");
      asm volatile("cp.async.commit_group;" ::: "memory");
      for (int i = 4 * n4 + threadIdx.x; i < n_el; i += kPolRows) sO[i] = src[i];
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    } else {
      for (int i = threadIdx.x; i < rows_here * OD; i += kPolRows) {
        const int r = fast_div(i, inv_od), d = i - r * OD;
        sO[r * (OD + 1) + d] = obs[row_base * OD + i];
      }
    }
    __syncthreads();
    const int r = threadIdx.x;
    if (r < rows_here) {
      const float* xo = sO + r * ld_o;
      float* out = sA + r * AD;
      for (int a0 = 0; a0 < AP; a0 += 24) {
        switch (min(6, (AP - a0) >> 2)) {
          case 1: stub_policy_row<1>(xo, sP, AP, OD, a0, AD, out); break;
          case 2: stub_policy_row<2>(xo, sP, AP, OD, a0, AD, out); break;
          case 3: stub_policy_row<3>(xo, sP, AP, OD, a0, AD, out); break;
          case 4: stub_policy_row<4>(xo, sP, AP, OD, a0, AD, out); break;
          case 5: stub_policy_row<5>(xo, sP, AP, OD, a0, AD, out); break;
          default: stub_policy_row<6>(xo, sP, AP, OD, a0, AD, out); break;
        }
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < rows_here * AD; i += kPolRows) act[row_base * AD + i] = sA[i];
  }
}

}  // namespace cmbpo

using namespace cmbpo;

static thread_local char g_roll_err[160];
extern "C" size_t cmbpo_rollout_scratch_bytes(int64_t n_paths) {
  const size_t nb = (size_t)((n_paths + kRB - 1) / kRB) + 1;
  // certain[n] + keep[n] (u8) + two int arrays of per-block counts + snapshot word
  return (size_t)n_paths * 2 + 64 + nb * 2 * sizeof(int) + 64;
}

namespace cmbpo {
// E: stable compaction of the surviving rows: next_obs -> new cur_obs, path ids follow
__global__ void __launch_bounds__(kRB) k_compact_dev(cmbpo_rollout_io io, const uint8_t* keep, const int* blk_off,
                                                      const int* n_before_dev) {
  const int n_before = *n_before_dev;
  const int i = blockIdx.x * kRB + threadIdx.x;
  const int k = (i < n_before) ? keep[i] : 0;
  const int dst = blk_off[blockIdx.x] + block_exclusive_scan(k, nullptr);
  __shared__ int s_dst[kRB];
  s_dst[threadIdx.x] = k ? dst : -1;
  const int E = (io.mode == 2 && io.n_members > 1) ? io.n_members : 1;
  const size_t cap = (size_t)io.n_paths;
  const int e0 = blockIdx.y;                        // grid.y = members: every member's block moves in parallel
  if (k && e0 == 0) static_cast<int*>(io.new_path_id_dev)[dst] = static_cast<const int*>(io.path_id_dev)[i];
  if (k && io.new_val_dev && e0 < E) {   // V(s_t+1) of a survivor is V(s_t) of its next step: the critics run once per state
    io.new_val_dev[e0 * cap + dst] = io.val_next_dev[e0 * cap + i];
    io.new_cval_dev[e0 * cap + dst] = io.cval_next_dev[e0 * cap + i];
  }
  __syncthreads();
  const int OD = io.obs_dim;
  const float inv_od = 1.0f / (float)OD;
  const int row_base = blockIdx.x * kRB;
  const int n_here = min(kRB, n_before - row_base);
  {   // iv_gae: every member's own state follows (model_sampler.py:444,:463-464)
    const int e = e0;
    const float* src = static_cast<const float*>(io.next_obs_dev) + (e * cap + row_base) * OD;
    float* dstp = static_cast<float*>(io.new_obs_dev) + e * cap * OD;
    const float* const src1[1] = {src};
    float* const dst1[1] = {dstp};
    scatter_rows<1>(src1, dst1, n_here * OD, OD, inv_od, [&](int r) -> long long { return s_dst[r]; });
  }
}
}  // namespace cmbpo

namespace cmbpo {
// finish_all_paths (model_sampler.py:550-570): one thread per live row; the last thread block clears the alive count
__global__ void __launch_bounds__(kRB) k_finish_all(cmbpo_rollout_io io) {
  RolloutState* st = static_cast<RolloutState*>(io.state_dev);
  const int n = st->n_alive, ptr = st->ptr;
  const int E = (io.mode == 2 && io.n_members > 1) ? io.n_members : 1;
  const size_t cap = (size_t)io.n_paths;
  for (int i = blockIdx.x * kRB + threadIdx.x; i < n; i += gridDim.x * kRB) {
    const int pid = static_cast<const int*>(io.path_id_dev)[i];
    io.finish_ptr_dev[pid] = ptr;
    for (int e = 0; e < E; ++e) {
      io.last_val_dev[e * cap + pid] = io.val_cur_dev[e * cap + i];
      io.last_cval_dev[e * cap + pid] = io.cval_cur_dev[e * cap + i];
    }
  }
  __threadfence();
  __syncthreads();
  __shared__ bool s_last;
  if (threadIdx.x == 0) s_last = atomicAdd(&st->ticket_a, 1u) == gridDim.x - 1;
  __syncthreads();
  if (s_last && threadIdx.x == 0) { st->n_alive = 0; st->finished = 1; st->ticket_a = 0; }
}
}  // namespace cmbpo

extern "C" int cmbpo_rollout_finish_all(const cmbpo_rollout_io* io_, void* stream_) {
  if (!io_ || !io_->state_dev || !io_->path_id_dev || !io_->finish_ptr_dev || !io_->val_cur_dev || !io_->cval_cur_dev ||
      !io_->last_val_dev || !io_->last_cval_dev) return CMBPO_ERR_ARG;
  const cmbpo_rollout_io io = *io_;
  if (io.n_paths <= 0) return CMBPO_OK;
  long long nb = (io.n_paths + kRB - 1) / kRB;
  if (nb > 148 * 8) nb = 148 * 8;
  // every block reads n_alive before the last one clears it: blocks only pass the ticket after their loop
  k_finish_all<<<(int)nb, kRB, 0, (cudaStream_t)stream_>>>(io);
  return cudaGetLastError() == cudaSuccess ? CMBPO_OK : CMBPO_ERR_CUDA;
}

// One sampler step after env.step: see the file header.  `n_paths` bounds the grid; the live row
// count is read from state_dev on the device, so the host never synchronises inside a rollout.
extern "C" int cmbpo_rollout_advance(const cmbpo_rollout_io* io_, void* stream_) {
  if (!io_ || !io_->state_dev || !io_->scratch_dev) return CMBPO_ERR_ARG;
  const cmbpo_rollout_io io = *io_;
  if (io.scratch_bytes < cmbpo_rollout_scratch_bytes(io.n_paths)) return CMBPO_ERR_WORKSPACE;
  if (io.n_paths <= 0) return CMBPO_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  const int nb = (int)((io.n_paths + kRB - 1) / kRB);
  uint8_t* certain = static_cast<uint8_t*>(io.scratch_dev);
  uint8_t* keep = certain + io.n_paths;
  size_t off = ((size_t)io.n_paths * 2 + 63) & ~(size_t)63;
  int* blk_a = reinterpret_cast<int*>(static_cast<uint8_t*>(io.scratch_dev) + off);
  int* blk_b = blk_a + nb + 1;
  int* snap = blk_b + nb;   // last word of blk_b's slack
  k_decide<<<nb, kRB, 0, stream>>>(io, certain, blk_a, snap, io.alive_hist_dev ? io.alive_hist_dev + io.step_index : nullptr);
  k_store<<<nb, kRB, 0, stream>>>(io, certain, blk_a, keep, blk_b);
  if (io.mode == 2 && io.n_members > 1) k_store_payload<<<dim3(nb, io.n_members), kRB, 0, stream>>>(io, certain, snap);
  k_compact_dev<<<dim3(nb, (io.mode == 2 && io.n_members > 1) ? io.n_members : 1), kRB, 0, stream>>>(io, keep, blk_b, snap);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { snprintf(g_roll_err, sizeof(g_roll_err), "rollout_advance: %s", cudaGetErrorString(e)); return CMBPO_ERR_CUDA; }
  return CMBPO_OK;
}

extern "C" int cmbpo_stub_policy(const void* obs_dev, const void* P_dev, void* act_dev, const void* n_rows_dev,
                                 int64_t n_rows, int32_t obs_dim, int32_t act_dim, int32_t n_slots, void* stream_) {
  if (!obs_dev || !P_dev || !act_dev || n_rows < 0) return CMBPO_ERR_ARG;
  if (n_rows == 0) return CMBPO_OK;
  if (n_slots < 1) n_slots = 1;
  long long blocks = ((n_rows + kPolRows - 1) / kPolRows) * n_slots;
  if (blocks > 148 * 6) blocks = 148 * 6;
  const size_t smem = ((size_t)obs_dim * ((act_dim + 3) & ~3) + (size_t)kPolRows * (obs_dim + 1 + act_dim)) * sizeof(float);
  if (smem > 200 * 1024) return CMBPO_ERR_UNSUPPORTED;
  static size_t smem_set = 48 * 1024;
  if (smem > smem_set) {
    if (cudaFuncSetAttribute(k_stub_policy, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return CMBPO_ERR_CUDA;
    smem_set = smem;
  }
  k_stub_policy<<<(int)blocks, kPolRows, smem, (cudaStream_t)stream_>>>(
      static_cast<const float*>(obs_dev), static_cast<const float*>(P_dev), static_cast<float*>(act_dev),
      static_cast<const int*>(n_rows_dev), n_rows, obs_dim, act_dim, n_slots);
  return cudaGetLastError() == cudaSuccess ? CMBPO_OK : CMBPO_ERR_CUDA;
}

// ---------------------------------------------------------------------------------------
// Packing of the populated transitions of a model pool into one dense record matrix (what
// ModelBuffer.get selects with populated_mask, softlearning/replay_pools/modelbuffer.py:348-405),
// in [path, t] order: the send buffer of the multi-GPU all-gather.
//   record = obs[E*OD] | act[E*AD] | next_obs[E*OD] | rew[E] | cost[E] | dyn_error | term[E] | path | t
// (E = 1 unless the pool is an 'iv_gae' pool; path = global start-state index and t are int32 bit
// patterns in the last two fp32 slots).  populated_mask rows are prefixes by construction (a finished
// path never stores again, modelbuffer.py:110-114), so the k-th record of a path is its column k.
// ---------------------------------------------------------------------------------------
namespace cmbpo {

__global__ void __launch_bounds__(kRB) k_pack_count(cmbpo_pack_io io, int* path_off, int* blk_cnt, int* path_len) {
  const long long b = (long long)blockIdx.x * kRB + threadIdx.x;
  const int T = io.max_path_length;
  int len = 0;
  if (b < io.n_paths) {
    const uint8_t* m = io.pool_populated_dev + (size_t)b;   // storage [T][n_paths]
    for (int t = 0; t < T; ++t) len += m[(size_t)t * io.n_paths] ? 1 : 0;
  }
  int tot;
  const int off = block_exclusive_scan_int(len, &tot);
  if (b < io.n_paths) { path_off[b] = off; path_len[b] = len; }
  if (threadIdx.x == 0) blk_cnt[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) k_pack_scan(int* blk_cnt, int n_blocks, long long* blk_off, long long* count_out) {
  __shared__ long long sh[1024];
  __shared__ long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n_blocks; base += 1024) {
    const int i = base + threadIdx.x;
    const long long v = (i < n_blocks) ? blk_cnt[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
      const long long t = (threadIdx.x >= off) ? sh[threadIdx.x - off] : 0;
      __syncthreads();
      sh[threadIdx.x] += t;
      __syncthreads();
    }
    const long long incl = sh[threadIdx.x];
    if (i < n_blocks) blk_off[i] = carry + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry += incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *count_out = carry;
}

constexpr int kPackSub = 4;              // write blocks per 256-path scan block (more blocks in flight: the
constexpr int kPackPaths = kRB / kPackSub;   // per-record chain lookup -> loads -> store is latency-bound)
__global__ void __launch_bounds__(kRB) k_pack_write(cmbpo_pack_io io, const int* path_off, const int* blk_cnt,
                                                    const long long* blk_off, int W) {
  __shared__ int s_off[kPackPaths + 1];
  const int blk = blockIdx.x / kPackSub, sub = blockIdx.x % kPackSub;
  const long long b0 = (long long)blk * kRB + (long long)sub * kPackPaths;
  const int blk_total = blk_cnt[blk];
  if (threadIdx.x <= kPackPaths) {
    const long long bp = b0 + threadIdx.x;
    const bool in_blk = (sub * kPackPaths + (int)threadIdx.x) < kRB && bp < io.n_paths;
    s_off[threadIdx.x] = in_blk ? path_off[bp] : blk_total;
  }
  __syncthreads();
  const int rec_lo = s_off[0], total = s_off[kPackPaths];
  const long long out0 = blk_off[blk];
  if (out0 + blk_total > io.capacity) return;   // caller sized the buffer too small: count_dev still reports the need
  const int E = io.n_members > 1 ? io.n_members : 1;
  const int OD = io.obs_dim, AD = io.act_dim, T = io.max_path_length;
  // column boundaries of the record; compact records have an empty next_obs block (c1 == c2)
  const int c0 = E * OD, c1 = c0 + E * AD, c2 = c1 + (io.compact ? 0 : E * OD), c3 = c2 + E, c4 = c3 + E, c5 = c4 + 1, c6 = c5 + E;
  const size_t B = (size_t)io.n_paths;
  float* out = io.records_dev + (size_t)out0 * W;
  // one warp per record (the path lookup is warp-uniform), lanes stride over the record's columns; kU records
  // are in flight per warp so kU independent global loads overlap (a single chain ran at ~0.9 TB/s)
  constexpr int kU = 4, kWarps = kRB / 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int rec0 = rec_lo + warp; rec0 < total; rec0 += kWarps * kU) {
    int tt[kU];
    size_t bb[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int rec = min(rec0 + u * kWarps, total - 1);
      int lo = 0, hi = kPackPaths;   // last path with s_off[path] <= rec (paths without records share their successor's offset)
      while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (s_off[mid] <= rec) lo = mid; else hi = mid; }
      tt[u] = rec - s_off[lo];
      bb[u] = (size_t)(b0 + lo);
    }
    for (int col = lane; col < W; col += 32) {
      // column -> (field base pointer, element stride decode) is the same for all kU records
      const float* base = nullptr;
      const uint8_t* tbase = nullptr;
      int e = 0, d = 0, width = 1, kind = 0;   // kind 0: float field, 1: term byte, 2: path id, 3: t
      if (col < c0) { e = (E == 1) ? 0 : col / OD; d = col - e * OD; width = OD; base = io.pool_obs_dev; }
      else if (col < c1) { const int k = col - c0; e = (E == 1) ? 0 : k / AD; d = k - e * AD; width = AD; base = io.pool_act_dev; }
      else if (col < c2) { const int k = col - c1; e = (E == 1) ? 0 : k / OD; d = k - e * OD; width = OD; base = io.pool_next_obs_dev; }
      else if (col < c3) { e = col - c2; base = io.pool_rew_dev; }
      else if (col < c4) { e = col - c3; base = io.pool_cost_dev; }
      else if (col < c5) { base = io.pool_dyn_error_dev; }
      else if (col < c6) { e = col - c5; tbase = io.pool_term_dev; kind = 1; }
      else kind = (col == c6) ? 2 : 3;
      float v[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const size_t cell = ((size_t)e * T + tt[u]) * B + bb[u];   // storage [E][T][n_paths]
        if (kind == 0) v[u] = __ldg(base + cell * width + d);
        else if (kind == 1) v[u] = __ldg(tbase + cell) ? 1.0f : 0.0f;
        else if (kind == 2) v[u] = __int_as_float((int)(io.row_offset + (long long)bb[u]));
        else v[u] = __int_as_float(tt[u]);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int rec = rec0 + u * kWarps;
        if (rec < total) out[(size_t)rec * W + col] = v[u];
      }
    }
  }
}
}  // namespace cmbpo

namespace cmbpo {
// compact records: next_obs of every path's last stored transition, [n_paths][E * OD] (zeros for empty paths)
__global__ void __launch_bounds__(256) k_pack_last(cmbpo_pack_io io, const int* path_len) {
  const int E = io.n_members > 1 ? io.n_members : 1, OD = io.obs_dim, T = io.max_path_length;
  const long long per = (long long)E * OD, total = io.n_paths * per;
  const size_t B = (size_t)io.n_paths;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long b = i / per;
    const int j = (int)(i - b * per), e = j / OD, d = j - e * OD;
    const int L = path_len[b];
    io.last_next_obs_dev[i] = L > 0 ? __ldg(io.pool_next_obs_dev + (((size_t)e * T + (L - 1)) * B + (size_t)b) * OD + d) : 0.f;
  }
}
}  // namespace cmbpo

extern "C" int32_t cmbpo_record_width_compact(int32_t obs_dim, int32_t act_dim, int32_t n_members) {
  const int E = n_members > 1 ? n_members : 1;
  return E * (obs_dim + act_dim + 3) + 1 + 2;
}

extern "C" int32_t cmbpo_record_width(int32_t obs_dim, int32_t act_dim, int32_t n_members) {
  const int E = n_members > 1 ? n_members : 1;
  return E * (2 * obs_dim + act_dim + 3) + 1 + 2;
}

extern "C" size_t cmbpo_pack_scratch_bytes(int64_t n_paths) {
  const size_t nb = (size_t)((n_paths + kRB - 1) / kRB) + 1;
  return (size_t)n_paths * 2 * sizeof(int) + nb * (sizeof(int) + sizeof(long long)) + 256;
}

extern "C" int cmbpo_pack_records(const cmbpo_pack_io* io_, void* stream_) {
  if (!io_ || !io_->records_dev || !io_->count_dev || !io_->scratch_dev || !io_->pool_populated_dev) return CMBPO_ERR_ARG;
  const cmbpo_pack_io io = *io_;
  if (io.n_paths < 0 || io.max_path_length <= 0) return CMBPO_ERR_ARG;
  if (io.scratch_bytes < cmbpo_pack_scratch_bytes(io.n_paths)) return CMBPO_ERR_WORKSPACE;
  cudaStream_t stream = (cudaStream_t)stream_;
  const int nb = (int)((io.n_paths + kRB - 1) / kRB);
  // scratch: [nb+1] int64 block offsets | [n_paths] int path offsets | [nb] int block counts
  long long* blk_off = static_cast<long long*>(io.scratch_dev);
  int* path_off = reinterpret_cast<int*>(blk_off + nb + 1);
  int* path_len = path_off + io.n_paths;
  int* blk_cnt = path_len + io.n_paths;
  if (io.compact && !io.last_next_obs_dev) return CMBPO_ERR_ARG;
  const int W = io.compact ? cmbpo_record_width_compact(io.obs_dim, io.act_dim, io.n_members)
                           : cmbpo_record_width(io.obs_dim, io.act_dim, io.n_members);
  if (nb > 0) k_pack_count<<<nb, kRB, 0, stream>>>(io, path_off, blk_cnt, path_len);
  k_pack_scan<<<1, 1024, 0, stream>>>(blk_cnt, nb, blk_off, reinterpret_cast<long long*>(io.count_dev));
  if (nb > 0) k_pack_write<<<nb * kPackSub, kRB, 0, stream>>>(io, path_off, blk_cnt, blk_off, W);
  if (nb > 0 && io.compact) {
    const long long tot = io.n_paths * (long long)(io.n_members > 1 ? io.n_members : 1) * io.obs_dim;
    long long gb = (tot + 255) / 256;
    if (gb > 148 * 16) gb = 148 * 16;
    k_pack_last<<<(int)gb, 256, 0, stream>>>(io, path_len);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { snprintf(g_roll_err, sizeof(g_roll_err), "pack_records: %s", cudaGetErrorString(e)); return CMBPO_ERR_CUDA; }
  return CMBPO_OK;
}
