// step_post.cu -- per-row post-processing of the ensemble heads: everything FakeEnv.step does
// after BNN.predict returns (mbpo/models/fake_env.py:156-222), plus the transposing
// "finalize" used by the plain BNN.predict boundary.
//
// One thread per rollout row.  The heads arrive tile-major (cmbpo_internal.cuh head_index; written coalesced by
// ensemble_fwd.cu), so every load below is a fully coalesced 128-byte warp access.  Row-major
// tensors (obs in, next_obs out) go through a shared-memory tile so global traffic stays
// coalesced too.
#include <cstdlib>
#include "cmbpo_internal.cuh"
#include "step_post.cuh"

namespace cmbpo {

// ---------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011), counter = (row, stream, 0, 0), key = seed
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t (&out)[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__device__ __forceinline__ float u32_to_unit(uint32_t x) {  // (0,1]
  return ((float)(x >> 8) + 1.0f) * (1.0f / 16777216.0f);
}

// standard normal from one Philox draw: Box-Muller on two 32-bit words (u1 in (0,1], u2 in (0,1])
__device__ __forceinline__ float philox_normal(unsigned long long row, uint32_t col, uint32_t stream, unsigned long long seed) {
  uint32_t rnd[4];
  philox4x32_10((uint32_t)row, (uint32_t)(row >> 32), stream, 0x9E3779B9u ^ col, (uint32_t)seed, (uint32_t)(seed >> 32), rnd);
  const float u1 = u32_to_unit(rnd[0]), u2 = u32_to_unit(rnd[1]);
  return sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
}

// softplus(x) = log(1 + e^x), stable for large |x| (np.logaddexp(0, x))
__device__ __forceinline__ float softplusf(float x) { return x > 0.f ? x + log1pf(expf(-x)) : log1pf(expf(x)); }

// np.clip(x, 0, 1e10) including NumPy's NaN propagation (mbpo/models/utils.py:24)
__device__ __forceinline__ float clip_kl_exact(float x) {
  const float r = fminf(fmaxf(x, 0.f), 1e10f);
  return (x != x) ? x : r;
}

// static termination / cost functions, mbpo/static/<env>.py.  `x` = next_obs features the
// functions look at; comparisons are done in fp32 exactly as NumPy does on fp32 arrays.
struct RowFeatures {
  float x0, x1, x2, x3, last;
  bool finite_all;      // np.isfinite(next_obs).all(-1)
  bool lt100_from1;     // (next_obs[..., 1:] < 100).all(-1)   (hopper, quirk Q2)
};

__device__ __forceinline__ bool termination(int env, const RowFeatures& f) {
  switch (env) {
    case ENV_ANT: {  // ant.py:9-14
      const bool nd = f.finite_all && (f.x0 >= 0.2f) && (f.x0 <= 1.0f);
      return !nd;
    }
    case ENV_ANTSAFE: {  // antsafe.py:9-21 -- precedence quirk Q1: (a*b*c*z_rot) >= -0.7
      const float zr = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(__fmul_rn(f.x2, f.x2), __fmul_rn(f.x3, f.x3))));
      const float gate = (f.finite_all && (f.x0 >= 0.2f) && (f.x0 <= 1.0f)) ? 1.0f : 0.0f;
      const bool nd = __fmul_rn(gate, zr) >= -0.7f;
      return !nd;
    }
    case ENV_HUMANOID:
    case ENV_HUMANOIDSAFE:  // humanoid.py:11-12
      return (f.x0 < 1.0f) || (f.x0 > 2.0f);
    case ENV_HOPPER:
    case ENV_HOPPERSAFE: {  // hopper.py:9-15
      const bool nd = f.finite_all && f.lt100_from1 && (f.x0 > 0.7f) && (fabsf(f.x1) < 0.2f);
      return !nd;
    }
    case ENV_WALKER2D: {  // walker2d.py:9-14
      const bool nd = (f.x0 > 0.8f) && (f.x0 < 2.0f) && (f.x1 > -1.0f) && (f.x1 < 1.0f);
      return !nd;
    }
    default:  // halfcheetah.py:9, halfcheetahsafe.py:9: never terminates
      return false;
  }
}
__device__ __forceinline__ float static_cost(int env, const RowFeatures& f, bool done) {
  switch (env) {
    case ENV_ANTSAFE: {  // antsafe.py:32-46
      const float obj = (fabsf(f.last) > 3.2f) ? 1.0f : 0.0f;
      return fminf(fmaxf((done ? 1.0f : 0.0f) + obj, 0.f), 1.f);
    }
    case ENV_HUMANOIDSAFE: {  // humanoidsafe.py:21-33
      const float obj = (fabsf(f.last) > 3.0f) ? 1.0f : 0.0f;
      return fminf(fmaxf((done ? 1.0f : 0.0f) + obj, 0.f), 1.f);
    }
    case ENV_HALFCHEETAHSAFE:  // halfcheetahsafe.py:17-19
      return (fabsf(__fmul_rn(f.last, 10.0f)) < 2.0f) ? 1.0f : 0.0f;
    case ENV_HOPPERSAFE:  // hoppersafe.py:24-25
      return (fabsf(f.last) < 1.0f) ? 1.0f : 0.0f;
    default:
      return 0.0f;
  }
}

template <int E>
__device__ __forceinline__ float pick(const float (&m)[E], int idx) {
  float r = m[0];
#pragma unroll
  for (int e = 1; e < E; ++e) r = (idx == e) ? m[e] : r;
  return r;
}

// e^x as one FMUL + MUFU.EX2: the .ftz form skips the denormal-input scaling (two FMULs and a compare per call) that
// __expf emits in a non-fast-math build; results below 2^-126 flush to zero, where var + 1e-10 is 1e-10 either way
__device__ __forceinline__ float exp_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {   // one MUFU.RCP, no range fix-up branch
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// One column d of one row, reference order of operations: all-pairs KL with the per-pair clip to [0, 1e10]
// (utils.py:14-57, quirk Q5), var_E(mean) (fake_env.py:160).  Taken by columns with non-finite / extreme
// heads and by columns where a pair could reach the upper clip; everything else goes through
// column_stats_fast below.
template <int E>
__device__ __noinline__ void column_stats_exact(const float (&m)[E], const float (&lv)[E], float (&v)[E],
                                                float& kl_out, float& epv_out) {
  float ls[E], vk[E];
#pragma unroll
  for (int e = 0; e < E; ++e) {
    v[e] = __expf(lv[e]);                       // bnn.py:898
    // log_std = clip(log(sqrt(var)), -100, 1e8); var is recomputed from it (utils.py:20,48-49)
    const float l = logf(sqrtf(v[e]));
    ls[e] = fminf(fmaxf(l, -100.f), 1e8f);
    if (l != l) ls[e] = l;
    vk[e] = expf(2.0f * ls[e]);
  }
  float kl = 0.f;
#pragma unroll
  for (int i = 0; i < E; ++i) {
#pragma unroll
    for (int j = 0; j < E; ++j) {
      const float d = m[j] - m[i];
      const float pre = 0.5f * ((d * d + vk[i]) / (vk[j] + 1e-10f) - 1.0f) + ls[j] - ls[i];
      kl += clip_kl_exact(pre);
    }
  }
  kl_out = kl / ((float)(E * (E - 1)) + 1e-10f);
  float mean = 0.f;
#pragma unroll
  for (int e = 0; e < E; ++e) mean += m[e];
  mean *= (1.0f / E);
  float s2 = 0.f;
#pragma unroll
  for (int e = 0; e < E; ++e) { const float t = m[e] - mean; s2 = fmaf(t, t, s2); }
  epv_out = s2 * (1.0f / E);
}

// Call site of the exact path: it works on copies, so only this (rare) branch puts the arrays on the stack.
template <int E>
__device__ __forceinline__ void column_stats_rare(const float (&m)[E], const float (&lv)[E], float (&v)[E],
                                                  float& kl_out, float& epv_out) {
  float mc[E], lc[E], vc[E];
#pragma unroll
  for (int e = 0; e < E; ++e) { mc[e] = m[e]; lc[e] = lv[e]; }
  column_stats_exact<E>(mc, lc, vc, kl_out, epv_out);
#pragma unroll
  for (int e = 0; e < E; ++e) v[e] = vc[e];
}

// The same statistics in O(E) for finite, moderate heads.  Summed over all ordered pairs i != j the
// log-std terms of KL(i||j) cancel, and with dl_i = mu_i - mean, S_c = sum dl_i^2, S_v = sum var_i:
//     sum_{i != j} (var_i + (mu_i - mu_j)^2) = S_v - var_j + S_c + E * dl_j^2          for each j
// so  sum_{i != j} KL(i||j) = sum_j (S_v - var_j + S_c + E dl_j^2) / (2 (var_j + EPS)) - E (E - 1) / 2.
// The differences are taken around the mean, so nearly identical members do not cancel catastrophically
// (same fp32 error class as the pairwise form, whose terms are "ratio/2 - 1/2" as well).  The per-pair clip
// of the reference cannot be applied here: the lower one only ever removes rounding noise (KL >= 0), and the
// function returns false -- the caller then takes column_stats_exact -- when some pair could come within a
// factor 10 of the upper one.
template <int E>
__device__ __forceinline__ bool column_stats_fast(const float (&m)[E], const float (&lv)[E], float (&v)[E],
                                                  float& kl_out, float& epv_out) {
  float inv[E], dl[E];
  float sv = 0.f, mean = 0.f, lvmin = lv[0], lvmax = lv[0];
#pragma unroll
  for (int e = 0; e < E; ++e) {
    v[e] = exp_fast(lv[e]);                     // bnn.py:898
    inv[e] = rcp_approx(v[e] + 1e-10f);
    sv += v[e];
    mean += m[e];
    lvmin = fminf(lvmin, lv[e]);
    lvmax = fmaxf(lvmax, lv[e]);
  }
  mean *= (1.0f / E);
  float s2 = 0.f, amax = 0.f;
#pragma unroll
  for (int e = 0; e < E; ++e) {
    dl[e] = m[e] - mean;
    s2 = fmaf(dl[e], dl[e], s2);
    amax = fmaxf(amax, fabsf(dl[e]));
  }
  const float d2max = amax * amax;
  const float base = sv + s2;
  float acc = -0.5f * (float)(E * (E - 1));
#pragma unroll
  for (int j = 0; j < E; ++j) {
    const float t = fmaf((float)E * dl[j], dl[j], base - v[j]);
    acc = fmaf(0.5f * inv[j], t, acc);
  }
  kl_out = fmaxf(acc, 0.f) * (1.0f / ((float)(E * (E - 1)) + 1e-10f));   // within 1 ulp of the division
  epv_out = s2 * (1.0f / E);
  // Guard (the caller takes the exact path when it fails): log-variances inside (-80, 80) -- fminf / fmaxf drop NaNs, the
  // comparisons below do not pass them; means finite and moderate (a NaN / Inf / > ~1e15 mean makes s2 NaN or huge); and
  // the largest possible pair term ((mu_i - mu_j)^2 + var_i) / (2 var_j) <= (4 d2max + vmax) * invmax / 2 below 1e9,
  // with vmax / invmax taken from the extreme log-variances (exp is monotone).
  const float vmax = exp_fast(lvmax), invmax = rcp_approx(exp_fast(lvmin) + 1e-10f);
  bool ok = (lvmin > -80.f) && (lvmax < 80.f) && (s2 < 1e30f);
#pragma unroll
  for (int e = 0; e < E; ++e) ok = ok && (lv[e] == lv[e]);     // a NaN log-variance would slip through fminf / fmaxf
  return ok && (4.f * d2max + vmax) * invmax < 1e9f;
}

// kPostParts warps share a block of kPostRows rows: warp `part` handles output columns d = part,
// part+kPostParts, ... of all 32 rows (lane = row), so head loads stay 128-byte coalesced while the serial
// per-thread chain -- the latency floor of the small sampler steps late in a rollout -- is a few columns
// long, and a small batch still spreads over all SMs (32-row blocks).  8 parts (4 blocks of 256 threads per
// SM) measured 5 % faster than 16 (2 x 512): half as many warps wait at each of the block's barriers.
constexpr int kPostRows = 32;
#ifndef CMBPO_POST_PARTS
#define CMBPO_POST_PARTS 8
#endif
constexpr int kPostParts = CMBPO_POST_PARTS;            // column parts of the 3-D kernel
constexpr int kPostThreads = kPostRows * kPostParts;
constexpr int kPostMaxParts = 10;                       // 2-D kernel: the launcher picks 6..10 parts so that D splits evenly

// Block partial sums of the four scalar diagnostics (fake_env.py:213-220); the last block to finish sums
// the partials in block order (deterministic) and writes the 4 means.
template <int E>
__device__ __forceinline__ void finish_scalars(const StepParams& p, const float (&vals)[4], long long n_rows,
                                               int n_slots_out) {
  __shared__ float red[4][kPostMaxParts];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float x = vals[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) red[k][warp] = x;
  }
  __syncthreads();
  if (tid < 4) {
    float x = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) x += red[tid][w];
    p.partials[(size_t)blockIdx.x * 4 + tid] = x;
  }
  // ---- the last block to finish sums the partials in block order (deterministic) -> 4 means
  __shared__ bool s_last;
  __shared__ double sh[4][kPostMaxParts];
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = atomicAdd(p.ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double s[4] = {0, 0, 0, 0};
  for (int b = tid; b < (int)gridDim.x; b += (int)blockDim.x)
    for (int k = 0; k < 4; ++k) s[k] += (double)__ldcg(p.partials + (size_t)b * 4 + k);
  for (int k = 0; k < 4; ++k) {
    double x = s[k];
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) sh[k][warp] = x;
  }
  __syncthreads();
  if (tid < 4) {
    double x = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) x += sh[tid][w];
    const double n = (double)n_rows;
    double denom = n;                                        // dkl_mean
    if (tid == 1) denom = n * (double)E * (double)p.D;       // ensemble_mean_var over [E,B,D]
    if (tid >= 2) denom = n * (double)n_slots_out;           // rew_mean / cost_mean
    p.scalars[tid] = (float)(denom > 0 ? x / denom : 0.0);
  }
  if (tid == 0) *p.ticket = 0;   // ready for the next launch
}

// 16-byte asynchronous global -> shared copy (LDGSTS, L1 bypass); completion through commit / wait groups
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Head staging: the E * 2D head lines of a 32-row block (128 bytes each, one per (member, column), inside one 128-row
// tile of the tile-major scratch) come in with 16-byte asynchronous copies, issued for row block k+1 as soon as the
// column arithmetic of block k has finished -- the loads are all in flight at once (no registers held) and overlap the
// block's termination / store tail and the other resident blocks' arithmetic.  The direct-from-global version of this
// kernel ran at 2.3-2.6 TB/s with load latency and the block barriers as its top stalls (profiles/r01_post_1M_*).
template <int E, bool OPT>   // OPT: the beyond-reference options (sample_mode / bound_logvar) are compiled in
__global__ void __launch_bounds__(kPostMaxParts * kPostRows, 3) step_post_kernel(const StepParams p) {
  extern __shared__ __align__(16) float dyn_smem[];
  float* sH = dyn_smem;                                   // [E * 2D][kPostRows] head lines of the current row block
  float* tile = dyn_smem + (size_t)E * 2 * p.D * kPostRows;   // [kPostRows * obs_dim + 32] obs in / next_obs out staging
  __shared__ float s_kl[kPostMaxParts][kPostRows], s_epv[kPostMaxParts][kPostRows];
  __shared__ float s_rew[kPostRows];
  __shared__ int s_eidx[kPostRows];
  __shared__ uint8_t s_flags[kPostMaxParts][kPostRows];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r = lane, part = warp;
  const int n_parts = (int)(blockDim.x >> 5), n_threads = (int)blockDim.x;   // launch_step_post_EO sizes the block to D
  const long long n_rows = p.n_rows_dev ? (long long)(*p.n_rows_dev) : p.n_rows;
  const unsigned long long rng_off = p.rng_offset + (p.rng_offset_dev ? *p.rng_offset_dev : 0ull);
  // 3-D mode: member slot e of obs/act/next_obs/rew/term/cost starts at row e * capacity (the array shapes
  // stay at capacity while the live row count shrinks during an 'iv_gae' rollout)
  const long long slot_rows = p.n_rows;
  const int D = p.D, OD = p.obs_dim;
  const int n_slots_out = p.per_slot ? E : 1;
  float acc_dkl = 0.f, acc_var = 0.f, acc_rew = 0.f, acc_cost = 0.f;

  const int n_lines = E * 2 * D;
  // head line L of the row block at row0: scratch element (row0 >> 7) * E*2D*128 + L * 128 + (row0 & 127), 32 floats
  auto issue_heads = [&](long long row0) {
    const float* src = p.heads + (size_t)(row0 >> 7) * (size_t)(n_lines * kTileRows) + (size_t)(row0 & (kTileRows - 1));
    for (int i = tid; i < n_lines * 8; i += n_threads) {
      const int L = i >> 3, q = i & 7;
      cp_async16(sH + L * kPostRows + q * 4, src + (size_t)L * kTileRows + q * 4);
    }
    cp_async_commit();
  };
  const long long row_step = (long long)gridDim.x * kPostRows;
  if ((long long)blockIdx.x * kPostRows < n_rows) issue_heads((long long)blockIdx.x * kPostRows);
  for (long long row0 = (long long)blockIdx.x * kPostRows; row0 < n_rows; row0 += row_step) {
    const long long g = row0 + r;
    const bool live = g < n_rows;
    const float* hrow = sH + r;   // member e, column j of this row at hrow[(e * 2D + j) * kPostRows]
    const int rows_here = (int)min((long long)kPostRows, n_rows - row0);
    int eidx = 0;
    if (!p.per_slot && live && part == 0) {   // one warp draws the members; the others pick them up from smem below
      if (p.elite_idx) {
        eidx = p.elite_idx[g];
      } else {  // random elite per row (bnn.py:437-439), counter-based so sharding cannot change it
        uint32_t rnd[4];
        const unsigned long long ctr = (unsigned long long)(p.row_offset + (p.row_ids ? (long long)p.row_ids[g] : g));
        philox4x32_10((uint32_t)ctr, (uint32_t)(ctr >> 32), (uint32_t)rng_off, 0u,
                      (uint32_t)p.seed, (uint32_t)(p.seed >> 32), rnd);
        eidx = p.elites[__umulhi(rnd[0], (uint32_t)p.n_elites)];
      }
      s_eidx[r] = eidx;
    }
    float kl_row = 0.f, epv_row = 0.f;
    // ---- pass over output slots (1 in 2-D mode; E in 3-D mode where every member keeps its own row)
    for (int so = 0; so < n_slots_out; ++so) {
      // stage obs rows of this block (slot `so` in 3-D mode) through smem, coalesced
      __syncthreads();
      {
        const float* src = p.obs + (p.per_slot ? (size_t)so * slot_rows * OD : 0) + (size_t)row0 * OD;
        for (int i = tid; i < rows_here * OD; i += n_threads) tile[i] = src[i];
      }
      cp_async_wait_all();     // this thread's share of the head lines has landed; the barrier publishes everybody's
      __syncthreads();
      if (!p.per_slot && live) eidx = s_eidx[r];
      bool finite_part = true, lt100_part = true;
      if (live) {
        for (int d = part; d < D; d += n_parts) {
          float m[E], lv[E], v[E];
          const float o = (d < OD) ? tile[r * OD + d] : 0.f;
#pragma unroll
          for (int e = 0; e < E; ++e) {
            m[e] = hrow[(e * 2 * D + d) * kPostRows];
            lv[e] = hrow[(e * 2 * D + D + d) * kPostRows];
          }
          if (d < OD) {
            if (p.per_slot) {
              // 3-D: member e's mean gets member e's own observation (fake_env.py:156)
              // obs of the other slots are read straight from global (L2-resident, just loaded)
#pragma unroll
              for (int e = 0; e < E; ++e)
                m[e] += (e == so) ? o : p.obs[((size_t)e * slot_rows + g) * OD + d];
            } else {
#pragma unroll
              for (int e = 0; e < E; ++e) m[e] += o;
            }
          }
          if (OPT && p.bound_logvar) {   // optional PETS bounding on the raw head (bnn.py:888-889, bnn_chua.py:400-401)
            const float sh = p.head_shift ? p.head_shift[D + d] : 0.f;
#pragma unroll
            for (int e = 0; e < E; ++e) {
              float raw = lv[e] - sh;
              raw = p.max_logvar - softplusf(p.max_logvar - raw);
              raw = p.min_logvar + softplusf(raw - p.min_logvar);
              lv[e] = raw + sh;
            }
          }
          float kl, epv;
          if (!column_stats_fast<E>(m, lv, v, kl, epv)) column_stats_rare<E>(m, lv, v, kl, epv);
          if (so == 0) {
            kl_row += kl; epv_row += epv;
            float sv = 0.f;
#pragma unroll
            for (int e = 0; e < E; ++e) sv += v[e];
            acc_var += sv;
            if (p.ep_var) p.ep_var[(size_t)g * D + d] = epv;
          }
          float x = p.per_slot ? pick<E>(m, so) : pick<E>(m, eidx);
          if (p.noise || (OPT && p.sample_mode == 1)) {  // optional north_star mode: mean + N(0,1)*std (off in the live reference)
            const float vs = p.per_slot ? pick<E>(v, so) : pick<E>(v, eidx);
            const unsigned long long grow = (unsigned long long)(p.row_offset + (p.row_ids ? (long long)p.row_ids[g] : g));
            const float z = p.noise ? p.noise[(p.per_slot ? (size_t)so * slot_rows * D : 0) + (size_t)g * D + d]
                                    : philox_normal(grow, (uint32_t)(so * 4096 + d + 1), (uint32_t)rng_off, p.seed);
            x = fmaf(z, sqrtf(vs), x);
          }
          if (d < OD) {
            tile[r * OD + d] = x;  // own element only: safe to overwrite the staged obs
            finite_part = finite_part && (fabsf(x) <= 3.402823466e38f);
            if (d >= 1) lt100_part = lt100_part && (x < 100.f);
          } else {
            s_rew[r] = x;  // reward is the last output column (fake_env.py:180)
          }
        }
      }
      s_flags[part][r] = (uint8_t)((finite_part ? 1 : 0) | (lt100_part ? 2 : 0));
      __syncthreads();
      // every warp is done with the head lines of this block: start fetching the next block's
      if (so == n_slots_out - 1 && row0 + row_step < n_rows) issue_heads(row0 + row_step);
      if (live && part == 0) {
        RowFeatures f;
        uint8_t fl = 3;
#pragma unroll
        for (int q = 0; q < n_parts; ++q) fl &= s_flags[q][r];
        f.finite_all = (fl & 1) != 0;
        f.lt100_from1 = (fl & 2) != 0;
        const float* xr = tile + r * OD;
        f.x0 = xr[0];
        f.x1 = OD > 1 ? xr[1] : 0.f;
        f.x2 = OD > 2 ? xr[2] : 0.f;
        f.x3 = OD > 3 ? xr[3] : 0.f;
        f.last = xr[OD - 1];
        const float rew = s_rew[r];
        const bool done = termination(p.env_id, f);
        const float cost = (p.cost_mode == 1) ? static_cost(p.env_id, f, done) : 0.f;
        const size_t oi = (size_t)so * slot_rows + g;
        p.rew[oi] = rew;
        p.term[oi] = done ? 1 : 0;
        if (p.cost) p.cost[oi] = cost;
        if (so == 0 || p.per_slot) { acc_rew += rew; acc_cost += cost; }
      }
      {
        float* dst = p.next_obs + (p.per_slot ? (size_t)so * slot_rows * OD : 0) + (size_t)row0 * OD;
        for (int i = tid; i < rows_here * OD; i += n_threads) dst[i] = tile[i];
      }
    }
    s_kl[part][r] = kl_row;
    s_epv[part][r] = epv_row;
    __syncthreads();
    if (live && part == 0) {
      float ks = 0.f, es = 0.f;
#pragma unroll
      for (int q = 0; q < n_parts; ++q) { ks += s_kl[q][r]; es += s_epv[q][r]; }
      const float dkl = ks / (float)D;  // np.mean(average_dkl(...), axis=-1), fake_env.py:158
      p.dkl_path[g] = dkl;
      if (p.ep_var_mean) p.ep_var_mean[g] = es / (float)D;
      acc_dkl += dkl;
    }
  }
  const float vals[4] = {acc_dkl, acc_var, acc_rew, acc_cost};
  finish_scalars<E>(p, vals, n_rows, n_slots_out);
}


// 3-D inputs ('iv_gae' rollouts, fake_env.py:147-153,:164-165): every member keeps its own sample, so a row
// has E observations in and E next observations out.  The column statistics (KL, var_E) are computed ONCE per
// (row, column) with member e's mean shifted by member e's own observation; all E member tiles live in
// (dynamic) shared memory so global traffic stays coalesced.  Warp `part` < E then evaluates the static
// termination / cost function of member `part`.
template <int E, bool OPT>
__global__ void __launch_bounds__(kPostThreads, 768 / kPostThreads) step_post3d_kernel(const StepParams p) {
  extern __shared__ float tiles[];                 // [E][kPostRows * OD]
  __shared__ float s_kl[kPostParts][kPostRows], s_epv[kPostParts][kPostRows];
  __shared__ float s_rew[E][kPostRows];
  __shared__ uint8_t s_fin[kPostParts][kPostRows], s_lt[kPostParts][kPostRows];   // bit e = member e
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r = lane, part = warp;
  const long long n_rows = p.n_rows_dev ? (long long)(*p.n_rows_dev) : p.n_rows;
  const unsigned long long rng_off = p.rng_offset + (p.rng_offset_dev ? *p.rng_offset_dev : 0ull);
  const long long slot_rows = p.n_rows;            // member stride = capacity
  const int D = p.D, OD = p.obs_dim;
  const int tile_elems = kPostRows * OD;
  float acc_dkl = 0.f, acc_var = 0.f, acc_rew = 0.f, acc_cost = 0.f;
  // 16-byte copies of the obs / next_obs tiles need every member slab 16-byte aligned (row0 * OD * 4 always is: 32 rows)
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(p.obs) | reinterpret_cast<uintptr_t>(p.next_obs)) & 15) == 0 &&
                      (((size_t)slot_rows * OD) & 3) == 0;
  const bool need_row_flags = p.env_id == ENV_ANT || p.env_id == ENV_ANTSAFE || p.env_id == ENV_HOPPER || p.env_id == ENV_HOPPERSAFE;

  for (long long row0 = (long long)blockIdx.x * kPostRows; row0 < n_rows; row0 += (long long)gridDim.x * kPostRows) {
    const long long g = row0 + r;
    const bool live = g < n_rows;
    // this row's entries of the tile-major head scratch (cmbpo_internal.cuh head_index): member e, column j at
    // hrow[(e * 2D + j) * kTileRows] -- 32-bit offsets, E * 2D * 128 elements per 128-row block
    const float* hrow = p.heads + (size_t)(g >> 7) * (size_t)(E * 2 * D * kTileRows) + (size_t)(g & (kTileRows - 1));
    const int rows_here = (int)min((long long)kPostRows, n_rows - row0);
    __syncthreads();
    const int n_el = rows_here * OD;
    if (vec_ok) {        // 16-byte cp.async straight into the tiles (the scalar loop below costs ~10 instructions per element)
      const int n4 = n_el >> 2;
#pragma unroll
      for (int e = 0; e < E; ++e) {
        const float* src = p.obs + ((size_t)e * slot_rows + row0) * OD;
        for (int q = tid; q < n4; q += kPostThreads) cp_async16(tiles + e * tile_elems + 4 * q, src + 4 * q);
      }
      cp_async_commit();
      const int tail = n_el & 3;
      if (tid < E * tail) {
        const int e = tid / tail, i = 4 * n4 + (tid - e * tail);
        tiles[e * tile_elems + i] = p.obs[((size_t)e * slot_rows + row0) * OD + i];
      }
      cp_async_wait_all();
    } else {
#pragma unroll
      for (int e = 0; e < E; ++e) {
        const float* src = p.obs + ((size_t)e * slot_rows + row0) * OD;
        for (int i = tid; i < n_el; i += kPostThreads) tiles[e * tile_elems + i] = src[i];
      }
    }
    __syncthreads();
    float kl_row = 0.f, epv_row = 0.f;
    uint32_t fin = (1u << E) - 1u, lt = (1u << E) - 1u;
    if (live) {
      // the 2E head values of the NEXT column of this warp are loaded before the current column is evaluated (register
      // double buffer): twice the loads in flight per thread, the kernel is bound by the latency of these strided reads
      float mn[E], lvn[E];
      if (part < D) {
#pragma unroll
        for (int e = 0; e < E; ++e) {
          mn[e] = hrow[(e * 2 * D + part) * kTileRows];
          lvn[e] = hrow[(e * 2 * D + D + part) * kTileRows];
        }
      }
      for (int d = part; d < D; d += kPostParts) {
        float m[E], lv[E], v[E];
#pragma unroll
        for (int e = 0; e < E; ++e) { m[e] = mn[e]; lv[e] = lvn[e]; }
        const int dn = d + kPostParts;
        if (dn < D) {
#pragma unroll
          for (int e = 0; e < E; ++e) {
            mn[e] = hrow[(e * 2 * D + dn) * kTileRows];
            lvn[e] = hrow[(e * 2 * D + D + dn) * kTileRows];
          }
        }
        if (d < OD) {
#pragma unroll
          for (int e = 0; e < E; ++e) m[e] += tiles[e * tile_elems + r * OD + d];   // fake_env.py:156
        }
        if (OPT && p.bound_logvar) {   // optional PETS bounding on the raw head (bnn.py:888-889, bnn_chua.py:400-401)
          const float sh = p.head_shift ? p.head_shift[D + d] : 0.f;
#pragma unroll
          for (int e = 0; e < E; ++e) {
            float raw = lv[e] - sh;
            raw = p.max_logvar - softplusf(p.max_logvar - raw);
            raw = p.min_logvar + softplusf(raw - p.min_logvar);
            lv[e] = raw + sh;
          }
        }
        float kl, epv;
        if (!column_stats_fast<E>(m, lv, v, kl, epv)) column_stats_rare<E>(m, lv, v, kl, epv);
        kl_row += kl; epv_row += epv;
        float sv = 0.f;
#pragma unroll
        for (int e = 0; e < E; ++e) sv += v[e];
        acc_var += sv;
        if (p.ep_var) p.ep_var[(size_t)g * D + d] = epv;
#pragma unroll
        for (int e = 0; e < E; ++e) {
          float x = m[e];
          if (p.noise) x = fmaf(p.noise[((size_t)e * slot_rows + g) * D + d], sqrtf(v[e]), x);
          else if (OPT && p.sample_mode == 1) {
            const unsigned long long grow = (unsigned long long)(p.row_offset + (p.row_ids ? (long long)p.row_ids[g] : g));
            x = fmaf(philox_normal(grow, (uint32_t)(e * 4096 + d + 1), (uint32_t)rng_off, p.seed), sqrtf(v[e]), x);
          }
          if (d < OD) {
            tiles[e * tile_elems + r * OD + d] = x;   // own element only
            if (need_row_flags) {                     // only the Ant / Hopper termination functions look at these
              if (!(fabsf(x) <= 3.402823466e38f)) fin &= ~(1u << e);
              if (d >= 1 && !(x < 100.f)) lt &= ~(1u << e);
            }
          } else {
            s_rew[e][r] = x;                           // reward = last output column (fake_env.py:180)
          }
        }
      }
    }
    s_fin[part][r] = (uint8_t)fin;
    s_lt[part][r] = (uint8_t)lt;
    s_kl[part][r] = kl_row;
    s_epv[part][r] = epv_row;
    __syncthreads();
    for (int so = part; so < E; so += kPostParts) {
      if (live) {
        RowFeatures f;
        uint32_t a = 0xffu, b = 0xffu;
#pragma unroll
        for (int q = 0; q < kPostParts; ++q) { a &= s_fin[q][r]; b &= s_lt[q][r]; }
        f.finite_all = ((a >> so) & 1u) != 0;
        f.lt100_from1 = ((b >> so) & 1u) != 0;
        const float* xr = tiles + so * tile_elems + r * OD;
        f.x0 = xr[0];
        f.x1 = OD > 1 ? xr[1] : 0.f;
        f.x2 = OD > 2 ? xr[2] : 0.f;
        f.x3 = OD > 3 ? xr[3] : 0.f;
        f.last = xr[OD - 1];
        const float rew = s_rew[so][r];
        const bool done = termination(p.env_id, f);
        const float cost = (p.cost_mode == 1) ? static_cost(p.env_id, f, done) : 0.f;
        const size_t oi = (size_t)so * slot_rows + g;
        p.rew[oi] = rew;
        p.term[oi] = done ? 1 : 0;
        if (p.cost) p.cost[oi] = cost;
        acc_rew += rew; acc_cost += cost;
      }
    }
    if (live && part == kPostParts - 1) {
      float ks = 0.f, es = 0.f;
#pragma unroll
      for (int q = 0; q < kPostParts; ++q) { ks += s_kl[q][r]; es += s_epv[q][r]; }
      const float dkl = ks / (float)D;
      p.dkl_path[g] = dkl;
      if (p.ep_var_mean) p.ep_var_mean[g] = es / (float)D;
      acc_dkl += dkl;
    }
    if (vec_ok) {
      const int n4 = n_el >> 2;
#pragma unroll
      for (int e = 0; e < E; ++e) {
        float* dst = p.next_obs + ((size_t)e * slot_rows + row0) * OD;
        for (int q = tid; q < n4; q += kPostThreads)
          *reinterpret_cast<float4*>(dst + 4 * q) = *reinterpret_cast<const float4*>(tiles + e * tile_elems + 4 * q);
      }
      const int tail = n_el & 3;
      if (tid < E * tail) {
        const int e = tid / tail, i = 4 * n4 + (tid - e * tail);
        p.next_obs[((size_t)e * slot_rows + row0) * OD + i] = tiles[e * tile_elems + i];
      }
    } else {
#pragma unroll
      for (int e = 0; e < E; ++e) {
        float* dst = p.next_obs + ((size_t)e * slot_rows + row0) * OD;
        for (int i = tid; i < n_el; i += kPostThreads) dst[i] = tiles[e * tile_elems + i];
      }
    }
  }
  const float vals[4] = {acc_dkl, acc_var, acc_rew, acc_cost};
  finish_scalars<E>(p, vals, n_rows, E);
}

// deterministic final reduction of the per-block partial sums -> 4 means
__global__ void reduce_scalars_kernel(const float* partials, int n_blocks, const int* n_rows_dev,
                                      long long n_rows_host, int E, int D, int n_slots_out, float* scalars) {
  __shared__ double sh[4][32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double s[4] = {0, 0, 0, 0};
  for (int b = tid; b < n_blocks; b += blockDim.x)
    for (int k = 0; k < 4; ++k) s[k] += (double)partials[(size_t)b * 4 + k];
  for (int k = 0; k < 4; ++k) {
    double x = s[k];
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) sh[k][warp] = x;
  }
  __syncthreads();
  if (tid < 4) {
    double x = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) x += sh[tid][w];
    const double n = (double)(n_rows_dev ? (long long)(*n_rows_dev) : n_rows_host);
    double denom = n;                                   // dkl_mean
    if (tid == 1) denom = n * (double)E * (double)D;    // ensemble_mean_var over [E,B,D]
    if (tid >= 2) denom = n * (double)n_slots_out;      // rew_mean / cost_mean
    scalars[tid] = (float)(denom > 0 ? x / denom : 0.0);
  }
}

// BNN.predict boundary: tile-major heads -> mean[E][B][D], var[E][B][D] row-major
// (bnn.py:713-717); var = exp(logvar) (bnn.py:898).  Non-probabilistic: mean only.
__global__ void finalize_predict_kernel(const float* __restrict__ heads, long long row_stride, long long n_rows,
                                        int E, int D, int probabilistic, float* mean, float* var) {
  const long long total = (long long)E * n_rows * D;
  const int n_head = probabilistic ? 2 * D : D;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(i % D);
    const long long g = (i / D) % n_rows;
    const int e = (int)(i / ((long long)D * n_rows));
    const float* h = heads + head_index(g, e, d, E, n_head);
    mean[i] = h[0];
    if (probabilistic && var) var[i] = __expf(h[(size_t)D * kTileRows]);
  }
}


// non-factored BNN.predict: ensemble mean [B][D]; variance = mean_E(var) + mean_E((mean-mean_E)^2) for
// probabilistic models (bnn.py:350-352), population var_E(mean) otherwise (bnn.py:736-743)
__global__ void aggregate_predict_kernel(const float* __restrict__ heads, long long row_stride, long long n_rows,
                                         int E, int D, int probabilistic, float* mean, float* var) {
  const long long total = n_rows * D;
  const int n_head = probabilistic ? 2 * D : D;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(i % D);
    const long long g = i / D;
    float m = 0.f;
    for (int e = 0; e < E; ++e) m += heads[head_index(g, e, d, E, n_head)];
    m /= (float)E;
    mean[i] = m;
    if (var) {
      float s2 = 0.f, sv = 0.f;
      for (int e = 0; e < E; ++e) {
        const float* h = heads + head_index(g, e, d, E, n_head);
        const float t = h[0] - m;
        s2 = fmaf(t, t, s2);
        if (probabilistic) sv += __expf(h[(size_t)D * kTileRows]);
      }
      var[i] = (probabilistic ? sv / (float)E : 0.f) + s2 / (float)E;
    }
  }
}

template <int E, bool OPT>
static cudaError_t launch_step_post_EO(const StepParams& p, int n_blocks, cudaStream_t stream) {
  if (p.per_slot && E > 1) {
    const size_t smem = (size_t)E * kPostRows * p.obs_dim * sizeof(float);
    static bool attr_set = false;   // same value for every launch of this instantiation (obs_dim <= 64)
    if (!attr_set) {
      cudaError_t e = cudaFuncSetAttribute(step_post3d_kernel<E, OPT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)((size_t)E * kPostRows * 64 * sizeof(float)));
      if (e != cudaSuccess) return e;
      attr_set = true;
    }
    step_post3d_kernel<E, OPT><<<n_blocks, kPostThreads, smem, stream>>>(p);
    return cudaGetLastError();
  }
  // 8 column parts (warps) per 32-row block unless that leaves the last column round mostly idle (HalfCheetah, D = 18:
  // 3 rounds of 8 at 75 % use -> 6 warps x 3 columns).  Measured at 1 M rows (tools/parts.sh, whole step): antsafe D = 30:
  // 6 / 8 / 10 parts = 2912 / 2863 / 2875 us; HalfCheetah: 6 / 9 = 2524 / 2573 us -- balanced columns (10 x 3, 9 x 2) do
  // not pay for the extra per-warp overhead per row block.
  const int rounds = (p.D + kPostParts - 1) / kPostParts;
  int parts = (p.D * 5 >= rounds * kPostParts * 4) ? kPostParts : (p.D + rounds - 1) / rounds;
  if (const char* ev = getenv("CMBPO_POST_PARTS_RT")) { const int c = atoi(ev); if (c >= 1 && c <= kPostMaxParts) parts = c; }
  const size_t smem = ((size_t)E * 2 * p.D * kPostRows + (size_t)kPostRows * p.obs_dim + 32) * sizeof(float);
  static bool attr_set = false;
  static int num_sms = 0;
  if (!attr_set) {   // obs_dim <= 64, D <= 65: at most E * 130 head lines + the obs tile
    cudaError_t e = cudaFuncSetAttribute(step_post_kernel<E, OPT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(((size_t)E * 2 * 65 * kPostRows + kPostRows * 64 + 32) * sizeof(float)));
    if (e != cudaSuccess) return e;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    attr_set = true;
  }
  // persistent grid: one wave of resident blocks (each block prefetches its next row block while it finishes the current)
  int per_sm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, step_post_kernel<E, OPT>, parts * kPostRows, smem);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) return cudaErrorInvalidConfiguration;
  const int cap = per_sm * (num_sms > 0 ? num_sms : 148);
  step_post_kernel<E, OPT><<<n_blocks < cap ? n_blocks : cap, parts * kPostRows, smem, stream>>>(p);
  return cudaGetLastError();
}

template <int E>
static cudaError_t launch_step_post_E(const StepParams& p, int n_blocks, cudaStream_t stream) {
  if (p.bound_logvar || p.sample_mode) return launch_step_post_EO<E, true>(p, n_blocks, stream);
  return launch_step_post_EO<E, false>(p, n_blocks, stream);
}

int step_post_blocks(long long n_rows, int num_sms) {
  long long tiles = (n_rows + kPostRows - 1) / kPostRows;
  long long cap = (long long)num_sms * 8;
  if (tiles < 1) tiles = 1;
  return (int)(tiles < cap ? tiles : cap);
}

cudaError_t launch_step_post(const StepParams& p, int E, int n_blocks, cudaStream_t stream) {
  switch (E) {
    case 1: return launch_step_post_E<1>(p, n_blocks, stream);
    case 2: return launch_step_post_E<2>(p, n_blocks, stream);
    case 3: return launch_step_post_E<3>(p, n_blocks, stream);
    case 4: return launch_step_post_E<4>(p, n_blocks, stream);
    case 5: return launch_step_post_E<5>(p, n_blocks, stream);
    case 6: return launch_step_post_E<6>(p, n_blocks, stream);
    case 7: return launch_step_post_E<7>(p, n_blocks, stream);
    case 8: return launch_step_post_E<8>(p, n_blocks, stream);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t launch_reduce_scalars(const float* partials, int n_blocks, const int* n_rows_dev, long long n_rows,
                                  int E, int D, int n_slots_out, float* scalars, cudaStream_t stream) {
  reduce_scalars_kernel<<<1, 256, 0, stream>>>(partials, n_blocks, n_rows_dev, n_rows, E, D, n_slots_out, scalars);
  return cudaGetLastError();
}

cudaError_t launch_aggregate_predict(const float* heads, long long row_stride, long long n_rows, int E, int D,
                                     int probabilistic, float* mean, float* var, int num_sms, cudaStream_t stream) {
  const long long total = n_rows * D;
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)num_sms * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  aggregate_predict_kernel<<<(int)blocks, 256, 0, stream>>>(heads, row_stride, n_rows, E, D, probabilistic, mean, var);
  return cudaGetLastError();
}

cudaError_t launch_finalize_predict(const float* heads, long long row_stride, long long n_rows, int E, int D,
                                    int probabilistic, float* mean, float* var, int num_sms, cudaStream_t stream) {
  const long long total = (long long)E * n_rows * D;
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)num_sms * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  finalize_predict_kernel<<<(int)blocks, 256, 0, stream>>>(heads, row_stride, n_rows, E, D, probabilistic, mean, var);
  return cudaGetLastError();
}

}  // namespace cmbpo
