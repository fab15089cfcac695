// gae_ops.cu -- ModelBuffer.finish_path_multiple on the device: generalised advantage
// estimation for the 'schedule' / 'uncertainty' pools and the inverse-variance weighted GAE + horizon trim of
// the 'iv_gae' pools (softlearning/replay_pools/modelbuffer.py:176-346; softlearning/policies/safe_utils/
// utils.py:18-188).
//
// One thread per path; every array is in the rollout's time-major storage ([E][T][n_paths], see
// include/cmbpo_b200.h), so a warp's loads of one column are one coalesced 128-byte access.  O(L) per path
// for plain GAE, O(E L^2) for the iv variant (L <= 64): HBM / latency bound, fp64 like the reference's
// NumPy (float64 lfilter / np.var on float32 buffers), results rounded to the float32 buffers at the same
// points as the reference.
#include <cstdint>
#include <cstdio>
#include <mutex>
#include <cuda_runtime.h>
#include "../../include/cmbpo_b200.h"

namespace cmbpo {

constexpr int kGaeMaxT = 64;
constexpr double kGaeEps = 1e-8;   // EPS, safe_utils/utils.py:5

// float32 delta of the reference: rews[:-1] + gamma * vals[1:] - vals[:-1] on float32 arrays (modelbuffer.py:210,:264)
__device__ __forceinline__ float delta_f32(float x, float g, float v_next, float v) {
  return __fsub_rn(__fadd_rn(x, __fmul_rn(g, v_next)), v);
}

// plain GAE of one signal: adv[t] = delta[t] + gamma*lam*adv[t+1] (discount_cumsum, utils.py:44-47)
__device__ void gae_path(const float* x, const float* v, float last, int L, size_t B, size_t b, double gamma, double lam,
                         float* adv, float* ret) {
  double run = 0.0;
  float v_next = last;
  const float g = (float)gamma;
  for (int t = L - 1; t >= 0; --t) {
    const float xt = x[(size_t)t * B + b], vt = v[(size_t)t * B + b];
    run = (double)delta_f32(xt, g, v_next, vt) + gamma * lam * run;
    const float a = (float)run;
    adv[(size_t)t * B + b] = a;
    ret[(size_t)t * B + b] = __fadd_rn(a, vt);          // R_t = A_t + V_t (modelbuffer.py:268)
    v_next = vt;
  }
}

// inverse-variance weighted GAE of one signal (modelbuffer.py:204-260 / :272-326)
template <int E>
__device__ void iv_gae_path(const float* x, const float* v, const float* last, int L, int T, size_t B, size_t b, int m,
                            double gamma_mat, double gamma, double lam, float* adv, float* ret, float* ret_var, float* roll) {
  const size_t es = (size_t)T * B;   // member stride
  double w[kGaeMaxT], var[kGaeMaxT];
  const double lg = lam * gamma * gamma;
  const float g32 = (float)gamma;
  for (int t = 0; t < L; ++t) {
    double s[E];
#pragma unroll
    for (int e = 0; e < E; ++e) s[e] = 0.0;
    double gm = 1.0, gh1 = gamma, lh = 1.0;   // gamma_mat^h, gamma^(h+1), (lam*gamma^2)^h
    const int H = L - t;
    for (int h = 0; h < H; ++h) {
      double r[E], mean = 0.0;
#pragma unroll
      for (int e = 0; e < E; ++e) {
        s[e] += gm * (double)x[e * es + (size_t)(t + h) * B + b];                       // disc_cumsum_matrix
        const double vn = (t + h + 1 < L) ? (double)v[e * es + (size_t)(t + h + 1) * B + b] : (double)last[e * B + b];
        r[e] = s[e] + vn * gh1;                                                        // + V_{t+h+1} * gamma^(h+1)
        mean += r[e];
      }
      mean /= E;
      double q = 0.0;
#pragma unroll
      for (int e = 0; e < E; ++e) { const double d = r[e] - mean; q += d * d; }
      var[h] = q / E;                                                                  // np.var(axis=0)
      w[h] = 1.0 / (var[h] + kGaeEps) * lh;
      gm *= gamma_mat; gh1 *= gamma; lh *= lg;
    }
    w[H - 1] *= 1.0 / (1.0 - lg + kGaeEps);                                            // last h of the row
    double tot = 0.0;
    for (int h = 0; h < H; ++h) tot += w[h];
    const double norm = 1.0 / tot;
    double a = 0.0, rv = 0.0, rl = 0.0, suffix = tot, gh = 1.0;
    for (int h = 0; h < H; ++h) {
      const int k = t + h;
      const float xk = x[m * es + (size_t)k * B + b], vk = v[m * es + (size_t)k * B + b];
      const float vn = (k + 1 < L) ? v[m * es + (size_t)(k + 1) * B + b] : last[m * B + b];
      a += gh * (double)delta_f32(xk, g32, vn, vk) * (suffix * norm);                  // d_weight_mat[t][h]
      rv += w[h] * norm * var[h];
      rl += (double)k * w[h];
      suffix -= w[h];
      gh *= gamma;
    }
    const float af = (float)a;
    adv[(size_t)t * B + b] = af;
    ret[(size_t)t * B + b] = __fadd_rn(af, v[m * es + (size_t)t * B + b]);
    ret_var[(size_t)t * B + b] = (float)rv;
    roll[(size_t)t * B + b] = (float)(rl * norm - (double)t);
  }
}

template <int E>
__global__ void __launch_bounds__(128) k_finish_paths(cmbpo_gae_io io) {
  const size_t B = (size_t)io.n_paths;
  const size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int T = io.max_path_length;
  int L = io.finish_ptr_dev[b];
  if (L <= 0) return;                 // modelbuffer.py:198: nothing stored, nothing computed
  if (L > T) L = T;
  if (io.mode != 2) {
    gae_path(io.rew_dev, io.val_dev, io.last_val_dev[b], L, B, b, io.gamma, io.lam, io.adv_dev, io.ret_dev);
    gae_path(io.cost_dev, io.cval_dev, io.last_cval_dev[b], L, B, b, io.cost_gamma, io.cost_lam, io.cadv_dev, io.cret_dev);
    return;
  }
  const int m = io.model_ind;
  iv_gae_path<E>(io.rew_dev, io.val_dev, io.last_val_dev, L, T, B, b, m, io.gamma, io.gamma, io.lam,
                 io.adv_dev, io.ret_dev, io.ret_var_dev, io.roll_lengths_dev);
  // the cost partial sums are discounted with gamma, the bootstrap with cost_gamma (modelbuffer.py:286-287)
  iv_gae_path<E>(io.cost_dev, io.cval_dev, io.last_cval_dev, L, T, B, b, m, io.gamma, io.cost_gamma, io.cost_lam,
                 io.cadv_dev, io.cret_dev, io.cret_var_dev, io.croll_lengths_dev);
  if (io.populated_dev) {
    // horizon = first column whose normalised return variance exceeds a limit, evaluated on the float32 buffers;
    // argmax of an all-False mask is 0, which drops the whole path (modelbuffer.py:329-343)
    const float r0 = __fadd_rn(io.ret_var_dev[b], 1e-8f), c0 = __fadd_rn(io.cret_var_dev[b], 1e-8f);
    int horizon = 0;
    for (int t = 0; t < L; ++t) {
      const bool ur = __fdiv_rn(io.ret_var_dev[(size_t)t * B + b], r0) > io.max_uncertainty_r;
      const bool uc = io.cares_about_cost && (__fdiv_rn(io.cret_var_dev[(size_t)t * B + b], c0) > io.max_uncertainty_c);
      if (ur || uc) { horizon = t; break; }
    }
    for (int t = horizon; t < T; ++t) io.populated_dev[(size_t)t * B + b] = 0;
  }
}

}  // namespace cmbpo

using namespace cmbpo;
static thread_local char g_gae_err[160];

extern "C" int cmbpo_finish_paths(const cmbpo_gae_io* io_, void* stream_) {
  if (!io_) return CMBPO_ERR_ARG;
  const cmbpo_gae_io io = *io_;
  if (io.n_paths < 0 || io.max_path_length <= 0) return CMBPO_ERR_ARG;
  if (!io.rew_dev || !io.val_dev || !io.cost_dev || !io.cval_dev || !io.finish_ptr_dev || !io.last_val_dev ||
      !io.last_cval_dev || !io.adv_dev || !io.ret_dev || !io.cadv_dev || !io.cret_dev) return CMBPO_ERR_ARG;
  if (io.mode == 2) {
    if (!io.ret_var_dev || !io.cret_var_dev || !io.roll_lengths_dev || !io.croll_lengths_dev) return CMBPO_ERR_ARG;
    if (io.max_path_length > kGaeMaxT || io.n_members < 1 || io.n_members > 8 || io.model_ind < 0 ||
        io.model_ind >= io.n_members) return CMBPO_ERR_UNSUPPORTED;
  }
  if (io.n_paths == 0) return CMBPO_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  const int nb = (int)((io.n_paths + 127) / 128);
  switch (io.mode == 2 ? io.n_members : 1) {
    case 1: k_finish_paths<1><<<nb, 128, 0, stream>>>(io); break;
    case 2: k_finish_paths<2><<<nb, 128, 0, stream>>>(io); break;
    case 3: k_finish_paths<3><<<nb, 128, 0, stream>>>(io); break;
    case 4: k_finish_paths<4><<<nb, 128, 0, stream>>>(io); break;
    case 5: k_finish_paths<5><<<nb, 128, 0, stream>>>(io); break;
    case 6: k_finish_paths<6><<<nb, 128, 0, stream>>>(io); break;
    case 7: k_finish_paths<7><<<nb, 128, 0, stream>>>(io); break;
    default: k_finish_paths<8><<<nb, 128, 0, stream>>>(io); break;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { snprintf(g_gae_err, sizeof(g_gae_err), "finish_paths: %s", cudaGetErrorString(e)); return CMBPO_ERR_CUDA; }
  return CMBPO_OK;
}

// ---------------------------------------------------------------------------------------
// ModelBuffer.get's advantage normalisation (modelbuffer.py:358-372): over the populated entries,
//   adv  <- (adv - mean(adv)) / (std(adv) + EPS)      (mpi_statistics_scalar: float32 mean and population std)
//   cadv <- cadv - mean(cadv)                         (centred, NOT rescaled)
// plus the means of ret / cret for the diagnostics.  Three small launches (sums -> squared deviations -> apply);
// partial sums in fp64, reduced by the last block to finish; the statistics are rounded to float32 where the
// reference holds them in float32.  stats_dev: [adv_mean, adv_std, cadv_mean, ret_mean, cret_mean, n].
// ---------------------------------------------------------------------------------------
namespace cmbpo {

constexpr int kNormThreads = 256;

__device__ __forceinline__ double block_sum(double x, double* sh) {
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = x;
  __syncthreads();
  double t = 0.0;
  for (int w = 0; w < kNormThreads / 32; ++w) t += sh[w];
  __syncthreads();
  return t;
}

template <int PHASE>
__global__ void __launch_bounds__(kNormThreads) k_adv_stats(cmbpo_norm_io io, double* partials, unsigned int* ticket) {
  __shared__ double sh[kNormThreads / 32];
  __shared__ bool s_last;
  const long long n = io.n_elems;
  constexpr int NV = PHASE == 1 ? 5 : 1;
  double acc[NV];
  for (int k = 0; k < NV; ++k) acc[k] = 0.0;
  const float mean = PHASE == 2 ? io.stats_dev[0] : 0.f;
  for (long long i = (long long)blockIdx.x * kNormThreads + threadIdx.x; i < n; i += (long long)gridDim.x * kNormThreads) {
    if (!io.populated_dev[i]) continue;
    if (PHASE == 1) {
      acc[0] += (double)io.adv_dev[i]; acc[1] += (double)io.cadv_dev[i];
      acc[2] += io.ret_dev ? (double)io.ret_dev[i] : 0.0; acc[3] += io.cret_dev ? (double)io.cret_dev[i] : 0.0;
      acc[4] += 1.0;
    } else {
      const float d = __fsub_rn(io.adv_dev[i], mean);     // (x - mean)**2 on float32 (mpi_tools.py:85)
      acc[0] += (double)__fmul_rn(d, d);
    }
  }
  for (int k = 0; k < NV; ++k) {
    const double t = block_sum(acc[k], sh);
    if (threadIdx.x == 0) partials[(size_t)blockIdx.x * NV + k] = t;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double tot[NV];
  for (int k = 0; k < NV; ++k) {
    double x = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += kNormThreads) x += __ldcg(partials + (size_t)b * NV + k);
    tot[k] = block_sum(x, sh);
  }
  if (threadIdx.x == 0) {
    if (PHASE == 1) {
      const double cnt = tot[4];
      io.stats_dev[0] = cnt > 0 ? (float)((float)tot[0] / (float)cnt) : 0.f;   // float32 sum / float32 n (mpi_tools.py:82-83)
      io.stats_dev[2] = cnt > 0 ? (float)((float)tot[1] / (float)cnt) : 0.f;
      io.stats_dev[3] = cnt > 0 ? (float)(tot[2] / cnt) : 0.f;
      io.stats_dev[4] = cnt > 0 ? (float)(tot[3] / cnt) : 0.f;
      io.stats_dev[5] = (float)cnt;
    } else {
      const float cnt = io.stats_dev[5];
      io.stats_dev[1] = cnt > 0 ? sqrtf((float)tot[0] / cnt) : 0.f;             // :86
    }
    *ticket = 0;
  }
}

__global__ void __launch_bounds__(kNormThreads) k_adv_apply(cmbpo_norm_io io) {
  const float mean = io.stats_dev[0], denom = __fadd_rn(io.stats_dev[1], 1e-8f), cmean = io.stats_dev[2];
  if (io.stats_dev[5] <= 0.f) return;                                           // modelbuffer.py:356 `if self.size>0`
  for (long long i = (long long)blockIdx.x * kNormThreads + threadIdx.x; i < io.n_elems; i += (long long)gridDim.x * kNormThreads) {
    if (!io.populated_dev[i]) continue;
    io.adv_dev[i] = __fdiv_rn(__fsub_rn(io.adv_dev[i], mean), denom);           // :361
    io.cadv_dev[i] = __fsub_rn(io.cadv_dev[i], cmean);                          // :367
  }
}
}  // namespace cmbpo

extern "C" size_t cmbpo_normalize_scratch_bytes(void) { return 1024 * 5 * sizeof(double) + 64; }

extern "C" int cmbpo_normalize_advantages(const cmbpo_norm_io* io_, void* stream_) {
  if (!io_ || !io_->adv_dev || !io_->cadv_dev || !io_->populated_dev || !io_->stats_dev || !io_->scratch_dev) return CMBPO_ERR_ARG;
  const cmbpo_norm_io io = *io_;
  if (io.n_elems < 0) return CMBPO_ERR_ARG;
  if (io.scratch_bytes < cmbpo_normalize_scratch_bytes()) return CMBPO_ERR_WORKSPACE;
  cudaStream_t stream = (cudaStream_t)stream_;
  long long nb = (io.n_elems + kNormThreads - 1) / kNormThreads;
  if (nb > 1024) nb = 1024;
  if (nb < 1) nb = 1;
  unsigned int* ticket = static_cast<unsigned int*>(io.scratch_dev);            // must be zero (first call: memset below)
  double* partials = reinterpret_cast<double*>(static_cast<uint8_t*>(io.scratch_dev) + 64);
  cudaMemsetAsync(ticket, 0, 64, stream);
  k_adv_stats<1><<<(int)nb, kNormThreads, 0, stream>>>(io, partials, ticket);
  k_adv_stats<2><<<(int)nb, kNormThreads, 0, stream>>>(io, partials, ticket);
  k_adv_apply<<<(int)nb, kNormThreads, 0, stream>>>(io);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { snprintf(g_gae_err, sizeof(g_gae_err), "normalize_advantages: %s", cudaGetErrorString(e)); return CMBPO_ERR_CUDA; }
  return CMBPO_OK;
}

// ---------------------------------------------------------------------------------------
// cmbpo_normalize_advantages_static: the same three launches as cmbpo_normalize_advantages, with the scratch (ticket word +
// per-block partial sums) in a file-scope device array instead of caller memory.  No counterpart in the reference; added
// at the request of the run's operators for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).  The entry point allocates nothing: no cudaMalloc / cudaFree / allocating library call.
//  * every GPU that loads the library holds the kNormStaticBytes of the array until exit;
//  * the array is shared by every call and keeps the last call's contents: the ticket word is cleared at the start of each
//    call (the partial sums are fully rewritten before they are read), and two calls that overlap on the GPU -- different
//    streams, or different host threads -- race on it.  Serialise them, or use cmbpo_normalize_advantages with own scratch;
//  * in host code `g_norm_static` is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress,
//    looked up once per device with that device current.
// ---------------------------------------------------------------------------------------
namespace cmbpo {
constexpr size_t kNormStaticBytes = CMBPO_NORMALIZE_STATIC_SCRATCH_BYTES;
static_assert(kNormStaticBytes % 256 == 0 && kNormStaticBytes >= 1024 * 5 * sizeof(double) + 64, "static scratch too small");
static __device__ __align__(256) unsigned char g_norm_static[kNormStaticBytes];
constexpr int kNormStaticMaxDevices = 64;
static void* g_norm_static_ptr[kNormStaticMaxDevices];      // device address of g_norm_static, per device ordinal
static std::mutex g_norm_static_mu;
}  // namespace cmbpo

extern "C" int cmbpo_normalize_advantages_static(const cmbpo_norm_io* io_, void* stream_, char* err, size_t err_len) {
  auto fail = [&](int code, const char* fmt, auto... args) {
    if (err && err_len) snprintf(err, err_len, fmt, args...);
    return code;
  };
  if (err && err_len) err[0] = 0;
  if (!io_ || !io_->adv_dev || !io_->cadv_dev || !io_->populated_dev || !io_->stats_dev)
    return fail(CMBPO_ERR_ARG, "normalize_advantages_static: adv_dev, cadv_dev, populated_dev and stats_dev must be given");
  const cmbpo_norm_io io = *io_;      // scratch_dev / scratch_bytes are ignored
  if (io.n_elems < 0) return fail(CMBPO_ERR_ARG, "normalize_advantages_static: negative n_elems");
  if (io.n_elems > (int64_t)CMBPO_NORMALIZE_STATIC_MAX_ELEMS)
    return fail(CMBPO_ERR_UNSUPPORTED, "normalize_advantages_static: n_elems %lld exceeds the maximum of %lld entries; use "
                "cmbpo_normalize_advantages with caller scratch", (long long)io.n_elems, (long long)CMBPO_NORMALIZE_STATIC_MAX_ELEMS);
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(CMBPO_ERR_CUDA, "normalize_advantages_static: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kNormStaticMaxDevices)
    return fail(CMBPO_ERR_UNSUPPORTED, "normalize_advantages_static: device ordinal %d beyond %d", dev, kNormStaticMaxDevices);
  void* base = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_norm_static_mu);
    if (!g_norm_static_ptr[dev]) {      // set-up, once per device, with that device current
      void* ptr = nullptr;
      e = cudaGetSymbolAddress(&ptr, g_norm_static);
      if (e != cudaSuccess || !ptr)
        return fail(CMBPO_ERR_CUDA, "normalize_advantages_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      g_norm_static_ptr[dev] = ptr;
    }
    base = g_norm_static_ptr[dev];
  }
  cudaStream_t stream = (cudaStream_t)stream_;
  long long nb = (io.n_elems + kNormThreads - 1) / kNormThreads;
  if (nb > 1024) nb = 1024;
  if (nb < 1) nb = 1;
  unsigned int* ticket = static_cast<unsigned int*>(base);
  double* partials = reinterpret_cast<double*>(static_cast<uint8_t*>(base) + 64);
  e = cudaMemsetAsync(ticket, 0, 64, stream);       // the array keeps the last call's contents: the ticket must start at zero
  if (e != cudaSuccess) return fail(CMBPO_ERR_CUDA, "normalize_advantages_static: cudaMemsetAsync: %s", cudaGetErrorString(e));
  k_adv_stats<1><<<(int)nb, kNormThreads, 0, stream>>>(io, partials, ticket);
  k_adv_stats<2><<<(int)nb, kNormThreads, 0, stream>>>(io, partials, ticket);
  k_adv_apply<<<(int)nb, kNormThreads, 0, stream>>>(io);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail(CMBPO_ERR_CUDA, "normalize_advantages_static: %s", cudaGetErrorString(e));
  return CMBPO_OK;
}
