// archive_ops.cu -- start-state sampling for the model rollouts: the Boltzmann-over-epochs
// distribution of CPOBuffer.boltz_dist and the weighted draw + gather of
// CPOBuffer.distributed_batch_from_archive (softlearning/replay_pools/cpobuffer.py:531-542, :559-610;
// mbpo/algorithms/cmbpo.py:328-333) on an archive that lives in HBM, so the start states of a rollout never
// touch the host.
//
// np.random.choice(n, size, p) is: cdf = cumsum(float64(p)); cdf /= cdf[-1]; u = random_sample(size);
// idx = searchsorted(cdf, u, side='right').  Here: one block builds the float64 cdf (per-thread sequential
// chunks + block scan), then one thread per sample draws u (injected, or 53-bit Philox4x32-10 doubles keyed by
// the global sample id), binary-searches the cdf and gathers its row of every requested field.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include "../../include/cmbpo_b200.h"

namespace cmbpo {

__device__ __forceinline__ void philox4x32_10_a(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                                                uint32_t (&out)[4]) {
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

constexpr int kCdfThreads = 1024;

// p[i] = epoch[i] >= 0 ? sample_p[epoch[i]] : 0 (cpobuffer.py:540-541), cdf = cumsum(p) / sum(p)
__global__ void __launch_bounds__(kCdfThreads) k_build_cdf(cmbpo_sample_io io) {
  __shared__ double s_tot[kCdfThreads];
  const long long n = io.archive_size;
  const long long chunk = (n + kCdfThreads - 1) / kCdfThreads;
  const long long lo = (long long)threadIdx.x * chunk, hi = min(n, lo + chunk);
  auto prob = [&](long long i) -> double {
    if (io.prob_dev) return (double)io.prob_dev[i];
    const int e = io.epoch_dev[i];
    return (e >= 0 && e < io.n_epochs) ? (double)io.epoch_prob_dev[e] : 0.0;
  };
  double run = 0.0;
  for (long long i = lo; i < hi; ++i) { run += prob(i); io.cdf_dev[i] = run; }
  s_tot[threadIdx.x] = run;
  __syncthreads();
  for (int off = 1; off < kCdfThreads; off <<= 1) {   // inclusive scan of the chunk totals
    const double t = (threadIdx.x >= off) ? s_tot[threadIdx.x - off] : 0.0;
    __syncthreads();
    s_tot[threadIdx.x] += t;
    __syncthreads();
  }
  const double base = threadIdx.x ? s_tot[threadIdx.x - 1] : 0.0;
  const double total = s_tot[kCdfThreads - 1];
  for (long long i = lo; i < hi; ++i) io.cdf_dev[i] = (io.cdf_dev[i] + base) / total;
}

__global__ void __launch_bounds__(256) k_choice_gather(cmbpo_sample_io io) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= io.batch_size) return;
  double u;
  if (io.uniform_dev) {
    u = io.uniform_dev[b];
  } else {   // 53-bit double like NumPy's random_sample: (a >> 5) * 2^26 + (b >> 6), over 2^53
    uint32_t r[4];
    const unsigned long long ctr = (unsigned long long)(io.sample_offset + b);
    philox4x32_10_a((uint32_t)ctr, (uint32_t)(ctr >> 32), (uint32_t)io.rng_offset, 0x51ED270Bu,
                    (uint32_t)io.seed, (uint32_t)(io.seed >> 32), r);
    u = ((double)(r[0] >> 5) * 67108864.0 + (double)(r[1] >> 6)) * (1.0 / 9007199254740992.0);
  }
  long long lo = 0, hi = io.archive_size;          // searchsorted(cdf, u, side='right'): first i with cdf[i] > u
  while (lo < hi) {
    const long long mid = (lo + hi) >> 1;
    if (io.cdf_dev[mid] <= u) lo = mid + 1; else hi = mid;
  }
  if (lo >= io.archive_size) lo = io.archive_size - 1;
  if (io.index_dev) io.index_dev[b] = lo;
  for (int f = 0; f < CMBPO_MAX_FIELDS; ++f) {
    const float* src = io.field_dev[f];
    if (!src) continue;
    const int w = io.field_width[f];
    float* dst = io.out_dev[f] + (size_t)b * w;
    for (int d = 0; d < w; ++d) dst[d] = src[(size_t)lo * w + d];
  }
}

}  // namespace cmbpo

using namespace cmbpo;

extern "C" int cmbpo_sample_archive(const cmbpo_sample_io* io_, void* stream_) {
  if (!io_ || !io_->cdf_dev) return CMBPO_ERR_ARG;
  const cmbpo_sample_io io = *io_;
  if (io.archive_size <= 0 || io.batch_size < 0) return CMBPO_ERR_ARG;
  if (!io.prob_dev && (!io.epoch_dev || !io.epoch_prob_dev || io.n_epochs <= 0)) return CMBPO_ERR_ARG;
  for (int f = 0; f < CMBPO_MAX_FIELDS; ++f)
    if (io.field_dev[f] && (!io.out_dev[f] || io.field_width[f] <= 0)) return CMBPO_ERR_ARG;
  cudaStream_t stream = (cudaStream_t)stream_;
  if (io.rebuild_cdf) k_build_cdf<<<1, kCdfThreads, 0, stream>>>(io);
  if (io.batch_size > 0) k_choice_gather<<<(int)((io.batch_size + 255) / 256), 256, 0, stream>>>(io);
  return cudaGetLastError() == cudaSuccess ? CMBPO_OK : CMBPO_ERR_CUDA;
}
