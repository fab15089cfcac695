// actor_ops.cu -- Gaussian policy head of the CPO actor inside the device rollout loop:
// softlearning/models/ac_network.py:105-117 (mlp_gaussian_policy: pi = mu + N(0,1) * exp(log_std), logp_pi =
// gaussian_likelihood(pi, mu, log_std), :50-52), as CPOPolicy.get_action_outs evaluates it every sampler step
// (softlearning/policies/cpo_policy.py:883-925, called at softlearning/samplers/model_sampler.py:275-294).  The actor MLP
// itself (tanh, e.g. 256 x 256) runs on the ensemble forward kernels as a 1-member non-probabilistic model.
#include <cstdint>
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

// One thread per row: the act_dim normals of the row are drawn once and applied to every slot of the row (the reference
// tiles one draw over the ensemble axis, ac_network.py:135-138); the log-probability is summed over the action columns
// in column order (deterministic).  The arrays are a few tens of bytes per row: the kernel is latency-, not bandwidth-bound.
__global__ void __launch_bounds__(256) k_actor_head(cmbpo_actor_io io) {
  const long long n = io.n_rows_dev ? (long long)(*io.n_rows_dev) : io.n_rows;
  const int A = io.act_dim, S = io.n_slots > 1 ? io.n_slots : 1;
  const unsigned long long rng = io.rng_offset + (io.rng_offset_dev ? *io.rng_offset_dev : 0ull);
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const unsigned long long ctr = (unsigned long long)(io.row_offset + (io.row_id_dev ? (long long)io.row_id_dev[r] : r));
    for (int s = 0; s < S; ++s) {
      const size_t o = ((size_t)s * io.n_rows + (size_t)r) * A;
      float lp = 0.f;
      for (int a = 0; a < A; ++a) {
        float z = 0.f;
        if (!io.deterministic) {
          if (io.noise_dev) {
            z = io.noise_dev[(size_t)r * A + a];
          } else {   // stream id 0x41 ("A") in counter word 3 keeps these draws apart from the step kernel's
            uint32_t rnd[4];
            philox4x32_10_a((uint32_t)ctr, (uint32_t)(ctr >> 32), (uint32_t)rng, 0x41000000u ^ (uint32_t)a,
                            (uint32_t)io.seed, (uint32_t)(io.seed >> 32), rnd);
            const float u1 = ((float)(rnd[0] >> 8) + 1.0f) * (1.0f / 16777216.0f);
            const float u2 = ((float)(rnd[1] >> 8) + 1.0f) * (1.0f / 16777216.0f);
            z = sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
          }
        }
        const float ls = io.log_std_dev[a], std = expf(ls);
        const float mu = io.mu_dev[o + a];
        const float pi = io.deterministic ? mu : fmaf(z, std, mu);
        io.pi_dev[o + a] = pi;
        const float q = (pi - mu) / (std + 1e-8f);                                  // ac_network.py:51, EPS = 1e-8
        lp += -0.5f * (q * q + 2.0f * ls + 1.8378770664093453f);                    // log(2 pi)
      }
      if (io.logp_pi_dev) io.logp_pi_dev[(size_t)s * io.n_rows + (size_t)r] = lp;
    }
  }
}

// np.random.permutation(num_networks) of FakeEnv.forward_shuffle (mbpo/models/fake_env.py:309-316) drawn on the device:
// one thread, Fisher-Yates from the front; stream id 0x50 ("P") in counter word 3, counter word 0 = step.
__global__ void k_draw_permutation(int n, unsigned long long seed, unsigned long long rng_offset,
                                   const unsigned long long* rng_offset_dev, unsigned long long step, int* perm) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const unsigned long long rng = rng_offset + (rng_offset_dev ? *rng_offset_dev : 0ull);
  int v[CMBPO_MAX_MEMBERS];
  for (int i = 0; i < n; ++i) v[i] = i;
  for (int i = 0; i + 1 < n; ++i) {
    uint32_t rnd[4];
    philox4x32_10_a((uint32_t)step, (uint32_t)(step >> 32), (uint32_t)rng, 0x50000000u ^ (uint32_t)i,
                    (uint32_t)seed, (uint32_t)(seed >> 32), rnd);
    const int j = i + (int)(((unsigned long long)rnd[0] * (unsigned long long)(n - i)) >> 32);
    const int t = v[i]; v[i] = v[j]; v[j] = t;
  }
  for (int i = 0; i < n; ++i) perm[i] = v[i];
}

}  // namespace cmbpo

extern "C" int cmbpo_draw_permutation(int32_t n, uint64_t seed, uint64_t rng_offset, const uint64_t* rng_offset_dev,
                                      uint64_t step, int32_t* perm_dev, void* stream_) {
  if (n < 1 || n > CMBPO_MAX_MEMBERS || !perm_dev) return CMBPO_ERR_ARG;
  cmbpo::k_draw_permutation<<<1, 32, 0, (cudaStream_t)stream_>>>(n, seed, rng_offset,
      reinterpret_cast<const unsigned long long*>(rng_offset_dev), step, perm_dev);
  return cudaGetLastError() == cudaSuccess ? CMBPO_OK : CMBPO_ERR_CUDA;
}

extern "C" int cmbpo_actor_head(const cmbpo_actor_io* io_, void* stream_) {
  if (!io_ || !io_->mu_dev || !io_->log_std_dev || !io_->pi_dev) return CMBPO_ERR_ARG;
  const cmbpo_actor_io io = *io_;
  if (io.n_rows < 0 || io.act_dim < 1) return CMBPO_ERR_ARG;
  if (io.n_rows == 0) return CMBPO_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  long long blocks = (io.n_rows + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  cmbpo::k_actor_head<<<(int)blocks, 256, 0, stream>>>(io);
  return cudaGetLastError() == cudaSuccess ? CMBPO_OK : CMBPO_ERR_CUDA;
}
