// step_post.cuh -- parameters of the per-row post-processing kernels (step_post.cu)
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace cmbpo {

// static_fns registry, mbpo/static/__init__.py:7-24 (the files the hot path supports)
enum EnvId : int {
  ENV_NONE = 0, ENV_ANT = 1, ENV_ANTSAFE = 2, ENV_HALFCHEETAH = 3, ENV_HALFCHEETAHSAFE = 4,
  ENV_HUMANOID = 5, ENV_HUMANOIDSAFE = 6, ENV_HOPPER = 7, ENV_HOPPERSAFE = 8, ENV_WALKER2D = 9
};

struct StepParams {
  const float* heads;        // tile-major [block][E][2D][128] scaled mean | shifted logvar (ensemble_fwd.cu)
  long long row_stride;      // rows the scratch was sized for (multiple of 128)
  int D;                     // obs_dim + 1 (fake_env.py:58)
  int obs_dim;
  int per_slot;              // 1 = 3-D inputs: every member keeps its own trajectory (fake_env.py:164-165)
  const float* obs;          // [B][obs] or [E][B][obs]
  const int* elite_idx;      // [B] member index per row (bnn.py:437-439) or nullptr => Philox
  const int* elites;         // elite list (bnn.py:428-431), used when elite_idx == nullptr
  int n_elites;
  unsigned long long seed, rng_offset;
  const unsigned long long* rng_offset_dev;   // optional device word added to rng_offset (graph replays advance it)
  long long row_offset;      // global index of local row 0 (sharding must not change the draw)
  const int* row_ids;        // optional [B] path id per row (replaces g in row_offset + g)
  const float* noise;        // optional standard normals [B][D] / [E][B][D]; nullptr = reference behaviour
  int sample_mode;           // 1: in-kernel Philox Gaussian sampling mean + N(0,1)*std (0 = reference: means)
  int bound_logvar;          // 1: soft-bound the raw log-variance head to [min_logvar, max_logvar]
  float max_logvar, min_logvar;
  const float* head_shift;   // [2D] output-scaler shifts (2*log s_out for the logvar columns), for bound_logvar
  int env_id;                // EnvId
  int cost_mode;             // 0 zeros (fake_env.py:202), 1 static cost_f (fake_env.py:189)
  const int* n_rows_dev;     // optional device-resident row count
  long long n_rows;
  float* next_obs;           // [B][obs] or [E][B][obs]
  float* rew;                // [B] or [E][B]
  uint8_t* term;             // [B] or [E][B]
  float* cost;               // [B] or [E][B] (may be nullptr)
  float* dkl_path;           // [B]   ensemble_dkl_path (fake_env.py:158)
  float* ep_var_mean;        // [B]   mean_d(ensemble_ep_var) (model_sampler.py:428) (may be nullptr)
  float* ep_var;             // [B][D] ensemble_ep_var (fake_env.py:160) (may be nullptr)
  float* partials;           // [n_blocks][4] partial sums: dkl, var, rew, cost
  unsigned int* ticket;      // zero-initialised block counter: the last block to finish reduces the partials
  float* scalars;            // [4] ensemble_dkl_mean, ensemble_mean_var, rew_mean, cost_mean (fake_env.py:213-220)
};

int step_post_blocks(long long n_rows, int num_sms);
cudaError_t launch_step_post(const StepParams& p, int E, int n_blocks, cudaStream_t stream);
cudaError_t launch_reduce_scalars(const float* partials, int n_blocks, const int* n_rows_dev, long long n_rows,
                                  int E, int D, int n_slots_out, float* scalars, cudaStream_t stream);
cudaError_t launch_aggregate_predict(const float* heads, long long row_stride, long long n_rows, int E, int D,
                                     int probabilistic, float* mean, float* var, int num_sms, cudaStream_t stream);
cudaError_t launch_finalize_predict(const float* heads, long long row_stride, long long n_rows, int E, int D,
                                    int probabilistic, float* mean, float* var, int num_sms, cudaStream_t stream);

}  // namespace cmbpo
