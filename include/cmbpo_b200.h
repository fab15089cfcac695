/* cmbpo_b200.h -- C ABI of the B200-native CMBPO model-rollout path (libcmbpo_b200.so).
 *
 * The reference (anyboby/ConstrainedMBPO) has no FFI for this path: the boundary is the Python
 * object surface construct_model / BNN.predict / FakeEnv.step.  Each entry point below cites
 * the reference interface it stands behind; the Python mirror of that surface
 * (constrainedmbpo_b200/{constructor,bnn,fake_env,rollout}.py) is a thin ctypes client of
 * this library.  Citations are relative to the reference repository root.
 *
 * Conventions: every `*_dev`/device pointer is a CUDA device pointer owned by the caller
 * (PyTorch tensors in the Python mirror); `stream` is a cudaStream_t passed as void*; no call
 * allocates, frees or synchronises except create/destroy/set_*; return value 0 = ok, <0 = error
 * (text via cmbpo_last_error()).  A handle is reentrant across handles, not thread-safe per
 * handle.  There is no CPU fallback: every call fails with CMBPO_ERR_CUDA if no sm_100 device
 * is usable.
 */
#ifndef CMBPO_B200_H_
#define CMBPO_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CMBPO_MAX_LAYERS 8
#define CMBPO_MAX_SEGMENTS 3
#define CMBPO_MAX_MEMBERS 16     /* ensemble members a handle can hold */

enum {
  CMBPO_OK = 0,
  CMBPO_ERR_ARG = -1,         /* invalid argument (mirrors ValueError("Invalid input dimension."), fc.py:111) */
  CMBPO_ERR_UNSUPPORTED = -2, /* shape outside what the resident-weight kernel supports */
  CMBPO_ERR_CUDA = -3,        /* CUDA runtime error / no sm_100 device */
  CMBPO_ERR_STATE = -4,       /* weights/scalers not set (mirrors RuntimeError, bnn.py:135-140,170-173) */
  CMBPO_ERR_WORKSPACE = -5    /* workspace too small */
};

/* FC._activations, mbpo/models/fc.py:14-21 (softmax is not supported on this path) */
enum { CMBPO_ACT_NONE = 0, CMBPO_ACT_SWISH = 1, CMBPO_ACT_RELU = 2, CMBPO_ACT_TANH = 3, CMBPO_ACT_SIGMOID = 4 };

/* mbpo/static/__init__.py:7-24 -- the static_fns modules whose termination_fn / cost_f are fused */
enum {
  CMBPO_ENV_NONE = 0, CMBPO_ENV_ANT = 1, CMBPO_ENV_ANTSAFE = 2, CMBPO_ENV_HALFCHEETAH = 3,
  CMBPO_ENV_HALFCHEETAHSAFE = 4, CMBPO_ENV_HUMANOID = 5, CMBPO_ENV_HUMANOIDSAFE = 6,
  CMBPO_ENV_HOPPER = 7, CMBPO_ENV_HOPPERSAFE = 8, CMBPO_ENV_WALKER2D = 9
};

/* Structure of one ensemble: what construct_model(...) fixes (mbpo/models/constructor.py:9-67)
 * after BNN.finalize doubled the head of a probabilistic model (mbpo/models/bnn.py:190-196). */
typedef struct cmbpo_cfg {
  int32_t num_members;                    /* num_networks                     constructor.py:13 */
  int32_t num_layers;                     /* len(hidden_dims) + 1             constructor.py:57-62 */
  int32_t in_dim;                         /* in_dim                           constructor.py:9 */
  int32_t out_dim;                        /* out_dim (per head)               constructor.py:10 */
  int32_t probabilistic;                  /* loss in {NLL, MSPE}              bnn.py:111-112 */
  int32_t layer_width[CMBPO_MAX_LAYERS];  /* FC output widths; last = 2*out_dim if probabilistic */
  int32_t layer_act[CMBPO_MAX_LAYERS];    /* CMBPO_ACT_* per layer; last must be NONE (bnn.py:194-196) */
  /* Tensor-core operand type (north_star: "tf32 or bf16 in, fp32 accumulate").  0 = bf16 (default, rtol 1e-2 vs the
   * reference's fp32 arithmetic), 1 = tf32: fp32 operand images read by tcgen05.mma kind::tf32 (10-bit mantissa, rtol
   * 1e-3), exact sigmoid / tanh in the epilogue instead of tanh.approx; half the tensor throughput, layer-at-a-time kernel. */
  int32_t operand_dtype;
} cmbpo_cfg;
enum { CMBPO_DTYPE_BF16 = 0, CMBPO_DTYPE_TF32 = 1 };

/* construct_model -> BNN(params) + FC stack + finalize   (constructor.py:41-67) */
int cmbpo_create(const cmbpo_cfg* cfg, int device, void** handle);
void cmbpo_destroy(void* handle);
const char* cmbpo_last_error(void);
/* 1 if a CUDA device with compute capability 10.x is visible, else 0 (never launches) */
int cmbpo_device_ok(int device);

/* FC weights [E][in][out] and biases [E][1][out], fp32 (mbpo/models/fc.py:154-185); read after
 * every BNN.train() (bnn.py:467-666) because training rewrites them.  Device pointers. */
int cmbpo_set_weights(void* handle, int32_t layer, const float* w_dev, const float* b_dev, void* stream);
/* TensorStandardScaler state (mbpo/models/utils.py:98-115): mu/var [dim] HOST pointers, NULL =
 * scaler absent (use_scaler_in/out False, bnn.py:201-207).  `var` is the variance (the TF
 * variable is only *named* _std); divisor = max(sqrt(var)*sc_factor, 1e-2) (utils.py:161). */
int cmbpo_set_scalers(void* handle, const float* mu_in, const float* var_in, const float* mu_out,
                      const float* var_out, float sc_factor, void* stream);
/* BNN._model_inds (bnn.py:98-102, :428-431): HOST pointer, duplicates allowed */
int cmbpo_set_elites(void* handle, const int32_t* elites, int32_t k, void* stream);

/* bytes of caller-provided device scratch needed for `max_rows` rows */
size_t cmbpo_workspace_bytes(void* handle, int64_t max_rows);

/* BNN.predict (mbpo/models/bnn.py:671-776).
 * Input rows are the concatenation of up to 3 row-major segments (fake_env.py:141-145,:194).
 * ndim 2: segments are [B][w]; ndim 3: [E][B][w], member e consumes slot e (fc.py:107-108).
 * agg 0 (factored): mean/var are [E][B][D]; agg 1: ensemble mean [B][D] and, for probabilistic
 * models, mean_E(var)+mean_E((mean-mean_E)^2) (bnn.py:350-352), else var_E (bnn.py:736-743).
 * var_dev may be NULL (inc_var=False). */
typedef struct cmbpo_predict_io {
  const float* seg_dev[CMBPO_MAX_SEGMENTS];
  int32_t seg_width[CMBPO_MAX_SEGMENTS];
  int64_t n_rows;
  int32_t ndim;
  int32_t agg;
  float* mean_dev;
  float* var_dev;
  void* workspace_dev;
  size_t workspace_bytes;
  const int32_t* n_rows_dev;   /* optional live row count on the device (rollout loop: n_rows is then the capacity; rows
                                * beyond the live count are not computed, their outputs are unspecified) */
} cmbpo_predict_io;
int cmbpo_predict(void* handle, const cmbpo_predict_io* io, void* stream);

/* Gaussian policy head (softlearning/models/ac_network.py:105-117, the actor CPOPolicy.get_action_outs runs every sampler
 * step, softlearning/policies/cpo_policy.py:883-925): given mu = the actor MLP's output (cmbpo_predict of a 1-member
 * tanh ensemble) and the state-independent log_std, pi = mu + N(0,1) * exp(log_std) and
 * logp_pi = sum_a -0.5 * (((pi - mu) / (exp(log_std) + 1e-8))^2 + 2 log_std + log(2 pi))  (gaussian_likelihood, :50-52).
 * noise_dev injects the normals ([n_rows][act]); NULL = Philox4x32-10 + Box-Muller keyed by (global row id, column,
 * rng_offset, seed).  n_slots > 1 ('iv_gae' states, [n_slots][n_rows][.] arrays): the SAME normals for every slot of a
 * row (ac_network.py:129-146 tiles one draw over the ensemble axis).  deterministic 1: pi = mu. */
typedef struct cmbpo_actor_io {
  const float* mu_dev;         /* [n_slots][n_rows][act] */
  const float* log_std_dev;    /* [act] */
  const float* noise_dev;      /* [n_rows][act] or NULL */
  int64_t n_rows;              /* rows per slot (capacity when n_rows_dev is given) */
  const int32_t* n_rows_dev;
  int32_t act_dim;
  int32_t n_slots;
  int32_t deterministic;
  const int32_t* row_id_dev;   /* optional [n_rows] global path id per row (Philox counter after compaction) */
  int64_t row_offset;
  uint64_t seed, rng_offset;
  const uint64_t* rng_offset_dev;
  float* pi_dev;               /* [n_slots][n_rows][act] */
  float* logp_pi_dev;          /* [n_slots][n_rows] */
} cmbpo_actor_io;
int cmbpo_actor_head(const cmbpo_actor_io* io, void* stream);

/* FakeEnv.step (mbpo/models/fake_env.py:131-222) for the dynamics ensemble of `handle`
 * (in_dim = obs+act, out_dim = obs+1). */
typedef struct cmbpo_step_io {
  const float* obs_dev;        /* [B][obs] or [E][B][obs] */
  const float* act_dev;        /* [B][act] or [E][B][act] */
  int64_t n_rows;              /* B (capacity when n_rows_dev is given) */
  const int32_t* n_rows_dev;   /* optional live row count on the device (rollout loop) */
  int32_t ndim;                /* 2 or 3 (fake_env.py:133-139, 1-D is broken upstream) */
  const int32_t* perm;         /* HOST [E]: forward_shuffle permutation (fake_env.py:309-316), NULL = identity */
  const int32_t* elite_idx_dev;/* [B] injected BNN.random_inds draw (bnn.py:437-439) or NULL => Philox */
  const int32_t* row_id_dev;   /* optional [B] global path id per row: Philox counter after compaction */
  const float* noise_dev;      /* optional N(0,1) [B][D] / [E][B][D]; NULL = reference behaviour */
  uint64_t seed, rng_offset;   /* Philox key / stream when elite_idx_dev == NULL */
  int64_t row_offset;          /* global index of row 0 (shard-invariant draws) */
  int32_t env_id;              /* CMBPO_ENV_* : static termination_fn (fake_env.py:182-185) */
  int32_t cost_mode;           /* 0: zeros (fake_env.py:202); 1: static cost_f (fake_env.py:189) */
  float* next_obs_dev;         /* [B][obs] or [E][B][obs] */
  float* rew_dev;              /* [B] or [E][B] */
  uint8_t* term_dev;           /* [B] or [E][B] */
  float* cost_dev;             /* [B] or [E][B], may be NULL */
  float* dkl_path_dev;         /* [B] ensemble_dkl_path (fake_env.py:158) */
  float* ep_var_mean_dev;      /* [B] mean over D of ensemble_ep_var, may be NULL */
  float* ep_var_dev;           /* [B][D] ensemble_ep_var (fake_env.py:160), may be NULL */
  float* scalars_dev;          /* [4]: ensemble_dkl_mean, ensemble_mean_var, rew_mean, cost_mean */
  void* workspace_dev;
  size_t workspace_bytes;
  /* Options beyond the live reference (both 0 = reference behaviour):
   * sample_mode 1: Gaussian sampling in the kernel, sample = mean + N(0,1)*std of the selected member (every member in
   *   3-D mode), normals from Philox4x32-10 (Box-Muller) keyed by (global row id, member, column, rng_offset, seed):
   *   independent of sharding and batch position.  noise_dev, when given, takes precedence (injected draws).
   * bound_logvar 1: logvar = max_logvar - softplus(max_logvar - logvar); logvar = min_logvar + softplus(logvar -
   *   min_logvar) on the raw head, before the inverse output scaler (the PETS bounding that is commented out in
   *   bnn.py:888-889; bnn_chua.py:400-401). */
  int32_t sample_mode;
  int32_t bound_logvar;
  float max_logvar, min_logvar;
  /* Optional device word added to rng_offset by the kernel (NULL = 0): a launch captured in a CUDA graph draws from a
   * fresh Philox stream on every replay when the graph also advances this word (ModelRollout.reset does). */
  const uint64_t* rng_offset_dev;
  /* Optional DEVICE [E] forward_shuffle permutation (3-D only); takes precedence over the host `perm`.  Filled by
   * cmbpo_draw_permutation inside the same stream / CUDA graph, it gives every replay of a captured 'iv_gae' rollout a
   * fresh permutation per step, like the reference's np.random.permutation per FakeEnv.step call (fake_env.py:314). */
  const int32_t* perm_dev;
} cmbpo_step_io;
int cmbpo_step(void* handle, const cmbpo_step_io* io, void* stream);

/* Uniform random permutation of 0..n-1 on the device (np.random.permutation(num_networks), fake_env.py:314): Fisher-Yates
 * driven by Philox4x32-10 words keyed by (seed, rng_offset + *rng_offset_dev, step); element i is swapped with
 * j = i + floor(u * (n - i)), u the 32-bit word / 2^32.  n <= CMBPO_MAX_MEMBERS.  perm_dev: int32[n]. */
int cmbpo_draw_permutation(int32_t n, uint64_t seed, uint64_t rng_offset, const uint64_t* rng_offset_dev, uint64_t step,
                           int32_t* perm_dev, void* stream);

/* ---- device-resident rollout bookkeeping (ModelSampler.sample around env.step,
 * softlearning/samplers/model_sampler.py:311-332,:386-474; ModelBuffer.store_multiple,
 * softlearning/replay_pools/modelbuffer.py:130-173) ---- */

/* lives in device memory; initialise n_alive = B and everything else to 0 (model_sampler.py:220-263) */
typedef struct cmbpo_rollout_state {
  int32_t n_alive;        /* rows of cur_obs / path_id that are alive */
  int32_t ptr;            /* pool column of the next store == path length of every alive path (modelbuffer.py:173) */
  int64_t total_samples;  /* model_sampler.py:386 */
  int32_t n_stored_last;  /* transitions the last advance wrote */
  int32_t finished;       /* 1 once no path is alive */
  int32_t tmp[6];         /* scratch of the advance kernels (must be zero at reset) */
} cmbpo_rollout_state;

/* One sampler step AFTER cmbpo_step produced next_obs/rew/term/cost/dkl_path/ep_var_mean for the
 * first state->n_alive rows.  mode 0 = 'schedule' (length mask only), 1 = 'uncertainty'
 * (accumulated ensemble KL vs dkl_lim, model_sampler.py:311-313), 2 = 'iv_gae' (every member rolls
 * its own state, model_sampler.py:225-226: cur_obs/act/next_obs/rew/term/cost are
 * [n_members][n_paths][.] with member stride n_paths rows, pools obs/act/next_obs/rew/cost/term are
 * [n_members][n_paths][T][.] (modelbuffer.py:48-51); no uncertainty cutoff and no max_samples cap
 * (:315,:327); a path ends when ANY member's state terminates (:445)).  Pools are indexed [path][t]
 * (modelbuffer.py:45-114), path = original start-state row; see the storage note at the pool pointers.  Survivors are compacted, order preserved,
 * into new_obs / new_path_id; state is updated on the device (no host synchronisation). */
typedef struct cmbpo_rollout_io {
  int64_t n_paths;             /* B: rows at reset, bounds every grid */
  int32_t max_path_length;     /* T (modelbuffer.py:19); a path ends when length >= T-1 (model_sampler.py:437) */
  int32_t mode;
  double dkl_lim;              /* model_sampler.py:179 set_rollout_dkl (a Python float upstream) */
  int64_t max_samples;         /* 0 = None (model_sampler.py:327) */
  int32_t step_index;          /* sampler step t: slot of alive_hist_dev written by this call */
  int32_t obs_dim, act_dim;
  const float* cur_obs_dev;    /* [n_alive][obs] */
  const float* act_dev;        /* [n_alive][act] */
  const float* next_obs_dev;   /* [n_alive][obs] */
  const float* rew_dev;        /* [n_alive] */
  const uint8_t* term_dev;     /* [n_alive] */
  const float* cost_dev;       /* [n_alive] or NULL */
  const float* dkl_path_dev;   /* [n_alive] */
  const float* ep_var_mean_dev;/* [n_alive] or NULL */
  const int32_t* path_id_dev;  /* [n_alive] original path of each alive row */
  void* state_dev;             /* cmbpo_rollout_state */
  double* dkl_accum_dev;       /* [B] _dyn_dkl_path: np.zeros => float64 upstream (model_sampler.py:244,:415) */
  /* Model pools.  LOGICAL index is the reference's [path b][column t] (modelbuffer.py:45-114); the STORAGE is
   * time-major, element (b, t) at ((t * n_paths) + b) * dim, so that the stores of one sampler step (all alive
   * paths, one column) are contiguous.  'iv_gae': one such [T][n_paths][dim] block per member. */
  float* pool_obs_dev; float* pool_act_dev; float* pool_next_obs_dev;   /* [T][B][dim] */
  float* pool_rew_dev; float* pool_cost_dev; uint8_t* pool_term_dev;    /* [T][B] */
  float* pool_dyn_error_dev; uint8_t* pool_populated_dev;               /* [T][B] */
  float* new_obs_dev;          /* [n_alive'][obs] compacted survivors */
  int32_t* new_path_id_dev;    /* [n_alive'] */
  void* scratch_dev; size_t scratch_bytes;   /* >= cmbpo_rollout_scratch_bytes(n_paths) */
  int32_t* alive_hist_dev;     /* optional [T]: rows alive when step `step_index` began (diagnostics / FLOP accounting) */
  int32_t n_members;           /* mode 2 only: E (leading axis of the step arrays and pools); 0/1 otherwise */
  /* Optional critic values (all NULL = not tracked).  The sampler stores V(s_t), VC(s_t) with every transition
   * (model_sampler.py:275-294, modelbuffer.py:141-146) and hands bootstrap values to finish_path_multiple when a
   * path ends (model_sampler.py:497-548).  val_cur/cval_cur: values of the current states, val_next/cval_next: of
   * next_obs ([n_alive], 2-D modes: already gathered per path with v_inds; mode 2: [E][n_paths]).  Survivors'
   * next values are compacted into new_val/new_cval (the next step's val_cur).  pool_val/pool_cval use the pool
   * storage; finish_ptr [n_paths], last_val / last_cval [E][n_paths] are what cmbpo_finish_paths consumes. */
  const float* val_cur_dev; const float* cval_cur_dev;
  const float* val_next_dev; const float* cval_next_dev;
  float* new_val_dev; float* new_cval_dev;
  float* pool_val_dev; float* pool_cval_dev;
  int32_t* finish_ptr_dev;
  float* last_val_dev; float* last_cval_dev;
} cmbpo_rollout_io;
size_t cmbpo_rollout_scratch_bytes(int64_t n_paths);
int cmbpo_rollout_advance(const cmbpo_rollout_io* io, void* stream);
/* ModelSampler.finish_all_paths (softlearning/samplers/model_sampler.py:550-570): the rollout loop was left with
 * paths still alive (CMBPO._train breaks on alive_ratio <= 0.1 or on the sample budget, mbpo/algorithms/cmbpo.py:339-348);
 * every live path is finished at the current pool column with V(current state), VC(current state) as bootstrap values
 * (it did not terminate by the environment).  Uses state_dev, path_id_dev, val_cur_dev / cval_cur_dev (values of the
 * current states of the live rows), finish_ptr_dev, last_val_dev / last_cval_dev, n_paths, mode, n_members of `io`;
 * afterwards no path is alive (state: n_alive = 0, finished = 1).  A no-op when nothing is alive. */
int cmbpo_rollout_finish_all(const cmbpo_rollout_io* io, void* stream);

/* Populated transitions of a model pool -> one dense fp32 record matrix in [path, t] order: what
 * ModelBuffer.get selects with populated_mask (softlearning/replay_pools/modelbuffer.py:348-405), laid out
 * as the send buffer of the multi-GPU all-gather.  Record (width cmbpo_record_width):
 *   obs[E*obs] | act[E*act] | next_obs[E*obs] | rew[E] | cost[E] | dyn_error | term[E] | path | t
 * E = n_members for 'iv_gae' pools, else 1 (pool storage as in cmbpo_rollout_io: time-major); path (= row_offset + pool row) and t are
 * int32 bit patterns.  count_dev receives the number of populated transitions; nothing is written for a
 * block of paths whose records would not fit `capacity` (compare count with capacity afterwards). */
typedef struct cmbpo_pack_io {
  int64_t n_paths;
  int32_t max_path_length;
  int32_t obs_dim, act_dim;
  int32_t n_members;
  int64_t row_offset;
  const float* pool_obs_dev; const float* pool_act_dev; const float* pool_next_obs_dev;
  const float* pool_rew_dev; const float* pool_cost_dev; const uint8_t* pool_term_dev;
  const float* pool_dyn_error_dev; const uint8_t* pool_populated_dev;
  float* records_dev;          /* [capacity][cmbpo_record_width] */
  int64_t capacity;
  int64_t* count_dev;          /* [1] */
  void* scratch_dev; size_t scratch_bytes;   /* >= cmbpo_pack_scratch_bytes(n_paths) */
  /* compact 1: inside a path next_obs[t] IS obs[t+1] (the sampler feeds next_obs back as the current observation,
   * model_sampler.py:463-467), so the record drops the next_obs block (width cmbpo_record_width_compact) and
   * last_next_obs_dev [n_paths][E*obs] receives the next_obs of every path's LAST stored transition (zeros for a path
   * without transitions): next_obs of record k = obs of record k+1 when both belong to one path, else last_next_obs[path].
   * 40 % fewer bytes to the host / over NVLink at the BASELINE shapes.  compact 0: the full record above. */
  int32_t compact;
  float* last_next_obs_dev;
} cmbpo_pack_io;
int32_t cmbpo_record_width(int32_t obs_dim, int32_t act_dim, int32_t n_members);
int32_t cmbpo_record_width_compact(int32_t obs_dim, int32_t act_dim, int32_t n_members);
size_t cmbpo_pack_scratch_bytes(int64_t n_paths);
int cmbpo_pack_records(const cmbpo_pack_io* io, void* stream);

/* ModelBuffer.finish_path_multiple for every path of a pool at once (softlearning/replay_pools/modelbuffer.py:
 * 176-346; softlearning/policies/safe_utils/utils.py:18-188).  Path b was finished when the pool column was
 * finish_ptr[b] (its slice is [0, finish_ptr[b]), modelbuffer.py:199) with bootstrap values last_val / last_cval
 * (model_sampler.py:497-548).  mode 0/1: GAE, adv = discount_cumsum(rew + gamma*V' - V, gamma*lam), ret = adv + V,
 * same for cost with cost_gamma / cost_lam.  mode 2 ('iv_gae'): inverse-variance weighted advantages of member
 * model_ind's trajectory, weights from the epistemic variance over members of the h-step bootstrapped returns,
 * plus ret_var / roll_lengths and the horizon trim of populated (first column whose normalised return variance
 * exceeds max_uncertainty_r / _c; a path with no such column is dropped entirely, as upstream, argmax quirk).
 * All arrays use the rollout's time-major storage: element (member e, path b, column t) at (e*T + t)*n_paths + b. */
typedef struct cmbpo_gae_io {
  int64_t n_paths;
  int32_t max_path_length;     /* T (<= 64 in mode 2) */
  int32_t mode;                /* 0 / 1: GAE; 2: 'iv_gae' */
  int32_t n_members;           /* mode 2: E */
  int32_t model_ind;           /* mode 2: ModelBuffer.model_ind (modelbuffer.py:26) */
  int32_t cares_about_cost;
  double gamma, lam, cost_gamma, cost_lam;            /* ModelBuffer.initialize (modelbuffer.py:33-43) */
  float max_uncertainty_r, max_uncertainty_c;         /* modelbuffer.py:14-15 */
  const float* rew_dev; const float* val_dev; const float* cost_dev; const float* cval_dev;   /* [E][T][B] */
  const int32_t* finish_ptr_dev;                      /* [B] */
  const float* last_val_dev; const float* last_cval_dev;                                      /* [E][B] */
  float* adv_dev; float* ret_dev; float* cadv_dev; float* cret_dev;                           /* [T][B] */
  float* ret_var_dev; float* cret_var_dev; float* roll_lengths_dev; float* croll_lengths_dev; /* [T][B], mode 2 */
  uint8_t* populated_dev;                             /* [T][B], trimmed in place in mode 2; may be NULL */
} cmbpo_gae_io;
int cmbpo_finish_paths(const cmbpo_gae_io* io, void* stream);

/* ModelBuffer.get's advantage normalisation over the populated entries (softlearning/replay_pools/modelbuffer.py:
 * 356-372; mpi_statistics_scalar, softlearning/policies/safe_utils/mpi_tools.py:71-93): adv <- (adv - mean) / (std + 1e-8)
 * with the float32 mean and population std, cadv <- cadv - mean(cadv) (centred, not rescaled), in place.
 * stats_dev [6] receives adv_mean, adv_std, cadv_mean, ret_mean, cret_mean, n.  Arrays are flat (any layout, the same
 * for all of them), n_elems entries; ret_dev / cret_dev may be NULL. */
typedef struct cmbpo_norm_io {
  int64_t n_elems;
  float* adv_dev; float* cadv_dev;
  const float* ret_dev; const float* cret_dev;
  const uint8_t* populated_dev;
  float* stats_dev;
  void* scratch_dev; size_t scratch_bytes;    /* >= cmbpo_normalize_scratch_bytes() */
} cmbpo_norm_io;
size_t cmbpo_normalize_scratch_bytes(void);
int cmbpo_normalize_advantages(const cmbpo_norm_io* io, void* stream);

/* cmbpo_normalize_advantages without any allocation on the caller's or the library's side.  This entry point has no
 * counterpart in the reference: it was added at the request of the run's operators, for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, launches and results as
 * cmbpo_normalize_advantages (bit-identical: same grid, same summation order); io->scratch_dev / scratch_bytes are ignored --
 * the scratch is a file-scope `static __device__` array of CMBPO_NORMALIZE_STATIC_SCRATCH_BYTES that every GPU loading the
 * library holds until exit.  Limits and rules:
 *   - n_elems <= CMBPO_NORMALIZE_STATIC_MAX_ELEMS, else CMBPO_ERR_UNSUPPORTED (nothing is launched);
 *   - all calls on a device share the array, which keeps the last call's contents (the call clears what must start at zero):
 *     calls that overlap on the GPU (other streams, other host threads) RACE on it -- serialise them;
 *   - the device current at the call is the one whose array is used: `stream` must belong to it.
 * err / err_len: optional buffer that receives the error text (empty on success); the call calls no cudaMalloc, cudaFree or
 * allocating library routine. */
#define CMBPO_NORMALIZE_STATIC_SCRATCH_BYTES 41216            /* 64 (ticket) + 1024 blocks x 5 doubles, rounded up to 256 */
#define CMBPO_NORMALIZE_STATIC_MAX_ELEMS (1LL << 26)          /* 67,108,864 pool entries ([T][B] flattened) */
int cmbpo_normalize_advantages_static(const cmbpo_norm_io* io, void* stream, char* err, size_t err_len);

/* Start-state sampling from the real-data archive (softlearning/replay_pools/cpobuffer.py:531-542 boltz_dist,
 * :559-610 distributed_batch_from_archive; mbpo/algorithms/cmbpo.py:328-333): weighted draw with replacement
 * (np.random.choice(archive_size, batch_size, p=dist): idx = searchsorted(cumsum(p)/sum(p), u, side='right')) and the
 * gather of the drawn rows, on an archive resident in device memory.  The row distribution is either given
 * (prob_dev, e.g. unif_dist / disc_dist) or built from the per-row epoch id and a per-epoch table
 * (boltz_dist: epoch_prob[e] = ep_probs[e] / rows_in_epoch[e], 0 for empty slots).  uniform_dev injects the
 * uniforms (parity); NULL = 53-bit Philox4x32-10 doubles keyed by (sample_offset + b, rng_offset, seed). */
#define CMBPO_MAX_FIELDS 4
typedef struct cmbpo_sample_io {
  int64_t archive_size;
  int64_t batch_size;
  const float* prob_dev;          /* [archive_size] or NULL */
  const int32_t* epoch_dev;       /* [archive_size] epoch id per row, < 0 = empty slot (cpobuffer epoch_archive) */
  const float* epoch_prob_dev;    /* [n_epochs] float32 like sample_p (cpobuffer.py:537-538) */
  int32_t n_epochs;
  int32_t rebuild_cdf;            /* 1: (re)build cdf_dev from the distribution; 0: reuse it */
  double* cdf_dev;                /* [archive_size] workspace / cache */
  const double* uniform_dev;      /* [batch_size] or NULL */
  uint64_t seed, rng_offset;
  int64_t sample_offset;          /* global index of sample 0 (shard-invariant draws) */
  const float* field_dev[CMBPO_MAX_FIELDS];   /* archive arrays [archive_size][width] (observations, mu, log_std, ...) */
  int32_t field_width[CMBPO_MAX_FIELDS];
  float* out_dev[CMBPO_MAX_FIELDS];           /* [batch_size][width] */
  int64_t* index_dev;             /* optional [batch_size] drawn row indices */
} cmbpo_sample_io;
int cmbpo_sample_archive(const cmbpo_sample_io* io, void* stream);

/* Synthetic benchmark policy a = tanh(obs @ P), P [obs][act] (not part of the reference; stands in
 * for policy.get_action_outs, model_sampler.py:275, in throughput runs).  n_slots > 1: obs / act are
 * [n_slots][n_rows][.] ('iv_gae' states), the live count applies to every slot. */
int cmbpo_stub_policy(const void* obs_dev, const void* P_dev, void* act_dev, const void* n_rows_dev,
                      int64_t n_rows, int32_t obs_dim, int32_t act_dim, int32_t n_slots, void* stream);

/* Measurement aid (no reference counterpart): while enabled, every ensemble-forward launch of this
 * handle is bracketed by CUDA events on its stream; cmbpo_profile_read waits for them, returns the
 * summed kernel time and the launch count, and clears the list.  Not usable under graph capture. */
int cmbpo_profile(void* handle, int enable);
int cmbpo_profile_read(void* handle, double* fwd_ms_total, int64_t* fwd_launches);

/* Which forward kernel this handle's net runs on (measurement aid, no reference counterpart):
 * 0 = ensemble_fwd_kernel (hidden layers <= 254 wide, two row-tile streams), 1 = wide_layer_kernel (one launch per
 * FC layer), 2 = fused_wide_kernel (hidden layers up to 512 bf16 / 256 tf32 columns, whole stack in one launch). */
int cmbpo_forward_path(void* handle);

/* watchdog word of the last launches on this handle (0 = clean), synchronises the stream */
int cmbpo_status(void* handle, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CMBPO_B200_H_ */
