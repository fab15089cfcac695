"""ctypes binding of libcmbpo_b200.so (C ABI declared in include/cmbpo_b200.h).

There is no fallback: if the library was not built (``python -c "import __graft_entry__ as g;
g.build()"`` or ``make``) or no sm_100 GPU is visible, every compute entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# CMBPO_LIB: debug builds of the same ABI (e.g. build/libcmbpo_b200_trace.so, `make trace`)
LIB_PATH = os.environ.get("CMBPO_LIB") or os.path.join(_HERE, "libcmbpo_b200.so")

MAX_LAYERS = 8
MAX_SEGMENTS = 3

ACT_IDS = {None: 0, "swish": 1, "ReLU": 2, "tanh": 3, "sigmoid": 4}  # fc.py:14-21

# mbpo/static/__init__.py:7-24
ENV_IDS = {
    None: 0, "none": 0, "ant": 1, "antsafe": 2, "halfcheetah": 3, "halfcheetahsafe": 4,
    "humanoid": 5, "humanoidsafe": 6, "hopper": 7, "hoppersafe": 8, "walker2d": 9,
}
ENVS_WITH_STATIC_COST = {"antsafe", "halfcheetahsafe", "humanoidsafe", "hoppersafe"}


class CmbpoError(RuntimeError):
    pass


DTYPE_IDS = {"bf16": 0, "tf32": 1}   # CMBPO_DTYPE_*: tensor-core operand type


class cmbpo_cfg(C.Structure):
    _fields_ = [
        ("num_members", C.c_int32), ("num_layers", C.c_int32), ("in_dim", C.c_int32),
        ("out_dim", C.c_int32), ("probabilistic", C.c_int32),
        ("layer_width", C.c_int32 * MAX_LAYERS), ("layer_act", C.c_int32 * MAX_LAYERS),
        ("operand_dtype", C.c_int32),
    ]


class cmbpo_predict_io(C.Structure):
    _fields_ = [
        ("seg_dev", C.c_void_p * MAX_SEGMENTS), ("seg_width", C.c_int32 * MAX_SEGMENTS),
        ("n_rows", C.c_int64), ("ndim", C.c_int32), ("agg", C.c_int32),
        ("mean_dev", C.c_void_p), ("var_dev", C.c_void_p),
        ("workspace_dev", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("n_rows_dev", C.c_void_p),
    ]


class cmbpo_step_io(C.Structure):
    _fields_ = [
        ("obs_dev", C.c_void_p), ("act_dev", C.c_void_p), ("n_rows", C.c_int64),
        ("n_rows_dev", C.c_void_p), ("ndim", C.c_int32), ("perm", C.POINTER(C.c_int32)),
        ("elite_idx_dev", C.c_void_p), ("row_id_dev", C.c_void_p), ("noise_dev", C.c_void_p),
        ("seed", C.c_uint64), ("rng_offset", C.c_uint64), ("row_offset", C.c_int64),
        ("env_id", C.c_int32), ("cost_mode", C.c_int32),
        ("next_obs_dev", C.c_void_p), ("rew_dev", C.c_void_p), ("term_dev", C.c_void_p),
        ("cost_dev", C.c_void_p), ("dkl_path_dev", C.c_void_p), ("ep_var_mean_dev", C.c_void_p),
        ("ep_var_dev", C.c_void_p), ("scalars_dev", C.c_void_p),
        ("workspace_dev", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("sample_mode", C.c_int32), ("bound_logvar", C.c_int32), ("max_logvar", C.c_float), ("min_logvar", C.c_float),
        ("rng_offset_dev", C.c_void_p), ("perm_dev", C.c_void_p),
    ]


class cmbpo_rollout_io(C.Structure):
    _fields_ = [
        ("n_paths", C.c_int64), ("max_path_length", C.c_int32), ("mode", C.c_int32),
        ("dkl_lim", C.c_double), ("max_samples", C.c_int64), ("step_index", C.c_int32),
        ("obs_dim", C.c_int32), ("act_dim", C.c_int32),
        ("cur_obs_dev", C.c_void_p), ("act_dev", C.c_void_p), ("next_obs_dev", C.c_void_p),
        ("rew_dev", C.c_void_p), ("term_dev", C.c_void_p), ("cost_dev", C.c_void_p),
        ("dkl_path_dev", C.c_void_p), ("ep_var_mean_dev", C.c_void_p),
        ("path_id_dev", C.c_void_p), ("state_dev", C.c_void_p), ("dkl_accum_dev", C.c_void_p),
        ("pool_obs_dev", C.c_void_p), ("pool_act_dev", C.c_void_p), ("pool_next_obs_dev", C.c_void_p),
        ("pool_rew_dev", C.c_void_p), ("pool_cost_dev", C.c_void_p), ("pool_term_dev", C.c_void_p),
        ("pool_dyn_error_dev", C.c_void_p), ("pool_populated_dev", C.c_void_p),
        ("new_obs_dev", C.c_void_p), ("new_path_id_dev", C.c_void_p),
        ("scratch_dev", C.c_void_p), ("scratch_bytes", C.c_size_t),
        ("alive_hist_dev", C.c_void_p), ("n_members", C.c_int32),
        ("val_cur_dev", C.c_void_p), ("cval_cur_dev", C.c_void_p), ("val_next_dev", C.c_void_p),
        ("cval_next_dev", C.c_void_p), ("new_val_dev", C.c_void_p), ("new_cval_dev", C.c_void_p),
        ("pool_val_dev", C.c_void_p), ("pool_cval_dev", C.c_void_p), ("finish_ptr_dev", C.c_void_p),
        ("last_val_dev", C.c_void_p), ("last_cval_dev", C.c_void_p),
    ]


class cmbpo_pack_io(C.Structure):
    _fields_ = [
        ("n_paths", C.c_int64), ("max_path_length", C.c_int32), ("obs_dim", C.c_int32), ("act_dim", C.c_int32),
        ("n_members", C.c_int32), ("row_offset", C.c_int64),
        ("pool_obs_dev", C.c_void_p), ("pool_act_dev", C.c_void_p), ("pool_next_obs_dev", C.c_void_p),
        ("pool_rew_dev", C.c_void_p), ("pool_cost_dev", C.c_void_p), ("pool_term_dev", C.c_void_p),
        ("pool_dyn_error_dev", C.c_void_p), ("pool_populated_dev", C.c_void_p),
        ("records_dev", C.c_void_p), ("capacity", C.c_int64), ("count_dev", C.c_void_p),
        ("scratch_dev", C.c_void_p), ("scratch_bytes", C.c_size_t),
        ("compact", C.c_int32), ("last_next_obs_dev", C.c_void_p),
    ]


class cmbpo_actor_io(C.Structure):
    _fields_ = [
        ("mu_dev", C.c_void_p), ("log_std_dev", C.c_void_p), ("noise_dev", C.c_void_p),
        ("n_rows", C.c_int64), ("n_rows_dev", C.c_void_p), ("act_dim", C.c_int32), ("n_slots", C.c_int32),
        ("deterministic", C.c_int32), ("row_id_dev", C.c_void_p), ("row_offset", C.c_int64),
        ("seed", C.c_uint64), ("rng_offset", C.c_uint64), ("rng_offset_dev", C.c_void_p),
        ("pi_dev", C.c_void_p), ("logp_pi_dev", C.c_void_p),
    ]


class cmbpo_gae_io(C.Structure):
    _fields_ = [
        ("n_paths", C.c_int64), ("max_path_length", C.c_int32), ("mode", C.c_int32), ("n_members", C.c_int32),
        ("model_ind", C.c_int32), ("cares_about_cost", C.c_int32),
        ("gamma", C.c_double), ("lam", C.c_double), ("cost_gamma", C.c_double), ("cost_lam", C.c_double),
        ("max_uncertainty_r", C.c_float), ("max_uncertainty_c", C.c_float),
        ("rew_dev", C.c_void_p), ("val_dev", C.c_void_p), ("cost_dev", C.c_void_p), ("cval_dev", C.c_void_p),
        ("finish_ptr_dev", C.c_void_p), ("last_val_dev", C.c_void_p), ("last_cval_dev", C.c_void_p),
        ("adv_dev", C.c_void_p), ("ret_dev", C.c_void_p), ("cadv_dev", C.c_void_p), ("cret_dev", C.c_void_p),
        ("ret_var_dev", C.c_void_p), ("cret_var_dev", C.c_void_p), ("roll_lengths_dev", C.c_void_p),
        ("croll_lengths_dev", C.c_void_p), ("populated_dev", C.c_void_p),
    ]


class cmbpo_norm_io(C.Structure):
    _fields_ = [
        ("n_elems", C.c_int64), ("adv_dev", C.c_void_p), ("cadv_dev", C.c_void_p), ("ret_dev", C.c_void_p),
        ("cret_dev", C.c_void_p), ("populated_dev", C.c_void_p), ("stats_dev", C.c_void_p),
        ("scratch_dev", C.c_void_p), ("scratch_bytes", C.c_size_t),
    ]


MAX_FIELDS = 4


class cmbpo_sample_io(C.Structure):
    _fields_ = [
        ("archive_size", C.c_int64), ("batch_size", C.c_int64), ("prob_dev", C.c_void_p), ("epoch_dev", C.c_void_p),
        ("epoch_prob_dev", C.c_void_p), ("n_epochs", C.c_int32), ("rebuild_cdf", C.c_int32), ("cdf_dev", C.c_void_p),
        ("uniform_dev", C.c_void_p), ("seed", C.c_uint64), ("rng_offset", C.c_uint64), ("sample_offset", C.c_int64),
        ("field_dev", C.c_void_p * MAX_FIELDS), ("field_width", C.c_int32 * MAX_FIELDS),
        ("out_dev", C.c_void_p * MAX_FIELDS), ("index_dev", C.c_void_p),
    ]


# every symbol include/cmbpo_b200.h declares: (restype, argtypes)
SYMBOLS = {
    "cmbpo_create": (C.c_int, [C.POINTER(cmbpo_cfg), C.c_int, C.POINTER(C.c_void_p)]),
    "cmbpo_destroy": (None, [C.c_void_p]),
    "cmbpo_last_error": (C.c_char_p, []),
    "cmbpo_device_ok": (C.c_int, [C.c_int]),
    "cmbpo_set_weights": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cmbpo_set_scalers": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_float, C.c_void_p]),
    "cmbpo_set_elites": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.c_int32, C.c_void_p]),
    "cmbpo_workspace_bytes": (C.c_size_t, [C.c_void_p, C.c_int64]),
    "cmbpo_predict": (C.c_int, [C.c_void_p, C.POINTER(cmbpo_predict_io), C.c_void_p]),
    "cmbpo_step": (C.c_int, [C.c_void_p, C.POINTER(cmbpo_step_io), C.c_void_p]),
    "cmbpo_draw_permutation": (C.c_int, [C.c_int32, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    "cmbpo_status": (C.c_int, [C.c_void_p, C.c_void_p]),
    "cmbpo_forward_path": (C.c_int, [C.c_void_p]),
    "cmbpo_profile": (C.c_int, [C.c_void_p, C.c_int]),
    "cmbpo_profile_read": (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "cmbpo_rollout_scratch_bytes": (C.c_size_t, [C.c_int64]),
    "cmbpo_rollout_advance": (C.c_int, [C.POINTER(cmbpo_rollout_io), C.c_void_p]),
    "cmbpo_actor_head": (C.c_int, [C.POINTER(cmbpo_actor_io), C.c_void_p]),
    "cmbpo_rollout_finish_all": (C.c_int, [C.POINTER(cmbpo_rollout_io), C.c_void_p]),
    "cmbpo_record_width": (C.c_int32, [C.c_int32, C.c_int32, C.c_int32]),
    "cmbpo_record_width_compact": (C.c_int32, [C.c_int32, C.c_int32, C.c_int32]),
    "cmbpo_pack_scratch_bytes": (C.c_size_t, [C.c_int64]),
    "cmbpo_pack_records": (C.c_int, [C.POINTER(cmbpo_pack_io), C.c_void_p]),
    "cmbpo_finish_paths": (C.c_int, [C.POINTER(cmbpo_gae_io), C.c_void_p]),
    "cmbpo_normalize_scratch_bytes": (C.c_size_t, []),
    "cmbpo_normalize_advantages": (C.c_int, [C.POINTER(cmbpo_norm_io), C.c_void_p]),
    "cmbpo_normalize_advantages_static": (C.c_int, [C.POINTER(cmbpo_norm_io), C.c_void_p, C.c_char_p, C.c_size_t]),
    "cmbpo_sample_archive": (C.c_int, [C.POINTER(cmbpo_sample_io), C.c_void_p]),
    "cmbpo_stub_policy": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                    C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
}

_lib = None


def load():
    """dlopen the library and bind prototypes (no GPU needed for this)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CmbpoError(
            f"{LIB_PATH} is missing: build it with `make` or `__graft_entry__.build()`; "
            "constrainedmbpo_b200 has no CPU or PyTorch fallback for the rollout path")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int):
    if rc != 0:
        msg = load().cmbpo_last_error()
        raise CmbpoError(f"libcmbpo_b200 error {rc}: {msg.decode() if msg else '?'}")


def require_device(device_index: int = 0):
    lib = load()
    if not lib.cmbpo_device_ok(device_index):
        raise CmbpoError("no sm_100 (B200) CUDA device visible: the rollout path is CUDA-only, "
                         "there is no CPU fallback")
    return lib


# ---- NumPy boundary -----------------------------------------------------------------------------
# The reference's callers (ModelSampler.sample, MBPO._rollout_model) hand NumPy arrays to FakeEnv.step / BNN.predict and
# get NumPy arrays back, every sampler step.  `tensor.cpu().numpy()` lands in freshly allocated pageable memory: the driver
# bounces the copy through its own pinned buffers and every first touch of the new pages faults (16.8 MB of predict outputs:
# ~4 ms; 133 MB of step outputs at 1 M rows: tens of ms).  Results therefore go to pinned host blocks of torch's caching
# host allocator (reused from call to call, handed back when the NumPy array dies): all copies are enqueued first, one
# synchronisation follows.
def to_host(tensors):
    """Device tensors -> NumPy arrays backed by cached pinned host memory (None stays None)."""
    import torch
    hosts, dev = [], None
    for t in tensors:
        if t is None:
            hosts.append(None)
            continue
        dev = t.device
        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        h.copy_(t, non_blocking=True)
        hosts.append(h)
    if dev is not None:
        torch.cuda.current_stream(dev).synchronize()
    return [None if h is None else h.numpy() for h in hosts]
