"""``ModelBuffer.finish_path_multiple`` on the device (softlearning/replay_pools/modelbuffer.py:176-346): GAE for
the 'schedule' / 'uncertainty' pools, inverse-variance weighted GAE + horizon trim for the 'iv_gae' pools.
A thin ctypes client of ``cmbpo_finish_paths``; PyTorch only owns the buffers."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

MODES = {"schedule": 0, "uncertainty": 1, "iv_gae": 2}


def _time_major(x, lead):
    """Logical [(E,) B, T] tensor -> storage [(E,) T, B] (no copy when it already is a ModelRollout pool view)."""
    s = x.transpose(lead, lead + 1)
    return s if s.is_contiguous() else s.contiguous()


def finish_paths(rew, val, cost, cval, finish_ptr, last_val, last_cval, rollout_mode, gamma=0.99, lam=0.95,
                 cost_gamma=0.99, cost_lam=0.95, model_ind=0, cares_about_cost=True, max_uncertainty_r=5.5,
                 max_uncertainty_c=5.5, populated=None):
    """All arguments are CUDA tensors shaped like the reference's buffers: ``rew/val/cost/cval`` ``[B, T]`` or, for
    'iv_gae', ``[E, B, T]``; ``finish_ptr`` int ``[B]`` (pool column when the path was finished); ``last_val`` /
    ``last_cval`` ``[B]`` or ``[E, B]`` (model_sampler.py:497-548); ``populated`` uint8/bool ``[B, T]`` (trimmed copy
    returned for 'iv_gae').  Returns a dict of float32 ``[B, T]`` tensors: adv, ret, cadv, cret (+ ret_var, cret_var,
    roll_lengths, croll_lengths, populated for 'iv_gae')."""
    lib = _lib.load()
    mode = MODES[rollout_mode]
    iv = mode == 2
    lead = 1 if iv else 0
    dev = rew.device
    B, T = rew.shape[-2:]
    f32 = lambda t: t.to(device=dev, dtype=torch.float32)
    srcs = [_time_major(f32(t), lead) for t in (rew, val, cost, cval)]
    fp = finish_ptr.to(device=dev, dtype=torch.int32).contiguous()
    lv, lc = f32(last_val).contiguous(), f32(last_cval).contiguous()
    names = ["adv", "ret", "cadv", "cret"] + (["ret_var", "cret_var", "roll_lengths", "croll_lengths"] if iv else [])
    outs = {k: torch.zeros(T, B, dtype=torch.float32, device=dev) for k in names}
    pop = None
    if iv and populated is not None:
        pop = _time_major(populated.to(device=dev, dtype=torch.uint8), 0).clone()
    io = _lib.cmbpo_gae_io()
    io.n_paths, io.max_path_length, io.mode = B, T, mode
    io.n_members = rew.shape[0] if iv else 1
    io.model_ind, io.cares_about_cost = int(model_ind), int(bool(cares_about_cost))
    io.gamma, io.lam, io.cost_gamma, io.cost_lam = gamma, lam, cost_gamma, cost_lam
    io.max_uncertainty_r, io.max_uncertainty_c = max_uncertainty_r, max_uncertainty_c
    io.rew_dev, io.val_dev, io.cost_dev, io.cval_dev = (t.data_ptr() for t in srcs)
    io.finish_ptr_dev, io.last_val_dev, io.last_cval_dev = fp.data_ptr(), lv.data_ptr(), lc.data_ptr()
    io.adv_dev, io.ret_dev, io.cadv_dev, io.cret_dev = (outs[k].data_ptr() for k in names[:4])
    if iv:
        io.ret_var_dev, io.cret_var_dev = outs["ret_var"].data_ptr(), outs["cret_var"].data_ptr()
        io.roll_lengths_dev, io.croll_lengths_dev = outs["roll_lengths"].data_ptr(), outs["croll_lengths"].data_ptr()
        io.populated_dev = pop.data_ptr() if pop is not None else None
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(lib.cmbpo_finish_paths(C.byref(io), stream))
    res = {k: v.t() for k, v in outs.items()}     # logical [B, T] views of the time-major storage
    if pop is not None:
        res["populated"] = pop.t()
    return res


def normalize_advantages_sharded(adv, cadv, populated, ret=None, cret=None, group=None):
    """The same normalisation when the pool is row-sharded over the ranks of ``group``: the reference's statistics are
    GLOBAL (``mpi_statistics_scalar`` all-reduces sum / count, then the squared deviations,
    softlearning/policies/safe_utils/mpi_tools.py:71-93), so the five partial sums and the count are all-reduced (fp64,
    6 + 1 numbers) and every rank standardises its shard with the global float32 mean / population std.  Plain torch ops
    on the tensors' device (NCCL for CUDA shards, gloo for the CPU tests): this runs once per rollout, not per step."""
    import torch.distributed as dist
    m = populated.bool()
    f64 = torch.float64
    zero = torch.zeros((), dtype=f64, device=adv.device)
    s = torch.stack([adv[m].to(f64).sum(), cadv[m].to(f64).sum(),
                     ret[m].to(f64).sum() if ret is not None else zero, cret[m].to(f64).sum() if cret is not None else zero,
                     m.sum().to(f64)])
    dist.all_reduce(s, group=group)
    n = s[4]
    if float(n) <= 0:
        return torch.zeros(6, dtype=torch.float32, device=adv.device)
    mean = (s[0] / n).to(torch.float32)
    cmean = (s[1] / n).to(torch.float32)
    sq = ((adv[m] - mean).to(f64) ** 2).sum().reshape(1)
    dist.all_reduce(sq, group=group)
    std = torch.sqrt(sq[0] / n).to(torch.float32)
    adv[m] = (adv[m] - mean) / (std + 1e-8)      # modelbuffer.py:361
    cadv[m] = cadv[m] - cmean                    # :367
    return torch.stack([mean, std, cmean, (s[2] / n).to(torch.float32), (s[3] / n).to(torch.float32), n.to(torch.float32)])


def normalize_advantages(adv, cadv, populated, ret=None, cret=None, group=None):
    """``ModelBuffer.get``'s advantage trick (modelbuffer.py:356-372) in place on CUDA tensors of one common layout
    (e.g. the ``[B, T]`` views ``finish_paths`` returns): ``adv`` standardised with the float32 mean / population std
    over the populated entries, ``cadv`` centred only.  Returns the statistics as a CUDA float32 tensor
    ``[adv_mean, adv_std, cadv_mean, ret_mean, cret_mean, n]``.  ``group``: the process group the rows are sharded
    over -- statistics then are global (``normalize_advantages_sharded``)."""
    if group is not None:
        import torch.distributed as dist
        if dist.get_world_size(group) > 1:
            return normalize_advantages_sharded(adv, cadv, populated, ret, cret, group)
    lib = _lib.load()
    dev = adv.device
    base = lambda t: t if t.is_contiguous() else (t.t() if t.dim() == 2 and t.t().is_contiguous() else None)
    ts = [base(t) for t in (adv, cadv, populated) + ((ret,) if ret is not None else ()) + ((cret,) if cret is not None else ())]
    if any(t is None for t in ts) or len({(tuple(t.shape), t.stride()) for t in ts}) != 1:
        raise ValueError("normalize_advantages: all arrays must share one dense layout")
    pop = ts[2]
    if pop.dtype != torch.uint8:
        raise ValueError("populated must be uint8")
    stats = torch.zeros(6, dtype=torch.float32, device=dev)
    scratch = torch.empty(int(lib.cmbpo_normalize_scratch_bytes()), dtype=torch.uint8, device=dev)
    io = _lib.cmbpo_norm_io()
    io.n_elems = adv.numel()
    io.adv_dev, io.cadv_dev, io.populated_dev = ts[0].data_ptr(), ts[1].data_ptr(), pop.data_ptr()
    io.ret_dev = ts[3].data_ptr() if ret is not None else None
    io.cret_dev = ts[3 + (ret is not None)].data_ptr() if cret is not None else None
    io.stats_dev, io.scratch_dev, io.scratch_bytes = stats.data_ptr(), scratch.data_ptr(), scratch.numel()
    _lib.check(lib.cmbpo_normalize_advantages(C.byref(io), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
    return stats
