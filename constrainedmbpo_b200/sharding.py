"""Row sharding of the rollout batch across the GPUs of one box and the all-gather of generated
transitions into the replicated model pool.

The reference has no multi-device path (mpi4py at world size 1, softlearning/policies/safe_utils/
mpi_tools.py:7-93), so the parity statement is: the rank-order concatenation of the per-rank
transition lists equals the single-device result (same kernels, same row -> Philox counter map via
``row_offset``).  Everything here is device-agnostic torch plumbing (NCCL on CUDA tensors, gloo on
CPU tensors in the tests); no arithmetic of the path happens in this file.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

RECORD_FIELDS = ("obs", "act", "next_obs", "rew", "cost", "dyn_error", "term")


def shard_bounds(n_rows: int, world: int, rank: int):
    """Contiguous row block of ``rank``: the first ``n_rows % world`` ranks get one extra row."""
    base, rem = divmod(int(n_rows), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def pack_records(pool: dict, row_offset: int = 0):
    """Populated transitions of a ``[B,T,.]`` pool (modelbuffer.py:45-114, filtered by
    ``populated_mask`` like ``ModelBuffer.get``, modelbuffer.py:348-405) as one fp32 matrix
    ``[n, 2*obs+act+6]``: obs | act | next_obs | rew | cost | dyn_error | term | path | t, in
    [path, t] order; ``path`` is the GLOBAL start-state index (exact in fp32 below 2**24 rows per
    box; the int64 twin ``coords`` carries it losslessly)."""
    mask = pool["populated"].bool()
    path, t = mask.nonzero(as_tuple=True)
    cols = []
    for k in RECORD_FIELDS:
        v = pool[k]
        if k != "dyn_error" and v.dim() > mask.dim() and tuple(v.shape[1:3]) == tuple(mask.shape):
            v = v.movedim(0, 2)      # 'iv_gae' pool [E,B,T,.] -> [B,T,E,.]: member-major columns per field
        v = v[mask]
        cols.append(v.reshape(v.shape[0], -1).to(torch.float32))
    rec = torch.cat(cols, dim=1)
    coords = torch.stack([path + int(row_offset), t], dim=1).to(torch.int64)
    return rec, coords


def unpack_records(rec: torch.Tensor, obs_dim: int, act_dim: int):
    o = 0
    out = {}
    for k, wdt in (("obs", obs_dim), ("act", act_dim), ("next_obs", obs_dim), ("rew", 1), ("cost", 1),
                   ("dyn_error", 1), ("term", 1)):
        v = rec[:, o:o + wdt]
        out[k] = v if wdt > 1 else v[:, 0]
        o += wdt
    out["term"] = out["term"] != 0
    return out


def _all_gather_rows(x: torch.Tensor, group=None):
    """Variable-length all-gather along dim 0 (counts exchanged first, payload padded to the max)."""
    world = dist.get_world_size(group)
    n = torch.tensor([x.shape[0]], dtype=torch.int64, device=x.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    m = max(counts) if counts else 0
    pad = torch.zeros((m,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    pad[: x.shape[0]] = x
    if x.is_cuda:
        full = torch.empty((world * m,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(full, pad, group=group)
        parts = [full[r * m: r * m + counts[r]] for r in range(world)]
    else:
        bufs = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(bufs, pad, group=group)
        parts = [bufs[r][: counts[r]] for r in range(world)]
    return torch.cat(parts, dim=0), counts


def gather_transitions(pool: dict, row_offset: int = 0, group=None):
    """All ranks' populated transitions, concatenated in rank order: ``(records, coords, counts)``."""
    rec, coords = pack_records(pool, row_offset)
    rec_all, counts = _all_gather_rows(rec, group)
    coords_all, _ = _all_gather_rows(coords, group)
    return rec_all, coords_all, counts


def broadcast_model_state(model, src=0, group=None):
    """Replicate a trained ensemble from rank ``src`` to every rank (weights / scalers / elites are
    replicated, one broadcast after each ``BNN.train()``): the fp32 master weights and biases go over NCCL as device
    tensors, scaler state and the elite list as a small object; every rank then re-packs its bf16 weight image
    (``load_state``).  ~3.9 MB at 4x200, ~91 MB at 4x1024."""
    rank = dist.get_rank(group)
    for t in list(model.weights) + list(model.biases):
        dist.broadcast(t, src=src, group=group)
    meta = [None]
    if rank == src:
        sc = lambda s: None if s is None else (s.cached_mu, s.cached_var)
        meta = [dict(scaler_in=sc(model.scaler_in), scaler_out=sc(model.scaler_out), elites=list(model._model_inds))]
    dist.broadcast_object_list(meta, src=src, group=group)
    m = meta[0]
    model.load_state(list(model.weights), list(model.biases), scaler_in=m["scaler_in"], scaler_out=m["scaler_out"],
                     elites=m["elites"])


def record_width(obs_dim, act_dim, n_members=1, compact=False):
    """Columns of one packed record (``cmbpo_record_width`` / ``_compact``): the RECORD_FIELDS payload + (path, t);
    compact records carry no next_obs block (see ``reconstruct_next_obs``)."""
    E = max(int(n_members), 1)
    return E * ((1 if compact else 2) * obs_dim + act_dim + 3) + 1 + 2


def pack_records_cuda(pool: dict, row_offset, obs_dim, act_dim, n_members, out, count, scratch, stream=None,
                      last_next_obs=None):
    """``cmbpo_pack_records``: the CUDA twin of ``pack_records`` + coords in one launch sequence, no host
    synchronisation.  ``out`` [capacity, W] fp32, ``count`` int64 [1], ``scratch`` uint8 (all CUDA).
    ``last_next_obs`` ([n_paths, E*obs] CUDA fp32): write COMPACT records (no next_obs block) plus the next_obs of every
    path's last transition into it."""
    import ctypes as C
    from . import _lib
    lib = _lib.load()
    B, T = pool["populated"].shape
    if pool["populated"].stride() != (1, B) or pool["obs"].stride()[-3:] != (obs_dim, B * obs_dim, 1):
        raise ValueError("pack_records_cuda expects ModelRollout pools (logical [B,T,.] views of time-major storage)")
    io = _lib.cmbpo_pack_io()
    io.n_paths, io.max_path_length, io.obs_dim, io.act_dim = B, T, obs_dim, act_dim
    io.n_members, io.row_offset = (n_members if n_members > 1 else 0), int(row_offset)
    io.pool_obs_dev, io.pool_act_dev, io.pool_next_obs_dev = pool["obs"].data_ptr(), pool["act"].data_ptr(), pool["next_obs"].data_ptr()
    io.pool_rew_dev, io.pool_cost_dev, io.pool_term_dev = pool["rew"].data_ptr(), pool["cost"].data_ptr(), pool["term"].data_ptr()
    io.pool_dyn_error_dev, io.pool_populated_dev = pool["dyn_error"].data_ptr(), pool["populated"].data_ptr()
    io.records_dev, io.capacity, io.count_dev = out.data_ptr(), out.shape[0], count.data_ptr()
    io.scratch_dev, io.scratch_bytes = scratch.data_ptr(), scratch.numel()
    if last_next_obs is not None:
        io.compact, io.last_next_obs_dev = 1, last_next_obs.data_ptr()
    st = stream if stream is not None else torch.cuda.current_stream(out.device)
    _lib.check(lib.cmbpo_pack_records(C.byref(io), C.c_void_p(st.cuda_stream)))


class PoolGatherer:
    """Per-rollout all-gather of the transitions a ``ModelRollout`` generated into the replicated model
    pool -- the only collective on the path.

    ``launch()`` (call it right after a rollout was enqueued) packs that rollout's populated transitions
    into a staging buffer on the compute stream (``cmbpo_pack_records``, no host sync) and then runs the
    NCCL collective of the PREVIOUS rollout on a side stream, so the transfer overlaps the rollout now
    executing.  Shards have different transition counts ('uncertainty' cutoffs, terminals): counts are
    all-gathered first (8 bytes per rank), then the payload goes straight into a dense
    ``[sum(counts), W]`` matrix -- ``all_gather_into_tensor`` when the counts are equal, else the
    uneven-size ``all_gather`` (one broadcast per rank inside one NCCL group).  ``wait()`` drains the
    pipeline and returns ``(records, counts)`` of the last rollout; ``records[:, -2:]`` viewed as
    int32 are the (global path, t) coordinates."""

    def __init__(self, rollout, world, group=None, capacity=None, pack_only=False):
        self.ro, self.world, self.group = rollout, world, group
        self.pack_only = pack_only   # measurement aid: skip the collective
        self.capacity = capacity
        self.last = None
        self._pending = None
        self._k = 0
        self._buf = None

    def _allocate(self):
        ro = self.ro
        dev = ro.device
        B, T = ro.pool["populated"].shape
        self.E = ro.env.num_networks if ro.mode == "iv_gae" else 1
        self.W = record_width(ro.env.obs_dim, ro.env.act_dim, self.E)
        cap = int(self.capacity) if self.capacity else B * max(T - 1, 1)   # a path stores at most T-1 (Q8)
        from . import _lib
        nscr = int(_lib.load().cmbpo_pack_scratch_bytes(B)) + 64
        self._buf = [dict(rec=torch.empty(cap, self.W, dtype=torch.float32, device=dev),
                          count=torch.zeros(1, dtype=torch.int64, device=dev),
                          count_host=torch.zeros(1, dtype=torch.int64).pin_memory(),
                          scratch=torch.empty(nscr, dtype=torch.uint8, device=dev),
                          packed=torch.cuda.Event(), sent=None) for _ in range(2)]
        self._side = torch.cuda.Stream(device=dev)

    def launch(self):
        if self._buf is None:
            self._allocate()
        ro = self.ro
        b = self._buf[self._k & 1]
        main = torch.cuda.current_stream(ro.device)
        if b["sent"] is not None:
            main.wait_event(b["sent"])          # the collective two rollouts ago has read this staging buffer
        pack_records_cuda(ro.pool, ro.row_offset, ro.env.obs_dim, ro.env.act_dim, self.E, b["rec"], b["count"],
                          b["scratch"], main)
        b["count_host"].copy_(b["count"], non_blocking=True)
        b["packed"].record(main)
        prev, self._pending = self._pending, b
        self._k += 1
        if prev is not None:
            self._collect(prev)

    def _collect(self, b):
        b["packed"].synchronize()
        n = int(b["count_host"][0])
        if n > b["rec"].shape[0]:
            raise RuntimeError(f"PoolGatherer: {n} transitions exceed the staging capacity {b['rec'].shape[0]}")
        if self.pack_only:
            self.last = (b["rec"][:n], [n])
            return
        with torch.cuda.stream(self._side):
            counts_dev = torch.empty(self.world, dtype=torch.int64, device=b["rec"].device)
            dist.all_gather_into_tensor(counts_dev, b["count"], group=self.group)
            counts = [int(c) for c in counts_dev.cpu()]
            total = sum(counts)
            full = torch.empty(total, self.W, dtype=torch.float32, device=b["rec"].device)
            if total > 0:
                if min(counts) == max(counts):
                    dist.all_gather_into_tensor(full, b["rec"][:n], group=self.group)
                else:
                    outs, o = [], 0
                    for c in counts:
                        outs.append(full[o:o + c])
                        o += c
                    dist.all_gather(outs, b["rec"][:n], group=self.group)
            ev = torch.cuda.Event()
            ev.record(self._side)
            b["sent"] = ev
        full.record_stream(self._side)
        self.last = (full, counts)

    def wait(self):
        if self._pending is not None:
            b, self._pending = self._pending, None
            self._collect(b)
        if self._buf is not None:
            self._side.synchronize()
        return self.last


class P2PPoolGatherer(PoolGatherer):
    """Same contract as ``PoolGatherer`` with the payload moved by the COPY ENGINES over NVLink peer
    mappings instead of NCCL kernels: the staging buffers are symmetric memory
    (``torch.distributed._symmetric_memory``), every rank pulls each peer's packed records straight into
    its dense ``[sum(counts), W]`` matrix with ``cudaMemcpyAsync`` on a side stream.  No SM is taken from
    the persistent forward kernel while the transfer overlaps the next rollout (an NCCL all-gather kernel
    holds 8-32 SMs for the whole transfer, and a forward cluster that cannot become resident delays its
    statically assigned tiles).  Cross-rank ordering: a symmetric-memory barrier after the pack ("records
    visible") and one after the pulls ("staging reusable")."""

    def _allocate(self):
        import torch.distributed._symmetric_memory as symm
        ro = self.ro
        dev = ro.device
        B, T = ro.pool["populated"].shape
        self.E = ro.env.num_networks if ro.mode == "iv_gae" else 1
        self.W = record_width(ro.env.obs_dim, ro.env.act_dim, self.E)
        cap = int(self.capacity) if self.capacity else B * max(T - 1, 1)
        group = self.group if self.group is not None else dist.group.WORLD
        self.rank = dist.get_rank(group)
        capt = torch.tensor([cap], dtype=torch.int64, device=dev)   # symmetric buffers: same size on every rank
        dist.all_reduce(capt, op=dist.ReduceOp.MAX, group=group)
        cap = self.cap = int(capt.item())
        from . import _lib
        nscr = int(_lib.load().cmbpo_pack_scratch_bytes(B)) + 64
        stage = symm.empty(2 * cap * self.W, dtype=torch.float32, device=dev)
        cnt = symm.empty(16, dtype=torch.int64, device=dev)
        cnt.zero_()
        self._stage_h = symm.rendezvous(stage, group)
        self._cnt_h = symm.rendezvous(cnt, group)
        self._stage, self._cnt = stage, cnt
        self._peer_stage = [[self._stage_h.get_buffer(r, (cap, self.W), torch.float32, s * cap * self.W)
                             for r in range(self.world)] for s in range(2)]
        self._peer_cnt = [self._cnt_h.get_buffer(r, (16,), torch.int64, 0) for r in range(self.world)]
        self._buf = [dict(rec=self._peer_stage[s][self.rank], count=cnt[s:s + 1], slot=s,
                          count_host=torch.zeros(1, dtype=torch.int64).pin_memory(),
                          scratch=torch.empty(nscr, dtype=torch.uint8, device=dev),
                          packed=torch.cuda.Event(), sent=None) for s in range(2)]
        self._side = torch.cuda.Stream(device=dev)

    def _collect(self, b):
        s = b["slot"]
        with torch.cuda.stream(self._side):
            self._side.wait_event(b["packed"])
            self._stage_h.barrier(channel=0)                     # every rank's records of this rollout are visible
            cnts = torch.stack([self._peer_cnt[r][s] for r in range(self.world)])
            counts = [int(c) for c in cnts.cpu()]
            if max(counts) > self.cap:
                raise RuntimeError(f"PoolGatherer: {max(counts)} transitions exceed the staging capacity {self.cap}")
            offs = [0]
            for c in counts:
                offs.append(offs[-1] + c)
            full = torch.empty(offs[-1], self.W, dtype=torch.float32, device=b["rec"].device)
            if not self.pack_only:
                for i in range(self.world):                      # own shard first, then peers, staggered by rank
                    r = (self.rank + i) % self.world
                    if counts[r]:
                        full[offs[r]:offs[r + 1]].copy_(self._peer_stage[s][r][:counts[r]], non_blocking=True)
            self._stage_h.barrier(channel=1)                     # every rank has pulled: staging slot reusable
            ev = torch.cuda.Event()
            ev.record(self._side)
            b["sent"] = ev
        full.record_stream(self._side)
        self.last = (full, counts)


class HostDrain(PoolGatherer):
    """The same pipeline with the HOST as destination (single rank, no collective): rollout k's packed
    records go to pinned host memory with one ``cudaMemcpyAsync`` on a side stream while rollout k+1
    runs.  ``wait()`` returns ``(host_records[:n], [n])``.  This is what ``ModelBuffer.get`` hands to the
    policy update when the consumer lives on the host (modelbuffer.py:348-405).
    ``compact=True`` (default): records without the next_obs block + ``last_next_obs`` [B, E*obs] per rollout
    (``self.last_next`` after ``wait()``; ``unpack_gathered(..., last_next_obs=...)`` rebuilds next_obs) -- 40 % fewer bytes
    over PCIe, which is what bounds the end-to-end rate of a 1 M-row rollout."""

    def __init__(self, rollout, capacity=None, compact=True):
        super().__init__(rollout, 1, None, capacity)
        self.compact = bool(compact)
        self._host = [None, None]
        self._host_last = [None, None]
        self.last_next = None

    def _allocate(self):
        super()._allocate()
        if self.compact:
            ro = self.ro
            B = ro.pool["populated"].shape[0]
            self.W = record_width(ro.env.obs_dim, ro.env.act_dim, self.E, compact=True)
            for b in self._buf:
                b["rec"] = b["rec"].view(-1)[: b["rec"].shape[0] * self.W].view(b["rec"].shape[0], self.W)
                b["last"] = torch.empty(B, self.E * ro.env.obs_dim, dtype=torch.float32, device=ro.device)

    def launch(self):
        if self._buf is None:
            self._allocate()
        if not self.compact:
            return super().launch()
        ro = self.ro
        b = self._buf[self._k & 1]
        main = torch.cuda.current_stream(ro.device)
        if b["sent"] is not None:
            main.wait_event(b["sent"])
        pack_records_cuda(ro.pool, ro.row_offset, ro.env.obs_dim, ro.env.act_dim, self.E, b["rec"], b["count"],
                          b["scratch"], main, last_next_obs=b["last"])
        b["count_host"].copy_(b["count"], non_blocking=True)
        b["packed"].record(main)
        prev, self._pending = self._pending, b
        self._k += 1
        if prev is not None:
            self._collect(prev)

    def _collect(self, b):
        b["packed"].synchronize()
        n = int(b["count_host"][0])
        if n > b["rec"].shape[0]:
            raise RuntimeError(f"HostDrain: {n} transitions exceed the staging capacity {b['rec'].shape[0]}")
        s = 0 if b is self._buf[0] else 1
        h = self._host[s]
        if h is None or h.shape[0] < n:
            h = self._host[s] = torch.empty(max(n, 1) + max(n, 1) // 8, self.W, dtype=torch.float32, pin_memory=True)
        if self.compact and self._host_last[s] is None:
            self._host_last[s] = torch.empty(tuple(b["last"].shape), dtype=torch.float32, pin_memory=True)
        with torch.cuda.stream(self._side):
            self._side.wait_event(b["packed"])
            h[:n].copy_(b["rec"][:n], non_blocking=True)
            if self.compact:
                self._host_last[s].copy_(b["last"], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._side)
            b["sent"] = ev
        self.last = (h[:n], [n])
        self.last_next = self._host_last[s] if self.compact else None

    def bytes_per_rollout(self, n):
        """device -> host bytes of a rollout with n transitions"""
        extra = self._buf[0]["last"].numel() * 4 if self.compact else 0
        return int(n) * self.W * 4 + extra


def reconstruct_next_obs(obs, path, last_next_obs, row_offset=0):
    """next_obs of compact records in [path, t] order: inside a path next_obs[t] is obs[t+1] (the sampler feeds
    next_obs back as the current observation, model_sampler.py:463-467); the last record of a path takes
    ``last_next_obs[path - row_offset]``.  ``obs`` [n, (E,) obs], ``path`` int [n] (global ids), torch or NumPy."""
    if isinstance(obs, torch.Tensor):
        n = obs.shape[0]
        nxt = torch.empty_like(obs)
        if n == 0:
            return nxt
        same = torch.zeros(n, dtype=torch.bool, device=obs.device)
        same[:-1] = path[1:] == path[:-1]
        nxt[:-1] = obs[1:]
        ends = ~same
        nxt[ends] = last_next_obs.reshape((last_next_obs.shape[0],) + tuple(obs.shape[1:]))[(path[ends] - row_offset).long()]
        return nxt
    import numpy as np
    n = obs.shape[0]
    nxt = np.empty_like(obs)
    if n == 0:
        return nxt
    same = np.zeros(n, bool)
    same[:-1] = path[1:] == path[:-1]
    nxt[:-1] = obs[1:]
    nxt[~same] = last_next_obs.reshape((last_next_obs.shape[0],) + obs.shape[1:])[path[~same] - row_offset]
    return nxt


def unpack_gathered(rec: torch.Tensor, obs_dim: int, act_dim: int, n_members: int = 1, last_next_obs=None, row_offset=0):
    """Split the gathered record matrix back into named arrays (+ int64 ``coords`` [n, 2]).  Compact records
    (``last_next_obs`` given): next_obs is rebuilt by ``reconstruct_next_obs``."""
    E = max(int(n_members), 1)
    compact = last_next_obs is not None
    o, out = 0, {}
    fields = [("obs", E * obs_dim), ("act", E * act_dim)] + ([] if compact else [("next_obs", E * obs_dim)]) + \
             [("rew", E), ("cost", E), ("dyn_error", 1), ("term", E)]
    for k, wdt in fields:
        v = rec[:, o:o + wdt]
        if E > 1 and k != "dyn_error":
            v = v.reshape(v.shape[0], E, -1)
            v = v[..., 0] if v.shape[-1] == 1 else v
        elif wdt == 1:
            v = v[:, 0]
        out[k] = v
        o += wdt
    out["term"] = out["term"] != 0
    out["coords"] = rec[:, o:o + 2].contiguous().view(torch.int32).to(torch.int64)
    if compact:
        out["next_obs"] = reconstruct_next_obs(out["obs"], out["coords"][:, 0], last_next_obs, row_offset)
    return out


def symmetric_pool_alloc(group=None):
    """``pool_alloc`` for ``ModelRollout``: the pool arena in symmetric memory (``torch.distributed._symmetric_memory``)
    so that every rank of ``group`` can read it through its NVLink peer mapping.  Returns the allocator; the arenas it
    handed out are rendezvoused by ``DensePoolGatherer``."""
    import torch.distributed._symmetric_memory as symm

    def alloc(nbytes):
        dev = torch.device("cuda", torch.cuda.current_device())
        n = torch.tensor([int(nbytes)], dtype=torch.int64, device=dev)      # symmetric buffers: same size on every rank
        dist.all_reduce(n, op=dist.ReduceOp.MAX, group=group)
        return symm.empty(int(n.item()), dtype=torch.uint8, device=dev)
    return alloc


class DensePoolGatherer:
    """All-gather of the model pool ITSELF, for rollouts whose pools are dense ('schedule' mode: every
    path stores every step): each rank's pool arena -- all fields, time-major, one contiguous block
    (``ModelRollout.pool_arena``) -- is pulled by every other rank with ONE copy-engine transfer per peer over the NVLink
    peer mapping.  No pack kernel, no count exchange, no host synchronisation, no SM taken from the rollout kernels.

    ``rollouts``: two ``ModelRollout`` objects (same shape, ``pool_alloc=symmetric_pool_alloc(group)``) used alternately,
    so the pulls of rollout k overlap rollout k+1, which writes the other arena.  ``launch(i)`` right after rollout i was
    enqueued; ``wait()`` drains.  ``pools(i)`` -> ``{field: [one logical [B_shard, T, ...] view per rank]}`` of the gathered
    pool of buffer i (modelbuffer.py:45-114): their concatenation over the ranks is the single-device pool when the shards
    are contiguous row blocks.  Only the peers' arenas are copied (the own shard is already resident), and only the first
    T-1 time blocks of each field (column T-1 is never populated, quirk Q8).  ``pack_only`` (measurement aid): no copies."""

    def __init__(self, rollouts, world, group=None, pack_only=False):
        import torch.distributed._symmetric_memory as symm
        assert len(rollouts) == 2
        self.ros, self.world, self.pack_only = rollouts, world, pack_only
        group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(group)
        ro = rollouts[0]
        dev = ro.device
        self.nbytes = ro.pool_arena.numel()
        assert rollouts[1].pool_arena.numel() == self.nbytes and ro.pool_layout == rollouts[1].pool_layout
        self._h = [symm.rendezvous(r.pool_arena, group) for r in rollouts]
        self._peer = [[h.get_buffer(r, (self.nbytes,), torch.uint8, 0) for r in range(world)] for h in self._h]
        self.full = [torch.zeros(world, self.nbytes, dtype=torch.uint8, device=dev) for _ in range(2)]   # column T-1 stays zero
        self._side = torch.cuda.Stream(device=dev)
        self._done = [None, None]      # rollout i's arena may be overwritten again after this event
        self._ready = [torch.cuda.Event(), torch.cuda.Event()]
        self._launched = [False, False]

    def before_rollout(self, i):
        """Call before enqueueing rollout i: its arena must not be overwritten while peers still pull the previous one."""
        if self._done[i] is not None:
            torch.cuda.current_stream(self.ros[i].device).wait_event(self._done[i])

    def _segments(self, ro):
        """Byte ranges of the arena worth moving: a path stores at most T-1 transitions (model_sampler.py:437, quirk Q8),
        so column T-1 of every time-major [T][B][.] field is never populated -- only the first T-1 time blocks travel
        (fields with a leading member axis are moved whole)."""
        segs = []
        for name, (o, shape, dt) in ro.pool_layout.items():
            nb = int(np.prod(shape)) * (4 if dt == torch.float32 else 1)
            lead = len(shape) - 2 - (1 if name in ("obs", "act", "next_obs") else 0)     # axes before T
            if lead == 0 and shape[0] > 1:
                nb = nb // shape[0] * (shape[0] - 1)
            segs.append((o, nb))
        return segs

    def launch(self, i):
        main = torch.cuda.current_stream(self.ros[i].device)
        self._ready[i].record(main)
        segs = self._segments(self.ros[i])
        with torch.cuda.stream(self._side):
            self._side.wait_event(self._ready[i])
            self._h[i].barrier(channel=0)                 # every rank's rollout i is complete
            if not self.pack_only:
                for k in range(1, self.world):            # the peers, staggered by rank; the own shard stays where it is
                    r = (self.rank + k) % self.world
                    for o, nb in segs:
                        self.full[i][r, o:o + nb].copy_(self._peer[i][r][o:o + nb], non_blocking=True)
            self._h[i].barrier(channel=1)                 # every rank has pulled: the arenas may be reused
            ev = torch.cuda.Event()
            ev.record(self._side)
            self._done[i] = ev
        self._launched[i] = True

    def wait(self):
        self._side.synchronize()

    def pools(self, i):
        """``{field: [per-rank logical pool views]}``: entry r is rank r's ``[B_shard, T, ...]`` pool (the own rank's entry
        is the rollout's own arena -- it is not copied); valid until rollout i runs again."""
        ro = self.ros[i]
        out = {}
        for name, (o, shape, dt) in ro.pool_layout.items():
            nb = int(np.prod(shape)) * (4 if dt == torch.float32 else 1)
            m = len(shape) - 2 - (1 if name in ("obs", "act", "next_obs") else 0)
            views = []
            for r in range(self.world):
                src = ro.pool_arena if r == self.rank else self.full[i][r]
                views.append(src[o:o + nb].view(dt).view(tuple(shape)).transpose(m, m + 1))
            out[name] = views
        return out
