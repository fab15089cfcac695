"""Start-state sampling for the model rollouts on a device-resident archive:
``CPOBuffer.boltz_dist`` + ``CPOBuffer.distributed_batch_from_archive`` (softlearning/replay_pools/cpobuffer.py:
531-542, :559-610), as ``CMBPO._train`` uses them (mbpo/algorithms/cmbpo.py:328-333).  A thin ctypes client of
``cmbpo_sample_archive``; the drawn start states stay on the device and feed ``ModelRollout.reset`` directly."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib


class DeviceArchive:
    """The part of the replay archive the rollout path reads: ``fields`` (name -> ``[N, w]`` float array, e.g.
    'observations' and the pi_info entries 'mu' / 'log_std') and ``epoch_archive`` (``[N]`` ints, < 0 = empty slot)."""

    def __init__(self, fields, epoch_archive, device="cuda:0", seed=None):
        self.device = torch.device(device)
        self.fields = {k: torch.as_tensor(np.asarray(v, dtype=np.float32)).reshape(len(v), -1).to(self.device).contiguous()
                       for k, v in fields.items()}
        if len(self.fields) > _lib.MAX_FIELDS:
            raise ValueError(f"at most {_lib.MAX_FIELDS} fields per draw")
        self.epoch_host = np.asarray(epoch_archive).astype(np.int64)
        self.archive_size = len(self.epoch_host)
        self.epoch = torch.as_tensor(self.epoch_host.astype(np.int32)).to(self.device)
        self.cdf = torch.empty(self.archive_size, dtype=torch.float64, device=self.device)
        self._seed = int(seed) if seed is not None else int(np.random.randint(2 ** 31 - 1))
        self._calls = 0
        self._lib = _lib.load()

    def boltz_table(self, kls, alpha=1):
        """Per-epoch row probability ``sample_p`` of ``boltz_dist`` (cpobuffer.py:535-538): float32, indexed by epoch."""
        ep_probs = np.exp(alpha * np.negative(kls))
        ep_probs /= np.sum(ep_probs)
        sample_p = np.bincount(self.epoch_host[self.epoch_host >= 0]).astype(np.float32)
        sample_p[sample_p > 0] = ep_probs / sample_p[sample_p > 0]
        return sample_p

    def boltz_dist(self, kls, alpha=1):
        """The full per-row distribution, as the reference returns it (cpobuffer.py:540-541); NumPy, for inspection."""
        sample_p = self.boltz_table(kls, alpha)
        return np.where(self.epoch_host >= 0, sample_p[self.epoch_host], 0)

    def distributed_batch_from_archive(self, batch_size, dist=None, epoch_table=None, fields=None, uniforms=None,
                                       sample_offset=0):
        """``np.random.choice(archive_size, batch_size, p=dist)`` + gather (cpobuffer.py:600-609) on the device.
        Give either ``dist`` (``[N]`` row probabilities, any of boltz / disc / unif) or ``epoch_table`` (the
        ``boltz_table``; the row distribution is then formed in the kernel).  ``uniforms`` injects the uniform draws
        (NumPy's ``random_sample``), None = Philox.  Returns a dict of CUDA tensors incl. 'index'."""
        names = list(self.fields) if fields is None else [f for f in fields if f != "pi_infos"]
        io = _lib.cmbpo_sample_io()
        io.archive_size, io.batch_size = self.archive_size, int(batch_size)
        keep = []
        if dist is not None:
            p = torch.as_tensor(np.asarray(dist, dtype=np.float32)).to(self.device).contiguous()
            keep.append(p)
            io.prob_dev = p.data_ptr()
        elif epoch_table is not None:
            t = torch.as_tensor(np.asarray(epoch_table, dtype=np.float32)).to(self.device).contiguous()
            keep.append(t)
            io.epoch_dev, io.epoch_prob_dev, io.n_epochs = self.epoch.data_ptr(), t.data_ptr(), t.numel()
        else:
            raise ValueError("give dist or epoch_table")
        io.rebuild_cdf, io.cdf_dev = 1, self.cdf.data_ptr()
        if uniforms is not None:
            u = torch.as_tensor(np.asarray(uniforms, dtype=np.float64)).to(self.device).contiguous()
            keep.append(u)
            io.uniform_dev = u.data_ptr()
        io.seed, io.rng_offset, io.sample_offset = self._seed, self._calls, int(sample_offset)
        self._calls += 1
        out = {}
        for i, k in enumerate(names):
            src = self.fields[k]
            out[k] = torch.empty(batch_size, src.shape[1], dtype=torch.float32, device=self.device)
            io.field_dev[i], io.field_width[i], io.out_dev[i] = src.data_ptr(), src.shape[1], out[k].data_ptr()
        out["index"] = torch.empty(batch_size, dtype=torch.int64, device=self.device)
        io.index_dev = out["index"].data_ptr()
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(self._lib.cmbpo_sample_archive(C.byref(io), stream))
        del keep
        return out
