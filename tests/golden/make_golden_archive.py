#!/usr/bin/env python
"""Golden fixture from the reference's OWN ``CPOBuffer.boltz_dist`` and ``distributed_batch_from_archive``
(softlearning/replay_pools/cpobuffer.py:531-542, :559-610), run here on a seeded archive (the buffer object is made
with ``__new__`` and handed the archive arrays; the two methods are the reference's, unmodified).  The uniforms
``np.random.choice`` consumed are recovered by re-seeding the global generator.
  python tests/golden/make_golden_archive.py     # writes tests/golden/archive_sampling.npz"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden_sampler import load_reference  # noqa: E402  (installs the import stubs, loads cpobuffer.py by file)


def main():
    load_reference()
    CPOBuffer = sys.modules["softlearning.replay_pools.cpobuffer"].CPOBuffer
    rng = np.random.default_rng(8)
    N, n_ep, B = 5000, 6, 3000
    epoch = np.repeat(np.arange(n_ep), [900, 700, 1100, 400, 800, 600])      # 4500 filled rows
    epoch = np.concatenate([epoch, -np.ones(N - len(epoch), int)])           # empty slots at the end
    obs = rng.standard_normal((N, 17)).astype(np.float32)
    mu, log_std = rng.standard_normal((N, 6)).astype(np.float32), rng.standard_normal((N, 6)).astype(np.float32)
    buf = CPOBuffer.__new__(CPOBuffer)
    buf.archive_size, buf.epoch_archive = N, epoch
    buf.arch_dict = {"observations": obs, "index": np.arange(N)[:, None]}
    buf.pi_info_archive = {"mu": mu, "log_std": log_std}
    kls = rng.uniform(0, 3, n_ep)
    alpha = 1.7
    dist = buf.boltz_dist(kls, alpha=alpha)
    np.random.seed(5)
    s = buf.distributed_batch_from_archive(B, dist, fields=["observations", "index", "pi_infos"])
    np.random.seed(5)
    u = np.random.random_sample(B)
    path = os.path.join(HERE, "archive_sampling.npz")
    np.savez_compressed(path, epoch=epoch, obs=obs, mu=mu, log_std=log_std, kls=kls, alpha=alpha, dist=dist, uniforms=u,
                        index=s["index"][:, 0], s_obs=s["observations"], s_mu=s["mu"], s_log_std=s["log_std"])
    print("dist sum", dist.sum(), dist.dtype, "->", path, os.path.getsize(path) // 1000, "kB; first indices", s["index"][:5, 0])


if __name__ == "__main__":
    main()
