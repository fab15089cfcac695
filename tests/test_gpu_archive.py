"""cmbpo_sample_archive (start-state sampling) against the reference's real CPOBuffer.boltz_dist /
distributed_batch_from_archive (tests/golden/archive_sampling.npz) and, for the Philox draws, against the distribution."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _archive():
    import constrainedmbpo_b200.archive as ar
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "archive_sampling.npz"))
    a = ar.DeviceArchive({"observations": g["obs"], "mu": g["mu"], "log_std": g["log_std"]}, g["epoch"], seed=11)
    return g, a


def test_weighted_draw_and_gather_match_reference_golden():
    g, a = _archive()
    dist = a.boltz_dist(g["kls"], float(g["alpha"]))
    assert np.array_equal(dist, g["dist"])
    for kw in (dict(dist=dist), dict(epoch_table=a.boltz_table(g["kls"], float(g["alpha"])))):
        s = a.distributed_batch_from_archive(len(g["uniforms"]), uniforms=g["uniforms"], **kw)
        assert np.array_equal(s["index"].cpu().numpy(), g["index"])          # integer work: exact
        assert np.array_equal(s["observations"].cpu().numpy(), g["s_obs"])
        assert np.array_equal(s["mu"].cpu().numpy(), g["s_mu"]) and np.array_equal(s["log_std"].cpu().numpy(), g["s_log_std"])


def test_philox_draws_follow_the_distribution_and_are_shard_invariant():
    g, a = _archive()
    table = a.boltz_table(g["kls"], float(g["alpha"]))
    n = 400_000
    s = a.distributed_batch_from_archive(n, epoch_table=table, fields=["observations"])
    idx = s["index"].cpu().numpy()
    ep = g["epoch"][idx]
    assert (ep >= 0).all()
    want = np.exp(-float(g["alpha"]) * g["kls"]); want /= want.sum()
    got = np.bincount(ep, minlength=len(want)) / n
    assert np.abs(got - want).max() < 4 * np.sqrt(want.max() / n) + 1e-3
    assert np.array_equal(s["observations"].cpu().numpy(), g["obs"][idx])
    # the same global sample ids give the same draws whatever the split
    a._calls -= 1
    lo = a.distributed_batch_from_archive(1000, epoch_table=table, fields=["observations"], sample_offset=5000)
    assert np.array_equal(lo["index"].cpu().numpy(), idx[5000:6000])
    # uniform over rows within an epoch: row frequencies of one epoch are flat
    rows = idx[ep == 2]
    cnt = np.bincount(rows - rows.min())
    assert cnt.std() < 3 * np.sqrt(cnt.mean())
