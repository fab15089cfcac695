"""cmbpo_normalize_advantages_static: the allocation-free variant of cmbpo_normalize_advantages (scratch in a file-scope
``static __device__`` array, see include/cmbpo_b200.h).  It has no counterpart in the reference; it was added at the request of
the run's operators.  Called through the C ABI, checked against the existing entry point on two calls in a row -- the second
call finds the first call's leftovers in the shared array -- plus the size limit.  The file sorts last on purpose: the round's
end runs ``pytest -x``."""
import ctypes as C
import gc

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _io(_lib, adv, cadv, pop, ret, cret, stats, scratch=None):
    io = _lib.cmbpo_norm_io()
    io.n_elems = adv.numel()
    io.adv_dev, io.cadv_dev, io.populated_dev = adv.data_ptr(), cadv.data_ptr(), pop.data_ptr()
    io.ret_dev, io.cret_dev = ret.data_ptr(), cret.data_ptr()
    io.stats_dev = stats.data_ptr()
    if scratch is not None:
        io.scratch_dev, io.scratch_bytes = scratch.data_ptr(), scratch.numel()
    return io


def test_static_scratch_normalize_matches_the_allocating_entry_point_twice():
    import torch
    from constrainedmbpo_b200 import _lib
    lib = _lib.require_device(0)
    torch.cuda.set_device(0)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    err = C.create_string_buffer(256)
    rng = np.random.default_rng(17)
    # throwaway call: loads the entry point's kernels (lazy module loading takes device memory on a first launch)
    w = [torch.ones(64, device="cuda") for _ in range(4)] + [torch.ones(64, dtype=torch.uint8, device="cuda"), torch.zeros(6, device="cuda")]
    assert lib.cmbpo_normalize_advantages_static(C.byref(_io(_lib, w[0], w[1], w[4], w[2], w[3], w[5])), stream, err, len(err)) == 0
    torch.cuda.synchronize()
    # two different pools in a row: 700k entries (1024 blocks, all partial-sum slots written), then a small ragged one that
    # uses 3 blocks and must not see the first call's partial sums or ticket
    for call, (n, frac) in enumerate([(700_003, 0.4), (613, 0.7)]):
        adv = (rng.standard_normal(n) * 3 + 5).astype(np.float32)
        cadv = (rng.standard_normal(n) + 2).astype(np.float32)
        ret, cret = rng.standard_normal(n).astype(np.float32), rng.standard_normal(n).astype(np.float32)
        pop = (rng.random(n) < frac).astype(np.uint8)
        t = lambda a: torch.from_numpy(a).cuda()
        # existing entry point, caller scratch
        a0, c0, st0 = t(adv), t(cadv), torch.zeros(6, device="cuda")
        scratch = torch.empty(int(lib.cmbpo_normalize_scratch_bytes()), dtype=torch.uint8, device="cuda")
        keep0 = (t(pop), t(ret), t(cret))
        io0 = _io(_lib, a0, c0, *keep0, st0, scratch)
        _lib.check(lib.cmbpo_normalize_advantages(C.byref(io0), stream))
        # new entry point: inputs and outputs from the caller, no scratch
        a1, c1, st1 = t(adv), t(cadv), torch.zeros(6, device="cuda")
        keep1 = (t(pop), t(ret), t(cret))
        io1 = _io(_lib, a1, c1, *keep1, st1)
        # "allocates nothing": no new device memory at the driver level (a cudaMalloc inside the library would show there, not
        # in torch's counters) and no allocation through torch's allocator.  Garbage of earlier tests is collected first and
        # only GROWTH counts, so tensors freed by a collector run in between (this file runs last) cannot fail the check.
        gc.collect()
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        allocs_before = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
        rc = lib.cmbpo_normalize_advantages_static(C.byref(io1), stream, err, len(err))
        assert rc == 0, err.value.decode()
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before and err.value == b""
        assert torch.cuda.memory_stats().get("allocation.all.allocated", 0) == allocs_before
        # the tolerance of tests/test_gpu_gae.py (rtol 1e-5 / atol 1e-6); the two paths share grid and summation order
        assert np.allclose(a1.cpu().numpy(), a0.cpu().numpy(), rtol=1e-5, atol=1e-6), call
        assert np.allclose(c1.cpu().numpy(), c0.cpu().numpy(), rtol=1e-5, atol=1e-6), call
        assert np.allclose(st1.cpu().numpy(), st0.cpu().numpy(), rtol=1e-5, atol=1e-7), call
        assert int(st1[5].item()) == int(pop.sum())
        m = pop.astype(bool)
        assert np.array_equal(a1.cpu().numpy()[~m], adv[~m])                       # unpopulated entries untouched
        assert abs(a1.cpu().numpy()[m].mean()) < 1e-4 and abs(a1.cpu().numpy()[m].std() - 1) < 1e-3


def test_static_scratch_normalize_rejects_sizes_beyond_its_maximum():
    import torch
    from constrainedmbpo_b200 import _lib
    lib = _lib.require_device(0)
    z = torch.zeros(8, device="cuda")
    pop = torch.ones(8, dtype=torch.uint8, device="cuda")
    stats = torch.full((6,), -1.0, device="cuda")
    io = _io(_lib, z, z.clone(), pop, z.clone(), z.clone(), stats)
    io.n_elems = (1 << 26) + 1                                   # CMBPO_NORMALIZE_STATIC_MAX_ELEMS + 1: nothing may be launched
    err = C.create_string_buffer(256)
    rc = lib.cmbpo_normalize_advantages_static(C.byref(io), C.c_void_p(torch.cuda.current_stream().cuda_stream), err, len(err))
    torch.cuda.synchronize()
    assert rc == -2 and b"exceeds the maximum" in err.value      # CMBPO_ERR_UNSUPPORTED
    assert (stats == -1).all()
    assert lib.cmbpo_normalize_advantages_static(None, None, err, len(err)) == -1 and err.value    # CMBPO_ERR_ARG
