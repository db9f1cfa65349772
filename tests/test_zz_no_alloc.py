"""dls_eval_wave_noalloc: the test wave for callers that cannot allocate device memory.

Who that is: anything that reaches the library through a thin foreign-function stub (the stub
forwards the model rows, the test images and two small result arrays, and has no device allocator
to size a workspace with), or a thread that wants an accuracy reading while the solver thread owns
the training workspace, or a caller that cannot afford cudaMalloc's implicit synchronisation in
the middle of a round.  For them the library keeps 32 MiB of scratch inside its own module image.

This file is named so that pytest collects it last.  It drives the entry point through the C-ABI
(ctypes, raw device pointers) and compares it with dls_eval_wave on the same inputs.  Both run
the same forward kernels in the same order, so the hit counts must be equal and the fp64 loss sums
equal to 1e-12 relative (tolerance of the other eval tests: 1 hit, 1e-4).  Two calls in a row with
DIFFERENT models and test-set sizes check that nothing left in the shared scratch by the first
call leaks into the second.  All result buffers are allocated before the calls; the entry point
itself must not allocate, which is checked with torch's allocator statistics.
"""
import ctypes

import numpy as np
import pytest
import torch

import dls_b200
from dls_b200 import _lib, kernels as K, synthetic
from dls_b200.layout import GNLENET as L
from dls_b200.models import initial_row

pytestmark = pytest.mark.gpu


def _models(n, seed, dev):
    base = initial_row(L, 42)
    g = torch.Generator().manual_seed(seed)
    X = base.repeat(n, 1)
    X[:, :L.P] += 0.02 * torch.randn((n, L.P), generator=g)
    return X.to(dev)


def _call_noalloc(X, rows, timg, tlab, correct, loss):
    return _lib.lib().dls_eval_wave_noalloc(
        L.arch_id, X.data_ptr(), X.stride(0), rows.data_ptr(), rows.numel(), timg.data_ptr(), tlab.data_ptr(),
        tlab.numel(), correct.data_ptr(), loss.data_ptr(), torch.cuda.current_stream().cuda_stream)


def test_noalloc_matches_eval_wave_twice_in_a_row(cuda):
    ws = K.Workspace(cuda)
    cases = [(5, 300, 1), (16, 1000, 2)]          # (models, test images, seed): second call is the maximum width
    for n_models, n_test, seed in cases:
        X = _models(n_models + 2, seed, cuda)
        rows = torch.arange(n_models, dtype=torch.int32, device=cuda).flip(0).contiguous() + 1
        ti, tl = synthetic.make_test_split(n_test, 32, 3, 10, seed=1234 + seed)
        timg, tlab = torch.from_numpy(ti.reshape(n_test, -1)).to(cuda), torch.from_numpy(tl).to(cuda)
        want_correct, want_loss = K.eval_wave(L, X, rows, timg, tlab, ws)
        correct = torch.full((n_models,), -7, dtype=torch.int32, device=cuda)
        loss = torch.full((n_models,), float("nan"), dtype=torch.float64, device=cuda)
        torch.cuda.synchronize()
        allocs_before = torch.cuda.memory_stats(cuda)["allocation.all.allocated"]
        rc = _call_noalloc(X, rows, timg, tlab, correct, loss)
        assert rc == 0, _lib.ERRORS.get(rc, rc)
        torch.cuda.synchronize()
        assert torch.cuda.memory_stats(cuda)["allocation.all.allocated"] == allocs_before
        assert torch.equal(correct, want_correct), (correct, want_correct)
        np.testing.assert_allclose(loss.cpu().numpy(), want_loss.cpu().numpy(), rtol=1e-12)
        assert int(want_correct.sum()) > 0 and bool(torch.isfinite(want_loss).all())
    # torch's counter above only sees torch's caching allocator.  A cudaMalloc made inside the library would
    # show up as less free device memory instead, so repeat the (now warm: every kernel is loaded) last call
    # and require that the device's free byte count did not drop.
    correct.fill_(-7)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(cuda)[0]
    rc = _call_noalloc(X, rows, timg, tlab, correct, loss)
    assert rc == 0, _lib.ERRORS.get(rc, rc)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(cuda)[0] >= free_before, "dls_eval_wave_noalloc allocated device memory"
    assert torch.equal(correct, want_correct)


def test_noalloc_rejects_oversize_and_bad_arguments(cuda):
    X = _models(2, 3, cuda)
    ti, tl = synthetic.make_test_split(64, 32, 3, 10)
    timg, tlab = torch.from_numpy(ti.reshape(64, -1)).to(cuda), torch.from_numpy(tl).to(cuda)
    rows17 = torch.zeros(17, dtype=torch.int32, device=cuda)
    correct = torch.zeros(17, dtype=torch.int32, device=cuda)
    loss = torch.zeros(17, dtype=torch.float64, device=cuda)
    assert _call_noalloc(X, rows17, timg, tlab, correct, loss) == -5           # DLS_E_TOO_BIG: > 16 models
    l = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    assert l.dls_eval_wave_noalloc(L.arch_id, X.data_ptr(), X.stride(0), rows17.data_ptr(), 1, timg.data_ptr(),
                                   tlab.data_ptr(), 4097, correct.data_ptr(), loss.data_ptr(), st) == -5
    assert l.dls_eval_wave_noalloc(L.arch_id, None, X.stride(0), rows17.data_ptr(), 1, timg.data_ptr(),
                                   tlab.data_ptr(), 64, correct.data_ptr(), loss.data_ptr(), st) == -1
    assert l.dls_eval_wave_noalloc(1, X.data_ptr(), X.stride(0), rows17.data_ptr(), 1, timg.data_ptr(),
                                   tlab.data_ptr(), 64, correct.data_ptr(), loss.data_ptr(), st) == -3
