"""Aggregate wave (CUDA, through the C-ABI) against the oracle: bit-exact."""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O
import dls_b200
from dls_b200 import kernels as K, synthetic
from dls_b200.layout import GNLENET as L
from katgen import kat_rows

pytestmark = pytest.mark.gpu


def _csr(models_lists, weights_lists=None):
    rowptr = [0]
    col, w = [], []
    for i, ms in enumerate(models_lists):
        ws = weights_lists[i] if weights_lists else [float(1.0 / len(ms))] * len(ms)
        col += list(ms)
        w += [np.float32(x) for x in ws]
        rowptr.append(len(col))
    return (np.array(rowptr, np.int32), np.array(col, np.int32), np.array(w, np.float32))


def _run(cuda, X, rowptr, col, w, P, variant, src_rows=None, out_rows=None, n_out=None):
    Xd = torch.from_numpy(X).to(cuda)
    n_out = n_out or (len(rowptr) - 1)
    Yd = torch.full((n_out, X.shape[1]), float("nan"), device=cuda)
    K.agg_spmm(Xd, Yd, torch.from_numpy(rowptr).to(cuda), torch.from_numpy(col).to(cuda),
               torch.from_numpy(w).to(cuda), P,
               src_rows=None if src_rows is None else torch.from_numpy(src_rows).to(cuda),
               out_rows=None if out_rows is None else torch.from_numpy(out_rows).to(cuda), variant=variant)
    torch.cuda.synchronize()
    return Yd.cpu().numpy()


@pytest.mark.parametrize("variant", [1, 2, 3, 4])
@pytest.mark.parametrize("tag,fan,weighted", [("f5", 5, False), ("f11", 11, False), ("f2w", 2, True), ("f5w", 5, True)])
def test_golden_fedavg(cuda, golden_dir, variant, tag, fan, weighted):
    """Same inputs as the reference-generated fixture: CUDA result == reference bits."""
    if variant == 4 and weighted:
        pytest.skip("variant 4 is for waves with one weight")
    g = np.load(os.path.join(golden_dir, "fedavg_kat.npz"))
    X = np.zeros((fan, L.ld), np.float32)
    X[:, :L.P] = kat_rows(fan, L.P, salt=fan + (100 if weighted else 0))
    ws = [list(g[tag + "_w"])] if weighted else None
    rowptr, col, w = _csr([list(range(fan))], ws)
    Y = _run(cuda, X, rowptr, col, w, L.P, variant)
    assert np.array_equal(Y[0, :L.P].view(np.uint32), g[tag + "_y"].view(np.uint32))


@pytest.mark.parametrize("variant", [1, 2, 3, 4])
@pytest.mark.parametrize("n,k", [(16, 4), (64, 6), (200, 7)])
def test_wave_vs_oracle(cuda, variant, n, k):
    """A whole D-PSGD aggregate wave on a k-regular topology, shuffled arrival order, own model last."""
    if (n * k) % 2:
        k += 1
    nb = synthetic.kreg_topology(n, k, seed=42)
    rs = np.random.RandomState(n)
    lists = []
    for i in range(n):
        l = list(nb[i])
        rs.shuffle(l)
        lists.append(l + [i])
    rowptr, col, w = _csr(lists)
    X = np.zeros((n, L.ld), np.float32)
    X[:, :L.P] = kat_rows(n, L.P, salt=n)
    Y = _run(cuda, X, rowptr, col, w, L.P, variant)
    Yo = O.fedavg_csr(X, rowptr, col, w, L.P)
    assert np.array_equal(Y[:, :L.P].view(np.uint32), Yo[:, :L.P].view(np.uint32))


@pytest.mark.parametrize("variant", [1, 2])
def test_ragged_rows_slots_and_special_values(cuda, variant):
    """Variable fan-in (churn), explicit weights, indirection through src_rows/out_rows, fan-in 1,
    fan-in > 64 (multiple staging passes), -0.0 / inf / nan propagation, P not a multiple of 4."""
    P, ld = 1001, 1024
    n_src = 70
    rs = np.random.RandomState(3)
    pool = np.zeros((100, ld), np.float32)
    pool[:, :P] = rs.standard_normal((100, P)).astype(np.float32)
    pool[7, :8] = [-0.0, np.inf, -np.inf, np.nan, 0.0, 1e-45, -1e38, 3e38]
    pool[9, :8] = [-0.0, 1.0, np.inf, 1.0, -0.0, 1e-45, -3e38, 3e38]
    src_rows = rs.permutation(100)[:n_src].astype(np.int32)
    src_rows[0], src_rows[1] = 7, 9
    lists = [[0], [0, 1], [1, 0, 5], list(range(n_src)), [3, 3, 3], list(rs.permutation(n_src)[:9])]
    weights = [[1.0], [0.5, 0.5], [0.2, 0.3, 0.5], [1.0 / n_src] * n_src, [0.1, 0.2, 0.7],
               list(rs.rand(9).astype(np.float64))]
    rowptr, col, w = _csr(lists, weights)
    out_rows = np.array([5, 0, 3, 1, 4, 2], np.int32)
    Y = _run(cuda, pool, rowptr, col, w, P, variant, src_rows=src_rows, out_rows=out_rows, n_out=6)
    Xs = pool[src_rows]
    Yo = O.fedavg_csr(Xs, rowptr, col, w, P, out_rows=out_rows, n_out=6)
    # NaN payload/sign bits are not defined by IEEE-754 arithmetic (x86 and the GPU pick different
    # canonical NaNs): NaNs must be in the same places, every other value bit-identical (incl. -0.0)
    nan_g, nan_o = np.isnan(Y[:, :P]), np.isnan(Yo[:, :P])
    assert np.array_equal(nan_g, nan_o) and nan_o.sum() >= 4
    assert np.array_equal(Y[:, :P].view(np.uint32)[~nan_o], Yo[:, :P].view(np.uint32)[~nan_o])


def test_empty_wave_and_unaligned_rows(cuda):
    X = torch.zeros((4, 64), device=cuda)
    Y = torch.zeros((4, 64), device=cuda)
    rp = torch.zeros(1, dtype=torch.int32, device=cuda)
    K.agg_spmm(X, Y, rp, rp, torch.zeros(1, device=cuda), 64)      # zero rows: no-op
    # unaligned leading dimension -> scalar path, still bit-exact
    P, ld = 37, 39
    rs = np.random.RandomState(0)
    Xn = rs.standard_normal((6, ld)).astype(np.float32)
    rowptr, col, w = _csr([[1, 2, 0], [5, 4]])
    Yd = _run(cuda, Xn, rowptr, col, w, P, 1)
    Yo = O.fedavg_csr(Xn, rowptr, col, w, P)
    assert np.array_equal(Yd[:2, :P].view(np.uint32), Yo[:, :P].view(np.uint32))


def test_linearity_and_fixed_point_full_size(cuda):
    """Size-independent properties at the full C2 size (256 nodes x GNLeNet): a wave over
    identical rows returns that row (up to the reference's rounding), and a permutation
    matrix reproduces the input bits."""
    n = 256
    Xd = torch.randn((n, L.ld), device=cuda)
    perm = torch.randperm(n, device=cuda).to(torch.int32)
    rowptr = torch.arange(n + 1, dtype=torch.int32, device=cuda)
    Y = torch.empty_like(Xd)
    K.agg_spmm(Xd, Y, rowptr, perm, torch.ones(n, device=cuda), L.P, variant=1)
    assert torch.equal(Y[:, :L.P], Xd[perm.long(), :L.P])
    Y2 = torch.empty_like(Xd)
    K.agg_spmm(Xd, Y2, rowptr, perm, torch.ones(n, device=cuda), L.P, variant=2)
    assert torch.equal(Y2[:, :L.P], Y[:, :L.P])
    # uniform fan-in 2 (node + one neighbour) at full size: variant 3 == variant 1 bit for bit
    col2 = torch.stack([perm, torch.arange(n, dtype=torch.int32, device=cuda)], 1).reshape(-1).contiguous()
    rp2 = torch.arange(n + 1, dtype=torch.int32, device=cuda) * 2
    w2 = torch.full((2 * n,), 0.5, device=cuda)
    Ya, Yb = torch.empty_like(Xd), torch.empty_like(Xd)
    K.agg_spmm(Xd, Ya, rp2, col2, w2, L.P, variant=1)
    K.agg_spmm(Xd, Yb, rp2, col2, w2, L.P, variant=3)
    assert torch.equal(Ya[:, :L.P], Yb[:, :L.P])


def test_copy_and_broadcast_rows(cuda):
    X = torch.randn((8, L.ld), device=cuda)
    Y = torch.zeros((8, L.ld), device=cuda)
    src = torch.tensor([3, 1], dtype=torch.int32, device=cuda)
    dst = torch.tensor([0, 7], dtype=torch.int32, device=cuda)
    K.copy_rows(X, Y, src, dst, L.P)
    assert torch.equal(Y[0, :L.P], X[3, :L.P]) and torch.equal(Y[7, :L.P], X[1, :L.P])
    K.broadcast_row(X[5], Y, torch.arange(8, dtype=torch.int32, device=cuda), L.P)
    assert torch.equal(Y[:, :L.P], X[5, :L.P].expand(8, L.P))


def test_metric_size_wave_all_variants_agree(cuda):
    """The metric's shape — 1024 nodes, 10-regular, fan-in 11, no explicit weights: the 16-bit-index staged kernel
    (variant 4, the one the executor picks there) gives the bits of the gather kernel, which the oracle checks at
    smaller sizes; 8 sampled rows are also compared with the oracle directly."""
    n, k = 1024, 10
    nb = synthetic.kreg_topology(n, k, seed=42)
    lists = [list(nb[i]) + [i] for i in range(n)]
    rowptr, col, w = _csr(lists)
    rs = np.random.RandomState(0)
    X = torch.from_numpy((rs.standard_normal((n, L.ld)) * np.exp2(rs.randint(-6, 6, (n, 1)))).astype(np.float32)).to(cuda)
    outs = {}
    for v in (1, 4):
        assert K.agg_variant_fits(v, n, n, len(col))
        Y = torch.full((n, L.ld), float("nan"), device=cuda)
        K.agg_spmm(X, Y, torch.from_numpy(rowptr).to(cuda), torch.from_numpy(col).to(cuda),
                   torch.from_numpy(w).to(cuda), L.P, n_src_rows=n, variant=v)
        torch.cuda.synchronize()
        outs[v] = Y[:, :L.P]
    assert torch.equal(outs[1].view(torch.int32), outs[4].view(torch.int32))
    Xh = X.cpu().numpy()
    for r in range(0, n, 128):
        want = O.fedavg_numpy([Xh[s, :L.P] for s in lists[r]])
        assert np.array_equal(outs[4][r].cpu().numpy().view(np.uint32), want.view(np.uint32))


def test_push_rows_to_device_addresses(cuda):
    """dls_push_rows (the multi-GPU halo push) with destinations given as raw device addresses — here rows of a
    second local tensor, on several GPUs rows of the peers' symmetric-memory halo.  Ragged tail (P % 4 != 0 is padded
    by the row pitch), many transfers and a single one."""
    X = torch.randn((40, L.ld), device=cuda)
    H = torch.full((64, L.ld), float("nan"), device=cuda)
    for src, dst in (([3], [7]), (list(range(0, 40, 3)), list(range(63, 63 - 14, -1)))):
        s = torch.tensor(src, dtype=torch.int32, device=cuda)
        d = torch.tensor([H.data_ptr() + r * L.ld * 4 for r in dst], dtype=torch.int64, device=cuda)
        K.push_rows(X, s, d, L.P)
        torch.cuda.synchronize()
        for a, b in zip(src, dst):
            assert torch.equal(H[b, :L.P], X[a, :L.P])
    untouched = [r for r in range(64) if r not in (7, *range(63, 63 - 14, -1))]
    assert bool(torch.isnan(H[untouched]).all())


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_push_rows_long_rows_and_column_ranges(cuda, mode):
    """Rows of FEMNIST length take the persistent form of the push kernel (one CTA per SM); the pipelined exchange
    pushes COLUMN RANGES of the rows (a view of X, destinations advanced by the same offset).  All kernel forms
    (dls_tune key 7) must move exactly the requested bytes."""
    from dls_b200 import _lib
    from dls_b200.layout import FEMNIST as F
    lib = _lib.lib()
    X = torch.randn((5, F.ld), device=cuda)
    H = torch.full((6, F.ld), float("nan"), device=cuda)
    src, dst = [4, 0, 2], [1, 5, 3]
    s = torch.tensor(src, dtype=torch.int32, device=cuda)
    d = torch.tensor([H.data_ptr() + r * F.ld * 4 for r in dst], dtype=torch.int64, device=cuda)
    cuts = [0, 425984, 847872, 1269760, F.P]
    lib.dls_tune(7, mode)
    try:
        for c0, c1 in zip(cuts[:-1], cuts[1:]):
            K.push_rows(X[:, c0:c1], s, d + 4 * c0, c1 - c0)
        torch.cuda.synchronize()
    finally:
        lib.dls_tune(7, 0)
    for a, b in zip(src, dst):
        assert torch.equal(H[b, :F.P], X[a, :F.P])
    assert bool(torch.isnan(H[[0, 2, 4]]).all())
