"""GNLeNet train / grad / test waves (CUDA, through the C-ABI) against the oracle.

Tolerances: precision 0 is fp32 CUDA-core math with a different summation order than ATen, so
gradients agree to ~1e-5 relative (L2 per tensor); per-round parameters are held to the
north-star bound of 1e-3 relative (we assert 3e-4 for fp32: discrete arg-max / ReLU decisions
amplify summation-order noise once the models differ in the last bits).  precision 1 runs the conv2/conv3
GEMMs on tcgen05 with bf16 operands and fp32 accumulation: operand rounding is 2^-9 relative, so
gradients are held to 3e-2 relative per tensor, per-round parameters to the 1e-3 bound."""
import json
import os
import pickle

import numpy as np
import pytest
import torch

from oracle import oracle as O
import dls_b200
from dls_b200 import kernels as K, synthetic
from dls_b200.layout import GNLENET as L

pytestmark = pytest.mark.gpu


def rel(a, b):
    return float(np.linalg.norm(a.astype(np.float64) - b.astype(np.float64)) /
                 max(np.linalg.norm(b.astype(np.float64)), 1e-30))


def perturbed_rows(n, seed=0):
    """Initial model plus per-node noise, with non-trivial GroupNorm scales (some negative)."""
    base = O.initial_row(L, 42)
    rs = np.random.RandomState(seed)
    X = np.tile(base, (n, 1))
    X[:, :L.P] += 0.02 * rs.standard_normal((n, L.P)).astype(np.float32)
    for key in ("features.2.weight", "features.5.weight", "features.9.weight"):
        off = L.offset(key)
        X[:, off:off + 8] *= -1.0
    return X.astype(np.float32)


@pytest.fixture(scope="module")
def data():
    imgs, labs = synthetic.make_partitions(6, 80, 32, 3, 10)
    return imgs, labs


CONV1_WGRAD_TC = True
WGRAD_TC = True      # set when the weight-gradient GEMMs also run on tensor cores
TOL = {0: dict(logit=2e-4, grad_t=2e-4, grad_all=5e-5, par=3e-4, upd=4e-3, loss=1e-4, hits=1),
       1: dict(logit=6e-2, grad_t=1.5e-1, grad_all=1e-1, par=1e-3, upd=1.5e-1, loss=2e-2, hits=6)}


@pytest.mark.parametrize("precision", [0, 1])
def test_eval_logits_accuracy_loss(cuda, data, precision):
    timg, tlab = synthetic.make_test_split(300, 32, 3, 10)     # 300: last group is partial
    X = perturbed_rows(3)
    Xd = torch.from_numpy(X).to(cuda)
    rows = torch.tensor([2, 0, 1], dtype=torch.int32, device=cuda)
    logits = torch.zeros((3, 300, 10), device=cuda)
    ws = K.Workspace(cuda)
    correct, loss = K.eval_wave(L, Xd, rows, torch.from_numpy(timg).to(cuda), torch.from_numpy(tlab).to(cuda),
                                ws, out_logits=logits, precision=precision)
    torch.cuda.synchronize()
    tol = TOL[precision]
    for j, r in enumerate([2, 0, 1]):
        ref = O.logits(L, X[r], timg)
        got = logits[j].cpu().numpy()
        assert np.abs(got - ref).max() < tol["logit"], np.abs(got - ref).max()
        acc, l, ncorrect = O.evaluate(L, X[r], timg, tlab)
        assert abs(int(correct[j]) - ncorrect) <= tol["hits"]     # argmax ties at rounding-noise level
        assert abs(float(loss[j]) / 300 - l) < tol["loss"]


@pytest.mark.parametrize("precision", [0, 1])
@pytest.mark.parametrize("batch,counts", [(32, [32, 32, 16, 1]), (8, [8, 5, 8, 0])])
def test_gradient_vs_autograd(cuda, data, batch, counts, precision):
    imgs, labs = data
    n = len(counts)
    X = perturbed_rows(n, seed=1)
    rs = np.random.RandomState(5)
    idx = -np.ones((n, batch), np.int32)
    for t in range(n):
        idx[t, :counts[t]] = t * 80 + rs.permutation(80)[:counts[t]]
    cnt = np.array(counts, np.int32)
    Xd = torch.from_numpy(X).to(cuda)
    G = torch.full((n, L.ld), float("nan"), device=cuda)
    loss = torch.zeros(n, device=cuda)
    ws = K.Workspace(cuda)
    K.grad_wave(L, Xd, G, torch.arange(n, dtype=torch.int32, device=cuda),
                torch.from_numpy(imgs.reshape(-1, 3072)).to(cuda), torch.from_numpy(labs.reshape(-1)).to(cuda),
                torch.from_numpy(idx).to(cuda), torch.from_numpy(cnt).to(cuda), batch, ws, loss_out=loss,
                precision=precision)
    torch.cuda.synchronize()
    tol = TOL[precision]
    Gh = G.cpu().numpy()
    flat_i, flat_l = imgs.reshape(-1, 32, 32, 3), labs.reshape(-1)
    for t in range(n):
        if counts[t] == 0:
            continue
        sel = idx[t, :counts[t]]
        gref, lref = O.gradient_of(L, X[t], flat_i[sel], flat_l[sel])
        assert abs(float(loss[t]) - lref) < tol["loss"]
        errs = {}
        for key, shp, off in L.entries:
            sz = int(np.prod(shp))
            errs[key] = rel(Gh[t, off:off + sz], gref[off:off + sz])
        assert max(errs.values()) < tol["grad_t"], (t, {k: round(v, 5) for k, v in errs.items()})
        assert rel(Gh[t, :L.P], gref) < tol["grad_all"]
        if precision == 1:
            # against the oracle that rounds the conv2/conv3 operands to bf16 the same way, the kernels
            # must agree tightly: what is left above is rounding, not an implementation error
            gemu, _ = O.gradient_of(L, X[t], flat_i[sel], flat_l[sel], emulate_bf16=True, wgrad_bf16=WGRAD_TC,
                                    conv1_wgrad_bf16=CONV1_WGRAD_TC)
            errs = {key: rel(Gh[t, off:off + int(np.prod(shp))], gemu[off:off + int(np.prod(shp))])
                    for key, shp, off in L.entries}
            # (residual: values that sit next to a bf16 rounding boundary flip with fp32 noise)
            assert max(errs.values()) < 2e-2, (t, {k: round(v, 5) for k, v in errs.items()})


@pytest.mark.parametrize("precision", [0, 1])
def test_train_kat_two_rounds(cuda, golden_dir, precision):
    """The reference-generated fixture: one peer, two consecutive train tasks (momentum carried
    over, partial last batch).  fp32 path: parameters within 3e-4 relative (bound: 1e-3)."""
    t = np.load(os.path.join(golden_dir, "train_kat.npz"))
    meta = json.loads(str(t["meta"]))
    imgs, labs = synthetic.make_partitions(meta["n_nodes"], meta["n_i"], 32, 3, 10)
    peer, n_i = meta["peer"], meta["n_i"]
    init = np.load(os.path.join(golden_dir, "init_gnlenet.npz"))["row"]
    X = torch.zeros((2, L.ld), device=cuda)
    X[1, :L.P] = torch.from_numpy(init).to(cuda)
    M = torch.zeros((meta["n_nodes"], L.ld), device=cuda)
    ws = K.Workspace(cuda)
    images = torch.from_numpy(imgs.reshape(-1, 3072)).to(cuda)
    labels = torch.from_numpy(labs.reshape(-1)).to(cuda)
    for rnd, key in ((1, "after_round1"), (2, "after_round2")):
        idx, cnt = synthetic.minibatch_index(meta["seed"], [peer], [rnd], n_i, 32, 5)
        gidx = np.where(idx >= 0, idx + peer * n_i, -1).astype(np.int32)
        mb_index = torch.from_numpy(np.ascontiguousarray(gidx.transpose(1, 0, 2))).to(cuda)   # [steps, tasks, B]
        mb_count = torch.from_numpy(np.ascontiguousarray(cnt.T)).to(cuda)
        first = torch.tensor([1 if rnd == 1 else 0], dtype=torch.uint8, device=cuda)
        K.train_wave(L, X, M, torch.tensor([1], dtype=torch.int32, device=cuda),
                     torch.tensor([peer], dtype=torch.int32, device=cuda), first, images, labels,
                     mb_index, mb_count, 32, 0.002, 0.9, 0.0, ws, precision=precision)
        torch.cuda.synchronize()
        got = X[1, :L.P].cpu().numpy()
        assert rel(got, t[key]) < TOL[precision]["par"], rel(got, t[key])
        # the update itself (difference from the init) must match too, not just the large weights
        assert rel(got - init, t[key] - init) < TOL[precision]["upd"]


def test_train_wave_many_nodes_vs_oracle(cuda, data):
    """Several nodes in one wave with different models, weight decay on, momentum from a
    previous buffer for some nodes and first-step for others."""
    imgs, labs = data
    n, n_i, steps, B = 5, 80, 3, 32
    X = perturbed_rows(n, seed=7)
    rs = np.random.RandomState(11)
    Mh = np.zeros((6, L.ld), np.float32)
    Mh[:, :L.P] = 0.01 * rs.standard_normal((6, L.P)).astype(np.float32)
    peers = [4, 0, 2, 5, 1]
    first = [1, 0, 0, 1, 0]
    idx, cnt = synthetic.minibatch_index(42, peers, [3] * n, n_i, B, steps)
    gidx = np.where(idx >= 0, idx + np.array(peers)[:, None, None] * n_i, -1).astype(np.int32)
    Xd = torch.from_numpy(X).to(cuda)
    Md = torch.from_numpy(Mh).to(cuda)
    ws = K.Workspace(cuda)
    loss = torch.zeros((steps, n), device=cuda)
    K.train_wave(L, Xd, Md, torch.arange(n, dtype=torch.int32, device=cuda),
                 torch.tensor(peers, dtype=torch.int32, device=cuda),
                 torch.tensor(first, dtype=torch.uint8, device=cuda),
                 torch.from_numpy(imgs.reshape(-1, 3072)).to(cuda), torch.from_numpy(labs.reshape(-1)).to(cuda),
                 torch.from_numpy(np.ascontiguousarray(gidx.transpose(1, 0, 2))).to(cuda),
                 torch.from_numpy(np.ascontiguousarray(cnt.T)).to(cuda), B, 0.01, 0.9, 1e-3, ws, loss_out=loss)
    torch.cuda.synchronize()
    Xg, Mg = Xd.cpu().numpy(), Md.cpu().numpy()
    for t in range(n):
        row, mom, losses = O.train_task(L, X[t], None if first[t] else Mh[peers[t]], imgs[peers[t]], labs[peers[t]],
                                        idx[t], cnt[t], 0.01, 0.9, 1e-3)
        assert rel(Xg[t, :L.P], row[:L.P]) < 3e-4
        assert rel(Xg[t, :L.P] - X[t, :L.P], row[:L.P] - X[t, :L.P]) < 2e-3
        assert rel(Mg[peers[t], :L.P], mom[:L.P]) < 2e-3
        np.testing.assert_allclose(loss[:, t].cpu().numpy(), np.array(losses), rtol=1e-3)
    # untouched momentum row stays untouched
    assert np.array_equal(Mg[3], Mh[3])


def test_kernel_variants_agree(cuda, data):
    """The tensor-core path's build-time alternatives, switched with dls_tune: the optimizer update fused into the
    weight-gradient epilogues must give the bits of the separate update kernel, and the plane / bulk-copy / tap-pair
    convolution kernels must agree with the round-1 kernels (one MMA per tap) to accumulation-order noise."""
    from dls_b200 import _lib
    imgs, labs = data
    n, n_i, steps, B = 5, 80, 3, 32
    peers = [4, 0, 2, 5, 1]
    idx, cnt = synthetic.minibatch_index(42, peers, [1] * n, n_i, B, steps)
    gidx = np.where(idx >= 0, idx + np.array(peers)[:, None, None] * n_i, -1).astype(np.int32)
    out = {}
    try:
        for name, key, val in (("default", 5, 0), ("unfused", 5, 1), ("round1_convs", 3, 1)):
            _lib.lib().dls_tune(3, 0); _lib.lib().dls_tune(5, 0)
            _lib.lib().dls_tune(key, val)
            Xd = torch.from_numpy(perturbed_rows(n, seed=5)).to(cuda)
            Md = torch.zeros((6, L.ld), device=cuda)
            K.train_wave(L, Xd, Md, torch.arange(n, dtype=torch.int32, device=cuda),
                         torch.tensor(peers, dtype=torch.int32, device=cuda), torch.ones(n, dtype=torch.uint8, device=cuda),
                         torch.from_numpy(imgs.reshape(-1, 3072)).to(cuda), torch.from_numpy(labs.reshape(-1)).to(cuda),
                         torch.from_numpy(np.ascontiguousarray(gidx.transpose(1, 0, 2))).to(cuda),
                         torch.from_numpy(np.ascontiguousarray(cnt.T)).to(cuda), B, 0.002, 0.9, 0.0, K.Workspace(cuda),
                         precision=1)
            torch.cuda.synchronize()
            out[name] = (Xd[:, :L.P].cpu().numpy(), Md[:, :L.P].cpu().numpy())
    finally:
        _lib.lib().dls_tune(3, 0); _lib.lib().dls_tune(5, 0)
    assert np.array_equal(out["default"][0], out["unfused"][0]) and np.array_equal(out["default"][1], out["unfused"][1])
    # different MMA grouping = different fp32 accumulation order; three momentum steps amplify it a little
    rx, rm = rel(out["default"][0], out["round1_convs"][0]), rel(out["default"][1], out["round1_convs"][1])
    assert rx < 3e-4 and rm < 3e-2, (rx, rm)
