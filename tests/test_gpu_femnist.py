"""FEMNIST CNN train / grad / test waves (CUDA fp32 path) against the oracle (oracle/oracle.py) and against
tests/golden/femnist_kat.npz, which the reference's own `CNN` and `SGDOptimizer` classes produced
(dasklearn/models/femnist.py:45-85 is importable once the missing module dasklearn.datasets.femnist is supplied
as a two-constant shim; the loop around it is model_trainer.py:99-126, which the reference's trainer itself
refuses to run for this dataset)."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
import dls_b200
from dls_b200 import kernels as K, synthetic
from dls_b200.dataset import DeviceDataset
from dls_b200.executor import WaveExecutor
from dls_b200.layout import FEMNIST as L
from dls_b200.settings import SessionSettings, LearningSettings
from dls_b200.solver import BatchedSolver

pytestmark = pytest.mark.gpu


def rel(a, b):
    return float(np.linalg.norm(a.astype(np.float64) - b.astype(np.float64)) /
                 max(np.linalg.norm(b.astype(np.float64)), 1e-30))


def rows(n, seed=0):
    base = O.initial_row(L, 42)
    rs = np.random.RandomState(seed)
    X = np.tile(base, (n, 1))
    X[:, :L.P] += 0.01 * rs.standard_normal((n, L.P)).astype(np.float32)
    return X.astype(np.float32)


@pytest.fixture(scope="module")
def data():
    return synthetic.make_partitions(4, 48, 28, 1, 62)


def test_layout_and_init():
    assert L.P == 1690046 and L.ld % 64 == 0
    from dls_b200.models import initial_row
    assert np.array_equal(initial_row(L, 42).numpy(), O.initial_row(L, 42))


@pytest.mark.parametrize("precision", [0, 1])
def test_eval(cuda, precision):
    timg, tlab = synthetic.make_test_split(100, 28, 1, 62)
    X = rows(2)
    Xd = torch.from_numpy(X).to(cuda)
    logits = torch.zeros((2, 100, 62), device=cuda)
    correct, loss = K.eval_wave(L, Xd, torch.tensor([1, 0], dtype=torch.int32, device=cuda),
                                torch.from_numpy(timg.reshape(100, -1)).to(cuda), torch.from_numpy(tlab).to(cuda),
                                K.Workspace(cuda), out_logits=logits, precision=precision)
    torch.cuda.synchronize()
    for j, r in enumerate([1, 0]):
        ref = O.logits(L, X[r], timg)
        assert np.abs(logits[j].cpu().numpy() - ref).max() < (5e-3 if precision else 2e-4)
        acc, l, nc = O.evaluate(L, X[r], timg, tlab)
        assert abs(int(correct[j]) - nc) <= (3 if precision else 1)
        assert abs(float(loss[j]) / 100 - l) < (2e-3 if precision else 1e-4)


# precision 1 = conv2 on bf16 tensor cores (fp32 accumulate): the fp32 autograd oracle is matched to bf16 rounding
# (2^-9 relative per operand), the fp32 path to summation-order noise.
@pytest.mark.parametrize("precision", [0, 1])
@pytest.mark.parametrize("batch,counts", [(32, [32, 16, 1]), (8, [8, 0, 5])])
def test_gradient_vs_autograd(cuda, data, batch, counts, precision):
    imgs, labs = data
    n = len(counts)
    X = rows(n, seed=1)
    rs = np.random.RandomState(2)
    idx = -np.ones((n, batch), np.int32)
    for t in range(n):
        idx[t, :counts[t]] = t * 48 + rs.permutation(48)[:counts[t]]
    Xd = torch.from_numpy(X).to(cuda)
    G = torch.full((n, L.ld), float("nan"), device=cuda)
    loss = torch.zeros(n, device=cuda)
    K.grad_wave(L, Xd, G, torch.arange(n, dtype=torch.int32, device=cuda),
                torch.from_numpy(imgs.reshape(-1, 784)).to(cuda), torch.from_numpy(labs.reshape(-1)).to(cuda),
                torch.from_numpy(idx).to(cuda), torch.tensor(counts, dtype=torch.int32, device=cuda), batch,
                K.Workspace(cuda), loss_out=loss, precision=precision)
    torch.cuda.synchronize()
    Gh = G.cpu().numpy()
    fi, fl = imgs.reshape(-1, 28, 28, 1), labs.reshape(-1)
    for t in range(n):
        if counts[t] == 0:
            continue
        sel = idx[t, :counts[t]]
        gref, lref = O.gradient_of(L, X[t], fi[sel], fl[sel])
        assert abs(float(loss[t]) - lref) < (2e-3 if precision else 1e-4)
        errs = {key: rel(Gh[t, off:off + int(np.prod(shp))], gref[off:off + int(np.prod(shp))])
                for key, shp, off in L.entries}
        # against fp32 autograd the bf16 path shows the rounding of a random-init network (heavy cancellation) ...
        assert max(errs.values()) < (1.5e-1 if precision else 2e-4), (t, errs)
        assert not np.isnan(Gh[t, :L.P]).any()
        if precision:
            # ... against the same network with conv2's operands rounded to bf16 only accumulation order is left
            gemu, lemu = O.gradient_of(L, X[t], fi[sel], fl[sel], emulate_bf16=True)
            assert abs(float(loss[t]) - lemu) < 1e-4
            errs = {key: rel(Gh[t, off:off + int(np.prod(shp))], gemu[off:off + int(np.prod(shp))])
                    for key, shp, off in L.entries}
            assert max(errs.values()) < 5e-3, (t, errs)


@pytest.mark.parametrize("precision", [0, 1])
@pytest.mark.parametrize("mu", [0.0, 0.9])
def test_train_wave_vs_oracle(cuda, data, mu, precision):
    imgs, labs = data
    n, n_i, steps, B = 3, 48, 3, 32                      # 48 samples: steps of 32 + 16, third step empty
    X = rows(n, seed=3)
    peers = [2, 0, 3]
    idx, cnt = synthetic.minibatch_index(42, peers, [1] * n, n_i, B, steps)
    gidx = np.where(idx >= 0, idx + np.array(peers)[:, None, None] * n_i, -1).astype(np.int32)
    Xd = torch.from_numpy(X).to(cuda)
    Md = torch.zeros((4, L.ld), device=cuda) if mu else None
    K.train_wave(L, Xd, Md, torch.arange(n, dtype=torch.int32, device=cuda),
                 torch.tensor(peers, dtype=torch.int32, device=cuda), torch.ones(n, dtype=torch.uint8, device=cuda),
                 torch.from_numpy(imgs.reshape(-1, 784)).to(cuda), torch.from_numpy(labs.reshape(-1)).to(cuda),
                 torch.from_numpy(np.ascontiguousarray(gidx.transpose(1, 0, 2))).to(cuda),
                 torch.from_numpy(np.ascontiguousarray(cnt.T)).to(cuda), B, 0.004, mu, 0.0, K.Workspace(cuda),
                 precision=precision)
    torch.cuda.synchronize()
    Xg = Xd.cpu().numpy()
    for t in range(n):
        row, mom, _ = O.train_task(L, X[t], None, imgs[peers[t]], labs[peers[t]], idx[t], cnt[t], 0.004, mu, 0.0)
        assert rel(Xg[t, :L.P], row[:L.P]) < (1e-4 if precision else 1e-5)     # north-star bound: 1e-3
        assert rel(Xg[t, :L.P] - X[t, :L.P], row[:L.P] - X[t, :L.P]) < (1.5e-1 if precision else 1e-3)
        if mu:
            assert rel(Md[peers[t], :L.P].cpu().numpy(), mom[:L.P]) < (1.5e-1 if precision else 1e-3)


def test_solver_dpsgd_femnist(cuda, tmp_path):
    """4 nodes on a ring, 2 rounds + test, reference defaults for femnist (lr 0.004, momentum 0,
    args.py:70-72) through the batched solver vs the oracle's sequential solve."""
    n, n_i, n_test, steps = 4, 64, 64, 2
    st = SessionSettings(participants=n, seed=42, dataset="femnist", model=None, data_dir=str(tmp_path),
                         learning=LearningSettings(0.004, 0.0, 0.0, 32, steps))
    dag = synthetic.dpsgd_dag(n, 2, synthetic.ring_topology(n), test_interval=2)
    ds = DeviceDataset.synthetic(L, n, n_i, n_test, cuda, seed=42)
    ex = WaveExecutor(st, ds)
    res = BatchedSolver(ex, dag).plan().run()
    imgs, labs = synthetic.make_partitions(n, n_i, 28, 1, 62)
    timg, tlab = synthetic.make_test_split(n_test, 28, 1, 62)
    sinks, acc = O.solve_dag(L, dag, imgs, labs, timg, tlab, seed=42, batch=32, local_steps=steps, lr=0.004, mu=0.0,
                             wd=0.0, n_i=n_i, minibatch_permutation=synthetic.minibatch_permutation)
    for name, row in sinks.items():
        assert rel(res[name][0].numpy(), row[:L.P]) < 1e-5
    for peer, rnd, t_us, a, l in acc:
        got = open(str(tmp_path / ("accuracies_%d.csv" % peer))).read().strip().split("\n")[1].split(",")
        assert got[1] == "femnist" and abs(float(got[7]) - a) <= 100.0 / n_test + 1e-9 and abs(float(got[8]) - l) < 1e-3


def test_pinned_to_reference_model_class(cuda, golden_dir):
    """The CUDA FEMNIST kernels against outputs of the reference's OWN `CNN` + `SGDOptimizer` classes
    (tests/golden/femnist_kat.npz, make_golden.py gen_femnist): logits, loss, gradient of one minibatch, and
    three optimizer steps with a partial last batch (32, 32, 20)."""
    import json, os
    g = np.load(os.path.join(golden_dir, "femnist_kat.npz"))
    meta = json.loads(str(g["meta"]))
    from dls_b200.models import initial_row
    init = initial_row(L, meta["seed"]).numpy()
    assert np.array_equal(init[:L.P:64], g["init_strided"])
    imgs, labs = synthetic.make_images(meta["n_i"], 28, 1, 62, seed=meta["image_seed"])
    order = synthetic.minibatch_permutation(meta["seed"], meta["peer"], meta["round"], meta["n_i"])
    B = meta["batch"]
    o1 = L.offset("fc1.weight")
    keep = np.ones(L.P, dtype=bool)
    keep[o1:o1 + 512 * 3136] = False
    keep[o1:o1 + 512 * 3136:meta["fc1_stride"]] = True
    d_img = torch.from_numpy(imgs.reshape(-1, 784)).to(cuda)
    d_lab = torch.from_numpy(labs.astype(np.int64)).to(cuda)
    Xd = torch.from_numpy(init[None, :].copy()).to(cuda)
    row0 = torch.zeros(1, dtype=torch.int32, device=cuda)
    # logits of the first minibatch through the test wave
    sel = order[:B]
    logits = torch.zeros((1, B, 62), device=cuda)
    K.eval_wave(L, Xd, row0, d_img[torch.from_numpy(sel.astype(np.int64)).to(cuda)].contiguous(),
                d_lab[torch.from_numpy(sel.astype(np.int64)).to(cuda)].contiguous(), K.Workspace(cuda),
                out_logits=logits)
    assert np.abs(logits[0].cpu().numpy() - g["logits"]).max() < 1e-5
    # gradient + loss of the first minibatch
    G = torch.zeros((1, L.ld), device=cuda)
    loss = torch.zeros(1, device=cuda)
    idx0 = torch.from_numpy(sel.astype(np.int32)[None, :].copy()).to(cuda)
    K.grad_wave(L, Xd, G, row0, d_img, d_lab, idx0, torch.tensor([B], dtype=torch.int32, device=cuda), B,
                K.Workspace(cuda), loss_out=loss)
    torch.cuda.synchronize()
    assert abs(float(loss[0]) - float(g["loss"])) < 1e-5
    Gh = G[0, :L.P].cpu().numpy()
    assert rel(Gh[keep], g["grad_kept"]) < 1e-4
    assert abs(np.linalg.norm(Gh.astype(np.float64)) - float(g["grad_l2"])) < 1e-4 * float(g["grad_l2"])
    # three optimizer steps (lr 0.004, momentum 0), partial last batch
    counts = meta["counts"]
    idx = -np.ones((3, 1, B), np.int32)
    for s, c in enumerate(counts):
        idx[s, 0, :c] = order[s * B: s * B + c]
    lo = torch.zeros((3, 1), device=cuda)
    K.train_wave(L, Xd, None, row0, row0, torch.ones(1, dtype=torch.uint8, device=cuda), d_img, d_lab,
                 torch.from_numpy(idx).to(cuda), torch.tensor(counts, dtype=torch.int32, device=cuda).reshape(3, 1), B,
                 meta["lr"], meta["momentum"], meta["wd"], K.Workspace(cuda), loss_out=lo)
    torch.cuda.synchronize()
    after = Xd[0, :L.P].cpu().numpy()
    assert rel(after[keep], g["after_kept"]) < 1e-5
    assert rel(after[keep] - init[:L.P][keep], g["after_kept"] - init[:L.P][keep]) < 1e-3
    assert abs(after.astype(np.float64).sum() - float(g["after_sum"])) < 1e-2
    assert np.abs(lo[:, 0].cpu().numpy() - g["step_losses"]).max() < 1e-4
