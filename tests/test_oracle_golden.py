"""The oracle (oracle/) against fixtures produced by the UNMODIFIED reference
(tests/golden/make_golden.py).  CPU only."""
import json
import os
import pickle

import numpy as np
import pytest
import torch

from oracle import oracle as O
import dls_b200
from dls_b200 import synthetic
from dls_b200.layout import GNLENET as L
from katgen import kat_rows

CASES = [("f5", 5, False), ("f11", 11, False), ("f2w", 2, True), ("f5w", 5, True)]


@pytest.fixture(scope="module", autouse=True)
def one_thread():
    n = torch.get_num_threads()
    torch.set_num_threads(1)      # the fixtures were made with 1 thread; bits depend on it
    yield
    torch.set_num_threads(n)


@pytest.mark.parametrize("tag,fan,weighted", CASES)
def test_fedavg_bit_exact_numpy_and_c(golden_dir, tag, fan, weighted):
    g = np.load(os.path.join(golden_dir, "fedavg_kat.npz"))
    X = kat_rows(fan, L.P, salt=fan + (100 if weighted else 0))
    ws = list(g[tag + "_w"]) if weighted else None
    y = O.fedavg_numpy([X[i] for i in range(fan)], ws)
    assert np.array_equal(y.view(np.uint32), g[tag + "_y"].view(np.uint32))
    wv = np.array(ws if ws else [1.0 / fan] * fan, dtype=np.float32)
    Y = O.fedavg_csr(X, np.array([0, fan]), np.arange(fan), wv, L.P)
    assert np.array_equal(Y[0, :L.P].view(np.uint32), g[tag + "_y"].view(np.uint32))


def test_fedavg_is_not_fma_or_mean(golden_dir):
    """The fixture distinguishes the reference's mul-then-add from an FMA / mean formulation."""
    g = np.load(os.path.join(golden_dir, "fedavg_kat.npz"))
    X = kat_rows(11, L.P, salt=11)
    mean = X.astype(np.float64).mean(axis=0).astype(np.float32)
    assert not np.array_equal(mean, g["f11_y"])


def test_initial_model(golden_dir):
    init = np.load(os.path.join(golden_dir, "init_gnlenet.npz"))["row"]
    r = O.initial_row(L, 42)
    assert np.array_equal(r[:L.P], init)
    assert np.all(r[L.P:] == 0)
    from dls_b200.models import initial_row
    assert np.array_equal(initial_row(L, 42).numpy()[:L.P], init)


def test_train_two_rounds_momentum_and_partial_batch(golden_dir):
    t = np.load(os.path.join(golden_dir, "train_kat.npz"))
    meta = json.loads(str(t["meta"]))
    imgs, labs = synthetic.make_partitions(meta["n_nodes"], meta["n_i"], 32, 3, 10)
    peer = meta["peer"]
    row, mom = O.initial_row(L, meta["seed"]), None
    for rnd, key in ((1, "after_round1"), (2, "after_round2")):
        idx, cnt = synthetic.minibatch_index(meta["seed"], [peer], [rnd], meta["n_i"], 32, 5)
        assert list(cnt[0]) == [32, 32, 16, 0, 0]           # min(local_steps, ceil(80/32)) = 3 steps
        row, mom, _ = O.train_task(L, row, mom, imgs[peer], labs[peer], idx[0], cnt[0], 0.002, 0.9, 0.0)
        # bit-exact on the fixture's torch build/thread count; 1e-6 otherwise
        if torch.__version__ == meta["torch"]:
            assert np.array_equal(row[:L.P], t[key])
        else:
            np.testing.assert_allclose(row[:L.P], t[key], rtol=1e-5, atol=1e-7)


def test_dpsgd_mini_full_solve(golden_dir):
    g = np.load(os.path.join(golden_dir, "dpsgd_mini.npz"))
    meta = json.loads(str(g["meta"]))
    with open(os.path.join(golden_dir, "dpsgd_mini_dag.pkl"), "rb") as f:
        dag = pickle.load(f)
    assert len(dag) == meta["n_tasks"] == 2 * meta["rounds"] * meta["n"] + meta["n"]
    imgs, labs = synthetic.make_partitions(meta["n"], meta["n_i"], 32, 3, 10)
    timg, tlab = synthetic.make_test_split(meta["n_test"], 32, 3, 10)
    sinks, acc = O.solve_dag(L, dag, imgs, labs, timg, tlab, seed=meta["seed"], batch=32,
                             local_steps=meta["local_steps"], lr=0.002, mu=0.9, wd=0.0, n_i=meta["n_i"],
                             minibatch_permutation=synthetic.minibatch_permutation)
    final = np.zeros((meta["n"], L.P), np.float32)
    for name, row in sinks.items():
        final[int(name.split("_")[1])] = row[:L.P]
    if torch.__version__ == meta["torch"]:
        assert np.array_equal(final[0], g["final_node0"])
        assert np.array_equal(final[5], g["final_node5"])
        assert np.array_equal(final[:, ::64], g["final_strided"])
    else:
        np.testing.assert_allclose(final[:, ::64], g["final_strided"], rtol=1e-5, atol=1e-7)
    # accuracy CSV rows in the reference's format (functions.py:121-128)
    for peer, rnd, t_us, a, l in acc:
        line = "dpsgd,cifar10,iid,%g,%d,%d,%.2f,%f,%f" % (1, peer, rnd, t_us / 1e6, a, l)
        assert line in str(g["acc_csv"][peer])


def test_compute_gradient_and_gradient_update_known_answers(golden_dir):
    """AD-PSGD entry points, pinned to the reference's functions.compute_gradient / gradient_update
    (tests/golden/make_golden.py gen_gradient): the gradient of the first scripted minibatch, and one
    step of a fresh SGD optimizer (no momentum carry-over) with a fresh and with a stale gradient."""
    t = np.load(os.path.join(golden_dir, "gradient_kat.npz"))
    meta = json.loads(str(t["meta"]))
    imgs, labs = synthetic.make_partitions(meta["n_nodes"], meta["n_i"], 32, 3, 10)
    peer = meta["peer"]
    init = O.initial_row(L, meta["seed"])
    exact = torch.__version__ == meta["torch"]

    def check(a, b):
        if exact:
            assert np.array_equal(a, b)
        else:
            np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-7)

    assert np.array_equal(init[:L.P], t["model_after_cg"])             # compute_gradient leaves the model alone
    perm = synthetic.minibatch_permutation(meta["seed"], peer, 1, meta["n_i"])[:32]
    g1, _ = O.gradient_of(L, init, imgs[peer][perm], labs[peer][perm])
    check(g1, t["grad1"])
    check(O.gradient_update_task(init, t["grad1"], L.P, meta["lr"], meta["wd"])[:L.P], t["upd1"])
    # momentum 0.9 is configured, yet the step is p - lr*g: the optimizer of gradient_update is fresh
    assert np.array_equal(t["upd1"], (t["model_after_cg"].astype(np.float64)
                                      - np.float64(np.float32(meta["lr"])) * t["grad1"].astype(np.float64)).astype(np.float32))
    check(O.gradient_update_task(t["trained"], t["grad1"], L.P, meta["lr"], meta["wd"])[:L.P], t["upd2"])   # stale gradient
    perm3 = synthetic.minibatch_permutation(meta["seed"], peer, 3, meta["n_i"])[:32]
    g2, _ = O.gradient_of(L, t["upd2"], imgs[peer][perm3], labs[peer][perm3])
    check(g2, t["grad2"])


def test_dpsgd_c1_full_solve(golden_dir):
    """BASELINE config C1 (16 nodes, 4-regular, 5 rounds x 5 local steps, test at round 5) as solved by the
    reference itself (make_golden.py gen_c1): the oracle reproduces every node's final model and CSV row."""
    g = np.load(os.path.join(golden_dir, "dpsgd_c1.npz"))
    meta = json.loads(str(g["meta"]))
    with open(os.path.join(golden_dir, "dpsgd_c1_dag.pkl"), "rb") as f:
        dag = pickle.load(f)
    n = meta["n"]
    assert (n, meta["k"], meta["rounds"]) == (16, 4, 5)
    assert len(dag) == meta["n_tasks"] == 2 * meta["rounds"] * n + n
    assert all(len(t["data"]["models"]) == meta["k"] + 1 for t in dag if t["func"] == "aggregate")
    imgs, labs = synthetic.make_partitions(n, meta["n_i"], 32, 3, 10)
    timg, tlab = synthetic.make_test_split(meta["n_test"], 32, 3, 10)
    sinks, acc = O.solve_dag(L, dag, imgs, labs, timg, tlab, seed=meta["seed"], batch=32,
                             local_steps=meta["local_steps"], lr=0.002, mu=0.9, wd=0.0, n_i=meta["n_i"],
                             minibatch_permutation=synthetic.minibatch_permutation)
    final = np.zeros((n, L.P), np.float32)
    for name, row in sinks.items():
        final[int(name.split("_")[1])] = row[:L.P]
    if torch.__version__ == meta["torch"]:
        assert np.array_equal(final[0], g["final_node0"])
        assert np.array_equal(final[11], g["final_node11"])
        assert np.array_equal(final[:, ::16], g["final_strided"])
    else:
        np.testing.assert_allclose(final[:, ::16], g["final_strided"], rtol=1e-5, atol=1e-7)
    for peer, rnd, t_us, a, l in acc:
        line = "dpsgd,cifar10,iid,%g,%d,%d,%.2f,%f,%f" % (1, peer, rnd, t_us / 1e6, a, l)
        assert line in str(g["acc_csv"][peer])


def femnist_kept_mask(LF, stride):
    o1 = LF.offset("fc1.weight")
    n1 = 512 * 3136
    keep = np.ones(LF.P, dtype=bool)
    keep[o1:o1 + n1] = False
    keep[o1:o1 + n1:stride] = True
    return keep


def test_femnist_pinned_to_reference_model_class(golden_dir):
    """FEMNIST CNN: init, logits, CrossEntropy loss, autograd gradients and three SGD steps (32, 32, 20
    samples) produced by the reference's own `CNN` + `SGDOptimizer` classes (make_golden.py gen_femnist)."""
    import hashlib
    from dls_b200.layout import FEMNIST as LF
    from dls_b200.models import initial_row
    g = np.load(os.path.join(golden_dir, "femnist_kat.npz"))
    meta = json.loads(str(g["meta"]))
    exact = torch.__version__ == meta["torch"]
    init = O.initial_row(LF, meta["seed"])
    assert hashlib.sha256(init[:LF.P].tobytes()).digest() == bytes(g["init_sha"])
    assert np.array_equal(initial_row(LF, meta["seed"]).numpy(), init)           # product-side init, same bits
    assert np.array_equal(init[:LF.P:64], g["init_strided"])
    imgs, labs = synthetic.make_images(meta["n_i"], 28, 1, 62, seed=meta["image_seed"])
    order = synthetic.minibatch_permutation(meta["seed"], meta["peer"], meta["round"], meta["n_i"])
    B = meta["batch"]
    sel = order[:B]
    keep = femnist_kept_mask(LF, meta["fc1_stride"])

    def check(a, b, tol=1e-5):
        if exact:
            assert np.array_equal(a, b)
        else:
            np.testing.assert_allclose(a, b, rtol=tol, atol=1e-7)

    check(O.logits(LF, init, imgs[sel]), g["logits"])
    grad, loss = O.gradient_of(LF, init, imgs[sel], labs[sel])
    check(np.float32(loss), g["loss"])
    check(grad[keep], g["grad_kept"])
    idx = -np.ones((3, B), np.int32)
    cnt = np.array(meta["counts"], np.int32)
    for s, c in enumerate(cnt):
        idx[s, :c] = order[s * B: s * B + c]
    row, mom, losses = O.train_task(LF, init, None, imgs, labs, idx, cnt, meta["lr"], meta["momentum"], meta["wd"])
    assert mom is None
    check(row[:LF.P][keep], g["after_kept"])
    np.testing.assert_allclose(losses, g["step_losses"], rtol=1e-6)
    check(O.logits(LF, row, imgs[order[:64]]), g["logits_after"])


def test_chunk_ops_known_answers(golden_dir):
    """functions.chunk / reconstruct_from_chunks of the reference (make_golden.py gen_chunks): the slice
    bounds and torch.mean(torch.stack(...), 0)'s summation order for 1..9 contributors, bit for bit."""
    g = np.load(os.path.join(golden_dir, "chunk_kat.npz"))
    meta = json.loads(str(g["meta"]))
    sets = json.loads(str(g["sets"]))
    X = kat_rows(meta["n_models"], L.P, salt=meta["salt"])
    bounds = O.chunk_bounds(L.P, meta["n_chunks"])
    assert [c1 - c0 for c0, c1 in bounds] == list(g["sizes"])
    assert np.array_equal(X[1][bounds[0][0]:bounds[0][1]], g["piece0_of_model1"])
    y = O.reconstruct_from_chunks(L.P, [[X[mi][c0:c1] for mi in sets[c]] for c, (c0, c1) in enumerate(bounds)])
    assert np.array_equal(y.view(np.uint32), g["y"].view(np.uint32))
    # the fixture separates torch's order from a plain sequential mean (tail columns of chunks with >= 5 pieces)
    c0, c1 = bounds[4]
    seq = np.zeros(c1 - c0, np.float32)
    for mi in sets[4]:
        seq = seq + X[mi][c0:c1]
    assert not np.array_equal(seq / np.float32(len(sets[4])), g["y"][c0:c1])


@pytest.mark.parametrize("n,k", [(2, 3), (7, 9), (33, 5), (100, 18), (257, 40)])
def test_mean_stack_matches_torch(n, k):
    rs = np.random.RandomState(n * 1000 + k)
    x = (rs.standard_normal((k, n)) * np.exp2(rs.randint(-8, 8, (k, n)))).astype(np.float32)
    y = torch.mean(torch.stack([torch.from_numpy(x[i].copy()) for i in range(k)]), dim=0).numpy()
    assert np.array_equal(y.view(np.uint32), O.mean_stack(list(x)).view(np.uint32))
