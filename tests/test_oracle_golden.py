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
