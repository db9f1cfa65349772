"""Host logic of the batched solver (no GPU): levelisation, slot assignment, reference counts,
reference DAG invariants (tests/test_dpsgd.py:26-35 of the reference: 2*R*N tasks, aggregate
fan-in k+1, train fan-out k+1)."""
import gzip
import os
import pickle

import numpy as np
import pytest

import dls_b200
from dls_b200 import synthetic
from dls_b200.settings import SessionSettings, LearningSettings
from dls_b200.solver import BatchedSolver


class PlanOnlyExecutor:
    """Stands in for WaveExecutor during planning: records what would be uploaded."""

    def __init__(self, n_slots, settings):
        self.settings = settings
        self.n_slots = n_slots
        self.free = list(range(n_slots - 1, -1, -1))
        self.refs = np.zeros(n_slots, dtype=np.int64)
        self.grown = 0

    def _grow(self, extra):
        self.grown += extra
        self.n_slots += extra
        self.refs = np.concatenate([self.refs, np.zeros(extra, dtype=np.int64)])

    def prepare_train(self, peers, rounds, steps):
        return ("train", list(peers), list(rounds), list(steps))

    def prepare_aggregate(self, model_slots, out_slots, weights):
        srcs = {s for ms in model_slots for s in ms}
        assert not (srcs & set(out_slots)), "aggregate wave overwrites its own input"
        return ("agg", [list(m) for m in model_slots], list(out_slots))


def settings(n):
    return SessionSettings(participants=n, learning=LearningSettings())


def test_reference_mini_dag_plan(golden_dir):
    dag = pickle.load(open(os.path.join(golden_dir, "dpsgd_mini_dag.pkl"), "rb"))
    ex = PlanOnlyExecutor(18, settings(8))
    s = BatchedSolver(ex, dag).plan()
    assert [w.func for w in s.waves] == ["train", "aggregate"] * 3 + ["test"]
    assert all(len(w.names) == 8 for w in s.waves)
    assert len(s.waves[0].init_slots) == 8 and not s.waves[0].copy_src
    # from round 2 on every train consumes an aggregate output exclusively: in place, no copies
    for w in s.waves[2::2][:2]:
        assert not w.copy_src and not w.init_slots
    # D-PSGD needs 2N rows: post-train and post-aggregate ping-pong
    assert ex.grown == 0
    used = {x for w in s.waves for x in w.out_slots}
    assert len(used) <= 16
    assert len(s.sink_slots) == 8
    # an aggregate never reads a row it writes, and fan-in = k+1 = 3
    for w in s.waves[1::2][:3]:
        assert all(len(m) == 3 for m in w.model_slots)


def test_reference_c2_dag_invariants(golden_dir):
    dag = pickle.load(gzip.open(os.path.join(golden_dir, "dpsgd_c2_dag.pkl.gz"), "rb"))
    n, k, rounds = 256, 8, 2
    assert len(dag) == 2 * rounds * n
    for t in dag:
        if t["func"] == "aggregate":
            assert len(t["inputs"]) == k + 1 and len(t["data"]["models"]) == k + 1
            assert t["data"]["models"][-1][0].startswith("train_")          # own model last
        else:
            assert t["func"] == "train"
            if t["data"]["round"] < rounds:
                assert len(t["outputs"]) == k + 1
    ex = PlanOnlyExecutor(2 * n + 2, settings(n))
    s = BatchedSolver(ex, dag).plan()
    assert [w.func for w in s.waves] == ["train", "aggregate", "train", "aggregate"]
    assert [len(w.names) for w in s.waves] == [n] * 4 and ex.grown == 0


def test_synthetic_dag_matches_reference_shape():
    n, k, rounds = 16, 4, 3
    nb = synthetic.kreg_topology(n, k, 42)
    dag = synthetic.dpsgd_dag(n, rounds, nb, test_interval=3)
    assert len(dag) == 2 * rounds * n + n
    names = {t["name"] for t in dag}
    for t in dag:
        for src, idx in t["inputs"]:
            assert src in names and idx == 0
        if t["func"] == "aggregate":
            assert len(t["data"]["models"]) == k + 1
            assert t["data"]["models"] == t["inputs"]
    ex = PlanOnlyExecutor(2 * n + 2, settings(n))
    s = BatchedSolver(ex, dag).plan()
    assert [w.func for w in s.waves] == ["train", "aggregate"] * rounds + ["test"]


def test_shared_input_forces_copy_and_churn_fan_in():
    """A model consumed by a train AND an aggregate must not be trained in place; missing
    neighbours (churn) just shrink the list."""
    dag = [
        {"name": "train_1", "func": "train", "data": {"model": (None, 0), "round": 1, "time": 0, "peer": 0},
         "inputs": [], "outputs": [("train_2", 0), ("agg_1", 0)]},
        {"name": "train_2", "func": "train", "data": {"model": ("train_1", 0), "round": 2, "time": 0, "peer": 0},
         "inputs": [("train_1", 0)], "outputs": [("agg_1", 0)]},
        {"name": "agg_1", "func": "aggregate", "data": {"models": [("train_1", 0), ("train_2", 0)], "round": 2,
                                                      "peer": 0, "weights": [0.25, 0.75]},
         "inputs": [("train_1", 0), ("train_2", 0)], "outputs": []},
    ]
    ex = PlanOnlyExecutor(2, settings(1))
    s = BatchedSolver(ex, dag).plan()
    w_train2 = s.waves[1]
    assert w_train2.copy_src and w_train2.copy_dst and w_train2.copy_src != w_train2.copy_dst
    assert s.waves[2].weights == [[0.25, 0.75]]
    assert ex.grown == 1           # three live rows were needed


def test_unknown_function_and_cycle():
    with pytest.raises(RuntimeError, match="not found"):
        BatchedSolver(PlanOnlyExecutor(2, settings(1)),
                      [{"name": "x", "func": "push_model", "data": {"peer": 0}, "inputs": [], "outputs": []}])
    cyc = [{"name": "a", "func": "test", "data": {"model": ("b", 0), "peer": 0}, "inputs": [("b", 0)], "outputs": [("b", 0)]},
           {"name": "b", "func": "test", "data": {"model": ("a", 0), "peer": 0}, "inputs": [("a", 0)], "outputs": [("a", 0)]}]
    with pytest.raises(RuntimeError, match="cycle"):
        BatchedSolver(PlanOnlyExecutor(2, settings(1)), cyc).plan()


def test_settings_from_coordinator_dict():
    d = {"algorithm": "dpsgd", "seed": 7, "work_dir": "/tmp", "dataset": "cifar10",
         "learning": {"learning_rate": 0.01, "momentum": 0.0, "weight_decay": 0.0, "batch_size": 16, "local_steps": 2},
         "participants": 4, "model": "gnlenet", "gradient_aggregation": 0, "unknown_future_field": 1,
         "data_dir": "/tmp/x", "k": 2, "topology": "ring"}
    s = SessionSettings.from_dict(d)
    assert s.learning.batch_size == 16 and s.seed == 7 and s.topology == "ring" and s.data_dir == "/tmp/x"


def test_minibatch_index_partial_batches():
    idx, cnt = synthetic.minibatch_index(42, [3, 4], [1, 1], 48, 32, 5)
    assert cnt.tolist() == [[32, 16, 0, 0, 0], [32, 16, 0, 0, 0]]      # reference defect 14: 2 steps of 32+16
    assert (idx[0, 1, 16:] == -1).all() and sorted(idx[0, :2].ravel()[idx[0, :2].ravel() >= 0]) == list(range(48))


def test_adpsgd_shaped_graph_solves_like_the_sequential_oracle(tmp_path):
    """compute_gradient / gradient_update tasks (bare-name `gradient_model` references, stale gradients,
    two input edges) through the batched solver with oracle arithmetic == the task-at-a-time oracle,
    bit for bit, planned and streamed; every gradient row is returned to the pool."""
    import sys
    import torch
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from cpu_standin import OracleExecutor
    from oracle import oracle as O
    from dls_b200.layout import GNLENET as L
    torch.set_num_threads(1)
    n, n_i = 4, 40
    dag = synthetic.adpsgd_dag(2, 2)
    funcs = [t["func"] for t in dag]
    assert funcs.count("compute_gradient") == 4 and funcs.count("gradient_update") == 4
    assert all(len(t["inputs"]) == 2 for t in dag if t["func"] == "gradient_update")
    imgs, labs = synthetic.make_partitions(n, n_i, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(16, 32, 3, 10)
    st = SessionSettings(participants=n, seed=42, data_dir=str(tmp_path), learning=LearningSettings(0.002, 0.9, 0.0, 32, 2))
    sinks, _ = O.solve_dag(L, dag, imgs, labs, timg, tlab, seed=42, batch=32, local_steps=2, lr=0.002, mu=0.9, wd=0.0,
                           n_i=n_i, minibatch_permutation=synthetic.minibatch_permutation)
    for streamed in (False, True):
        ex = OracleExecutor(st, imgs, labs, timg, tlab, list(range(n)), n_i)
        solver = BatchedSolver(ex, dag)
        if not streamed:
            solver.plan()
        res = solver.run()
        assert sorted(res) == sorted(sinks)
        for name, row in sinks.items():
            assert np.array_equal(ex.X[res[name][0].slot, :L.P].numpy(), row[:L.P]), name
        # only the four sink rows stay referenced: gradient rows and intermediate models were released
        assert int((ex.refs > 0).sum()) == 4
        assert [w.func for w in solver.waves].count("compute_gradient") == 2


def test_conflux_chunk_graph_against_oracle():
    """chunk / reconstruct_from_chunks waves (functions.py:136-146): the planner's aliasing of pieces to
    column ranges, per-edge reference counts and segment lists, executed with oracle arithmetic on CPU,
    must reproduce the oracle's task-by-task solve bit for bit."""
    import torch
    from oracle import oracle as O
    from cpu_standin import OracleExecutor
    from dls_b200.layout import GNLENET as L
    n, n_i = 6, 32
    dag = synthetic.conflux_dag(n, rounds=2, sample_size=3, n_chunks=5, seed=1)
    assert {t["func"] for t in dag} == {"train", "chunk", "reconstruct_from_chunks"}
    imgs, labs = synthetic.make_partitions(n, n_i, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(8, 32, 3, 10)
    st = SessionSettings(participants=n, seed=42, learning=LearningSettings(0.002, 0.9, 0.0, 32, 1))
    torch.set_num_threads(1)
    ex = OracleExecutor(st, imgs, labs, timg, tlab, list(range(n)), n_i)
    solver = BatchedSolver(ex, dag).plan()
    assert any(w.func == "reconstruct_from_chunks" and len(w.segments) == 5 * len(w.names) for w in solver.waves)
    res = solver.run()
    sinks, _ = O.solve_dag(L, dag, imgs, labs, timg, tlab, seed=42, batch=32, local_steps=1, lr=0.002, mu=0.9,
                           wd=0.0, n_i=n_i, minibatch_permutation=synthetic.minibatch_permutation)
    assert sorted(res) == sorted(sinks) and len(sinks) == 3
    for name, row in sinks.items():
        got = ex.X[res[name][0].slot].numpy()
        assert np.array_equal(got[:L.P].view(np.uint32), row[:L.P].view(np.uint32)), name
    # every row went back to the pool except the sinks
    assert int((ex.refs > 0).sum()) == len(sinks)
