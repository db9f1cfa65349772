"""Multi-rank solve (world_size 2, gloo, CPU): node partition, halo plan and the all_to_all
row exchange.  Rows and arithmetic are the oracle's (tests/cpu_standin.py), so the 2-rank result
must equal the single-process oracle solve bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import dls_b200
from dls_b200 import synthetic
from dls_b200.layout import GNLENET as L
from dls_b200.partition import DistributedSolver, block_partition, partition_nodes, topology_from_dag
from dls_b200.settings import SessionSettings, LearningSettings
from dls_b200.solver import BatchedSolver

N, K_REG, ROUNDS, N_I, STEPS, N_TEST = 8, 3, 2, 32, 1, 32


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _settings(tmp):
    return SessionSettings(participants=N, seed=42, data_dir=tmp, learning=LearningSettings(0.002, 0.9, 0.0, 32, STEPS))


def _dag():
    nb = synthetic.kreg_topology(N, K_REG + (1 if (N * K_REG) % 2 else 0), 42)
    return synthetic.dpsgd_dag(N, ROUNDS, nb, test_interval=ROUNDS, order_seed=5), nb


def _worker(rank, world, port, tmp, owner, out_q, streamed=False):
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from cpu_standin import OracleExecutor
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    dag, _ = _dag()
    imgs, labs = synthetic.make_partitions(N, N_I, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(N_TEST, 32, 3, 10)
    owned = [int(p) for p in np.nonzero(np.asarray(owner) == rank)[0]]
    ex = OracleExecutor(_settings(tmp), imgs, labs, timg, tlab, owned, N_I)
    solver = DistributedSolver(ex, dag, rank, world, owner)
    if not streamed:
        solver.plan()
    res = solver.run()              # unplanned: levels are planned, committed and exchanged one at a time
    out = {name: h[0].executor.X[h[0].slot, :L.P].numpy().copy() for name, h in res.items()}
    out_q.put((rank, out, solver.halo_bytes, solver.max_halo, solver.max_send))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("how", ["block", "bfs"])
def test_two_ranks_equal_single_process_oracle(tmp_path, how):
    from oracle import oracle as O
    torch.set_num_threads(1)
    dag, nb = _dag()
    owner = block_partition(N, 2) if how == "block" else partition_nodes(nb, 2)
    assert sorted(np.bincount(owner).tolist()) == [4, 4]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, str(tmp_path), owner.tolist(), q, how == "bfs"))
             for r in range(2)]
    for p in procs:
        p.start()
    got = {}
    halo = []
    for _ in range(2):
        rank, out, hb, mh, ms = q.get(timeout=240)
        got.update(out)
        halo.append((hb, mh, ms))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    imgs, labs = synthetic.make_partitions(N, N_I, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(N_TEST, 32, 3, 10)
    sinks, _ = O.solve_dag(L, dag, imgs, labs, timg, tlab, seed=42, batch=32, local_steps=STEPS, lr=0.002, mu=0.9,
                           wd=0.0, n_i=N_I, minibatch_permutation=synthetic.minibatch_permutation)
    assert sorted(got) == sorted(sinks)
    for name, row in sinks.items():
        assert np.array_equal(got[name], row[:L.P]), name
    assert all(h[0] > 0 for h in halo)                 # rows really crossed ranks
    # what one rank sends in total is what the other receives
    assert halo[0][0] + halo[1][0] > 0


def _gossip_worker(rank, world, port, tmp, owner, n, events, out_q):
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from cpu_standin import OracleExecutor
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    dag = synthetic.gossip_dag(n, events, seed=3)
    imgs, labs = synthetic.make_partitions(n, N_I, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(N_TEST, 32, 3, 10)
    owned = [int(p) for p in np.nonzero(np.asarray(owner) == rank)[0]]
    st = SessionSettings(participants=n, seed=42, data_dir=tmp, learning=LearningSettings(0.002, 0.9, 0.0, 32, STEPS))
    ex = OracleExecutor(st, imgs, labs, timg, tlab, owned, N_I)
    solver = DistributedSolver(ex, dag, rank, world, owner)
    res = solver.run()                       # streamed: one level at a time, every rank in every exchange
    out = {name: h[0].executor.X[h[0].slot, :L.P].numpy().copy() for name, h in res.items()}
    n_xchg = sum(1 for w in solver.waves if getattr(w, "xchg", None))
    idle = sum(1 for w in solver.waves if getattr(w, "xchg", None)
               and sum(w.xchg["send_counts"]) + sum(w.xchg["recv_counts"]) == 0)
    out_q.put((rank, out, n_xchg, idle))
    dist.barrier()
    dist.destroy_process_group()


def test_three_ranks_irregular_gossip_graph(tmp_path):
    """World size 3 on an asynchronous gossip graph: at most levels only SOME ranks have a cut edge.  The
    exchange is a collective, so every rank must enter it at every level where any rank exchanges (a rank
    that skipped it would hang or desynchronise the group); results must equal the sequential oracle."""
    from oracle import oracle as O
    torch.set_num_threads(1)
    n, events, world = 6, 9, 3
    dag = synthetic.gossip_dag(n, events, seed=3)
    owner = block_partition(n, world)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gossip_worker, args=(r, world, port, str(tmp_path), owner.tolist(), n, events, q))
             for r in range(world)]
    for p in procs:
        p.start()
    got, n_xchg, idle = {}, [], 0
    for _ in range(world):
        rank, out, nx, nidle = q.get(timeout=300)
        got.update(out)
        n_xchg.append(nx)
        idle += nidle
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert len(set(n_xchg)) == 1 and n_xchg[0] > 0      # the same number of collectives on every rank
    assert idle > 0                                      # and some rank took part without rows of its own
    imgs, labs = synthetic.make_partitions(n, N_I, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(N_TEST, 32, 3, 10)
    sinks, _ = O.solve_dag(L, dag, imgs, labs, timg, tlab, seed=42, batch=32, local_steps=STEPS, lr=0.002, mu=0.9,
                           wd=0.0, n_i=N_I, minibatch_permutation=synthetic.minibatch_permutation)
    assert sorted(got) == sorted(sinks)
    for name, row in sinks.items():
        assert np.array_equal(got[name], row[:L.P]), name


def test_partition_ring_has_two_cut_edges_per_rank():
    """The reference's ring relabels nodes randomly (dpsgd/simulation.py:37-41); the BFS partition
    recovers contiguity: each rank needs exactly 2 remote rows."""
    import random
    n, world = 64, 4
    r = random.Random(42)
    nodes = list(range(n))
    mapping = dict(zip(nodes, r.sample(nodes, n)))
    nb = [[] for _ in range(n)]
    for i in range(n):
        a, b = mapping[i], mapping[(i + 1) % n]
        nb[a].append(b); nb[b].append(a)
    owner = partition_nodes(nb, world)
    assert np.bincount(owner).tolist() == [16] * 4
    for rank in range(world):
        remote = {j for i in np.nonzero(owner == rank)[0] for j in nb[i] if owner[j] != rank}
        assert len(remote) == 2
    blk = block_partition(n, world)
    remote_blk = {j for i in np.nonzero(blk == 0)[0] for j in nb[i] if blk[j] != 0}
    assert len(remote_blk) > 2


def test_halo_plan_is_consistent_across_ranks(tmp_path):
    """Without running anything: rank r's send counts to q equal q's receive counts from r, and the
    aggregate CSR refers to halo rows after the slot pool."""
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_solver_plan import PlanOnlyExecutor

    class PlanHalo(PlanOnlyExecutor):
        halo_rows = 0
        layout = L

        def _grow(self, extra, halo_rows=None, send_rows=0):
            super()._grow(extra)
            self.halo_rows = max(self.halo_rows, halo_rows or 0)

        @property
        def halo_base(self):
            return self.n_slots

        def _dev_i32(self, a):
            return np.asarray(a, dtype=np.int32)

    world = 4
    n = 32
    nb = synthetic.kreg_topology(n, 5 + 1, 42)
    dag = synthetic.dpsgd_dag(n, 2, nb)
    assert topology_from_dag(dag, n) == [list(x) for x in nb]
    owner = block_partition(n, world)
    solvers = []
    for rank in range(world):
        ex = PlanHalo(2 * 8 + 2, _settings(str(tmp_path)))
        solvers.append(DistributedSolver(ex, dag, rank, world, owner).plan())
    for s in solvers:
        assert [w.func for w in s.waves] == ["train", "aggregate"] * 2
    for lvl in (1, 3):
        x = [next(w for w in s.waves if w.level == lvl).xchg for s in solvers]
        for r in range(world):
            for q in range(world):
                assert x[r]["send_counts"][q] == x[q]["recv_counts"][r]
            assert x[r]["send_counts"][r] == 0
    for s in solvers:
        w = s.waves[1]
        base = s.ex.halo_base
        srcs = {m for ms in w.model_slots for m in ms}
        n_halo = sum(w.xchg["recv_counts"])
        assert {m for m in srcs if m >= base} == set(range(base, base + n_halo))
        assert s.ex.halo_rows >= n_halo
