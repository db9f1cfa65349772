"""Push transport of the multi-GPU solve on CPU: the ranks run as threads of one process over a stand-in for the
symmetric allocation (tests/cpu_standin.PushOracleExecutor), so the planner's halo slot numbering, the alternating
halo halves, the every-rank-enters-every-exchange rule and the transfer order are exercised without a GPU.  Rows and
arithmetic are the oracle's: the result must equal the single-process oracle solve bit for bit."""
import os
import sys
import threading

import numpy as np
import pytest
import torch

import dls_b200
from dls_b200 import synthetic
from dls_b200.layout import GNLENET as L
from dls_b200.partition import DistributedSolver, block_partition, partition_nodes
from dls_b200.settings import SessionSettings, LearningSettings

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from cpu_standin import PushOracleExecutor  # noqa: E402

N_I, STEPS, N_TEST = 32, 1, 32


def _solve_threads(dag, n, world, owner, tmp):
    imgs, labs = synthetic.make_partitions(n, N_I, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(N_TEST, 32, 3, 10)
    st = SessionSettings(participants=n, seed=42, data_dir=tmp, learning=LearningSettings(0.002, 0.9, 0.0, 32, STEPS))
    barrier = threading.Barrier(world)
    execs = []
    counts = np.bincount(owner, minlength=world)
    for r in range(world):
        owned = [int(p) for p in np.nonzero(np.asarray(owner) == r)[0]]
        ex = PushOracleExecutor(st, imgs, labs, timg, tlab, owned, N_I)
        ex.setup_push(pool_cap=4 * int(counts.max()) + 8, halo_cap=DistributedSolver.HALO_VERSIONS * int(n - counts.min()),
                      barrier=barrier)
        execs.append(ex)
    ptrs = [ex.X.data_ptr() for ex in execs]
    for ex in execs:
        ex.peer_x_ptrs = ptrs
    out, errs, solvers = {}, [], [None] * world

    def run(r):
        try:
            s = DistributedSolver(execs[r], dag, r, world, owner, transport="push")
            solvers[r] = s
            res = s.run()                                      # streamed, one level at a time
            for name, h in res.items():
                out[name] = h[0].executor.X[h[0].slot, :L.P].numpy().copy()
        except BaseException as e:                             # noqa: BLE001 — reported by the main thread
            errs.append((r, e))
            barrier.abort()

    threads = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=600)
    assert not errs, errs
    return out, execs, solvers, (imgs, labs, timg, tlab)


def _oracle(dag, data):
    from oracle import oracle as O
    imgs, labs, timg, tlab = data
    sinks, _ = O.solve_dag(L, dag, imgs, labs, timg, tlab, seed=42, batch=32, local_steps=STEPS, lr=0.002, mu=0.9, wd=0.0,
                           n_i=N_I, minibatch_permutation=synthetic.minibatch_permutation)
    return sinks


def test_dpsgd_four_ranks_push_equals_oracle(tmp_path):
    torch.set_num_threads(1)
    n, world, rounds = 8, 4, 3
    nb = synthetic.kreg_topology(n, 4, 42)
    dag = synthetic.dpsgd_dag(n, rounds, nb, test_interval=rounds, order_seed=5)
    owner = partition_nodes(nb, world)
    got, execs, solvers, data = _solve_threads(dag, n, world, owner, str(tmp_path))
    sinks = _oracle(dag, data)
    assert sorted(got) == sorted(sinks)
    for name, row in sinks.items():
        assert np.array_equal(got[name], row[:L.P]), name
    row_bytes = L.ld * 4
    for r, ex in enumerate(execs):
        assert ex.pushed_to, "rank %d pushed nothing" % r
        # every push landed inside a peer's halo region, never in a pool row and never in the sender's own tensor
        for d in ex.pushed_to:
            q = [p for p in range(world) if execs[p].X.data_ptr() <= d < execs[p].X.data_ptr() + execs[p].X.numel() * 4]
            assert len(q) == 1 and q[0] != r
            slot = (d - execs[q[0]].X.data_ptr()) // row_bytes
            assert slot >= execs[q[0]].pool_cap and (d - execs[q[0]].X.data_ptr()) % row_bytes == 0
    # transfer order of one exchange: destinations interleaved round-robin starting at rank + 1 — the first transfers
    # of a list that reaches several peers go to different peers
    for r, s in enumerate(solvers):
        for w in s.waves:
            x = getattr(w, "xchg", None)
            if not x or len(x["dst_ptrs"]) < 2:
                continue
            peers = [next(p for p in range(world) if execs[p].X.data_ptr() <= d < execs[p].X.data_ptr() + execs[p].X.numel() * 4)
                     for d in x["dst_ptrs"]]
            n_dst = len(set(peers))
            assert len(set(peers[:n_dst])) == n_dst, (r, peers)
            first = min(set(peers), key=lambda p: (p - r) % world)
            assert peers[0] == first


def test_irregular_gossip_three_ranks_push(tmp_path):
    """Asynchronous gossip graph, three ranks: most levels have a cut edge on SOME ranks only; every rank still runs
    the barrier of every exchange level (a rank that skipped one would leave the others waiting: the threading barrier
    of the stand-in times out)."""
    torch.set_num_threads(1)
    n, events, world = 6, 9, 3
    dag = synthetic.gossip_dag(n, events, seed=3)
    owner = block_partition(n, world)
    got, execs, solvers, data = _solve_threads(dag, n, world, owner, str(tmp_path))
    sinks = _oracle(dag, data)
    assert sorted(got) == sorted(sinks)
    for name, row in sinks.items():
        assert np.array_equal(got[name], row[:L.P]), name
    n_x = [sum(1 for w in s.waves if getattr(w, "xchg", None)) for s in solvers]
    assert len(set(n_x)) == 1 and n_x[0] > 0
