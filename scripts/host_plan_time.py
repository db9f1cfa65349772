"""Host-side cost of planning one D-PSGD round (no GPU needed): where the e2e leg's host time goes.
    python scripts/host_plan_time.py [nodes] [world] [--profile]"""
import sys, time, cProfile, pstats
sys.path.insert(0, ".")
import numpy as np, torch
import dls_b200
from dls_b200 import synthetic
from dls_b200.solver import BatchedSolver
from dls_b200.partition import DistributedSolver, partition_nodes
from dls_b200.executor import WaveExecutor
from dls_b200.settings import SessionSettings, LearningSettings

N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
WORLD = int(sys.argv[2]) if len(sys.argv) > 2 else 1
K = int(np.log2(N))
nbrs = synthetic.kreg_topology(N, K, 42)
owner = partition_nodes(nbrs, WORLD) if WORLD > 1 else np.zeros(N, np.int32)
owned = [int(x) for x in np.nonzero(owner == 0)[0]]
NL = len(owned)


class FakeDS:
    def minibatches(self, peers, rounds, batch, steps):
        return synthetic.minibatch_index(42, peers, rounds, 320, batch, max(steps))


class FakeEx(WaveExecutor):
    def __init__(self):
        self.settings = SessionSettings(participants=N, learning=LearningSettings(0.002, 0.9, 0.0, 32, 5))
        from dls_b200.layout import GNLENET
        self.layout = GNLENET
        self.device = torch.device("cpu"); self.dataset = FakeDS()
        self.n_slots = 2 * NL; self.free = list(range(2 * NL - 1, NL - 1, -1)); self.refs = np.zeros(2 * NL, dtype=np.int64)
        self.refs[:NL] = 1
        self.halo_rows = 0; self.sendbuf = None
        self.node_ids = owned; self.local_of = {p: i for i, p in enumerate(owned)}

    def _grow(self, extra, halo_rows=None, send_rows=0):
        self.n_slots += extra
        self.refs = np.concatenate([self.refs, np.zeros(extra, dtype=np.int64)])

    def _upload(self, *arrays):
        return [torch.from_numpy(np.ascontiguousarray(a).copy()) for a in arrays]

    def _agg_variant(self, *a):
        return 1


ex = FakeEx()
dags = [synthetic.dpsgd_dag(N, 1, nbrs, name_offset=i * N) for i in range(20)]


def one(i):
    if WORLD > 1:
        s = DistributedSolver(ex, dags[i], 0, WORLD, owner)
    else:
        s = BatchedSolver(ex, dags[i])
    s.seed_models({p: j for j, p in enumerate(owned)})
    s.plan()
    return s


one(0)
t0 = time.perf_counter()
for i in range(1, 11):
    one(i)
print("nodes %d world %d: plan ms/round: %.2f" % (N, WORLD, (time.perf_counter() - t0) * 100))
if "--profile" in sys.argv:
    pr = cProfile.Profile(); pr.enable()
    for i in range(11, 20):
        one(i)
    pr.disable(); pstats.Stats(pr).sort_stats("tottime").print_stats(22)
