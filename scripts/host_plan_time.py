"""Host-side cost of planning one D-PSGD round (no GPU needed): where the e2e leg's host time goes."""
import sys, time, cProfile, pstats
sys.path.insert(0, ".")
import numpy as np, torch
import dls_b200
from dls_b200 import synthetic
from dls_b200.solver import BatchedSolver
from dls_b200.executor import WaveExecutor
from dls_b200.settings import SessionSettings, LearningSettings

N, K = 256, 8
nbrs = synthetic.kreg_topology(N, K, 42)

class FakeDS:
    def minibatches(self, peers, rounds, batch, steps):
        return synthetic.minibatch_index(42, peers, rounds, 320, batch, max(steps))

class FakeEx(WaveExecutor):
    def __init__(self):
        self.settings = SessionSettings.from_dict({"dataset": "cifar10", "model": "gnlenet", "participants": N,
            "learning": {"batch_size": 32, "local_steps": 5, "learning_rate": 0.002, "momentum": 0.9, "weight_decay": 0}}) \
            if hasattr(SessionSettings, "from_dict") else None
        self.device = torch.device("cpu"); self.dataset = FakeDS()
        self.n_slots = 2 * N; self.free = list(range(N, 2 * N)); self.refs = np.zeros(2 * N, dtype=np.int64)
        self.refs[:N] = 1
    def _grow(self, *a, **k): pass

ex = FakeEx()
dags = [synthetic.dpsgd_dag(N, 1, nbrs, name_offset=i * N) for i in range(20)]
def one(i):
    s = BatchedSolver(ex, dags[i]); s.seed_models({p: p for p in range(N)}); s.plan(); return s
one(0)
t0 = time.perf_counter()
for i in range(1, 11): one(i)
print("plan ms/round:", (time.perf_counter() - t0) * 100)
pr = cProfile.Profile(); pr.enable()
for i in range(11, 20): one(i)
pr.disable(); pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
