"""Runs a few D-PSGD rounds (for ncu launch lists): python scripts/one_round.py NODES ROUNDS [tune_key tune_val]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dls_b200
from dls_b200 import _lib, synthetic
from dls_b200.dataset import DeviceDataset
from dls_b200.executor import WaveExecutor
from dls_b200.layout import GNLENET as L
from dls_b200.settings import SessionSettings, LearningSettings
from dls_b200.solver import BatchedSolver
n, rounds = int(sys.argv[1]), int(sys.argv[2])
if len(sys.argv) > 4:
    _lib.lib().dls_tune(int(sys.argv[3]), int(sys.argv[4]))
dev = torch.device("cuda:0")
st = SessionSettings(participants=n, seed=42, learning=LearningSettings(0.002, 0.9, 0.0, 32, 2), precision=1)
ds = DeviceDataset.synthetic(L, n, 64, 16, dev, seed=42, per_node_seed=True)
ex = WaveExecutor(st, ds, device=dev)
BatchedSolver(ex, synthetic.dpsgd_dag(n, rounds, synthetic.kreg_topology(n, 8, 42))).plan().run()
torch.cuda.synchronize()
print("done")
