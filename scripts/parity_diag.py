"""Prints the raw parity numbers of the C1 fixture comparison (used to set test tolerances)."""
import json, os, pickle, sys, tempfile
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import dls_b200
from dls_b200.dataset import DeviceDataset
from dls_b200.executor import WaveExecutor
from dls_b200.layout import GNLENET as L
from dls_b200.settings import SessionSettings, LearningSettings
from dls_b200.solver import BatchedSolver
G = os.path.join(REPO, "tests", "golden")
g = np.load(os.path.join(G, "dpsgd_c1.npz")); meta = json.loads(str(g["meta"]))
dag = pickle.load(open(os.path.join(G, "dpsgd_c1_dag.pkl"), "rb"))
rel = lambda a, b: float(np.linalg.norm(a.astype(np.float64) - b) / np.linalg.norm(b.astype(np.float64)))
cuda = torch.device("cuda:0")
for precision in (0, 1):
    d = tempfile.mkdtemp()
    st = SessionSettings(participants=16, seed=42, data_dir=d, precision=precision, learning=LearningSettings(0.002, 0.9, 0.0, 32, 5))
    ds = DeviceDataset.synthetic(L, 16, meta["n_i"], meta["n_test"], cuda, seed=42)
    ex = WaveExecutor(st, ds)
    res = BatchedSolver(ex, dag).plan().run()
    final = {int(n.split("_")[1]): h[0].numpy() for n, h in res.items()}
    rs = [rel(final[p][::16], g["final_strided"][p]) for p in range(16)]
    l2 = [abs(np.sqrt((final[p].astype(np.float64) ** 2).sum()) - g["final_l2"][p]) / g["final_l2"][p] for p in range(16)]
    acc = []
    for p in range(16):
        got = open(os.path.join(d, "accuracies_%d.csv" % p)).read().strip().split("\n")[1].split(",")
        want = str(g["acc_csv"][p]).strip().split("\n")[1].split(",")
        acc.append((float(got[7]), float(want[7]), float(got[8]), float(want[8])))
    print("precision", precision, "node0", rel(final[0], g["final_node0"]), "node11", rel(final[11], g["final_node11"]))
    print("  strided rel max %.3e mean %.3e; l2 rel max %.3e" % (max(rs), np.mean(rs), max(l2)))
    print("  acc diff per node", [round(a - b, 2) for a, b, _, _ in acc], "mean diff", np.mean([a - b for a, b, _, _ in acc]))
    print("  loss diff max", max(abs(c - d_) for _, _, c, d_ in acc))
