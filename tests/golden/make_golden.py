"""Generates the fixtures in this directory by running the UNMODIFIED reference package.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
It imports `dasklearn` from /root/reference behind the import shims in oracle/refshim (packages
that are not installed: dataclasses_json, flwr_datasets, librosa, matplotlib) and injects a
scripted synthetic dataset in place of the network-backed CIFAR-10 loader.  No reference file is
modified or copied.  Fixtures:
  fedavg_kat.npz   functions.aggregate on 5 / 11 GNLeNet models (default and explicit weights)
  init_gnlenet.npz flat initial model under torch.manual_seed(42)
  train_kat.npz    functions.train for one peer, two consecutive rounds (momentum carry-over,
                   partial last batch: n_i=80 -> steps of 32, 32, 16)
  dpsgd_mini.*     full DES -> DAG -> sequential solve: 8 nodes, 2-regular, 3 rounds, test at round 3
  dpsgd_c1.*       BASELINE config C1 solved by the reference: 16 nodes, 4-regular, 5 rounds of 5 local
                   steps (n_i = 320), test at round 5
  femnist_kat.npz  the reference's own FEMNIST `CNN` class (dasklearn/models/femnist.py:45-85, importable
                   once the missing module `dasklearn.datasets.femnist` is supplied as a two-constant shim):
                   init under manual_seed(42), logits, CrossEntropy loss, autograd gradients and three
                   steps of the reference's SGDOptimizer (lr 0.004, momentum 0) on scripted minibatches
  chunk_kat.npz    functions.chunk / functions.reconstruct_from_chunks on GNLeNet models (1..9 contributors)
"""
import asyncio
import hashlib
import json
import os
import pickle
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(REPO, "oracle", "refshim"))
sys.path.insert(1, "/root/reference")
sys.path.insert(2, REPO)
sys.path.insert(3, HERE)

import numpy as np
import torch

torch.set_num_threads(1)

import dasklearn.functions as RF                       # noqa: E402  (the reference)
from dasklearn.models import create_model              # noqa: E402
from dasklearn.session_settings import LearningSettings  # noqa: E402
from dasklearn.simulation.dpsgd.settings import DPSGDSettings  # noqa: E402
from dasklearn.simulation.dpsgd.simulation import DPSGDSimulation  # noqa: E402
from dasklearn.simulation.simulation import Simulation  # noqa: E402
from dasklearn.tasks.dag import WorkflowDAG            # noqa: E402

import dls_b200                                        # noqa: E402
from dls_b200 import synthetic                         # noqa: E402
from dls_b200.layout import GNLENET                    # noqa: E402
from katgen import kat_rows                            # noqa: E402

Simulation.plot_compute_graph = lambda self: None
SEED = 42


def flat(model) -> np.ndarray:
    return torch.cat([p.detach().reshape(-1) for p in model.parameters()]).numpy().astype(np.float32)


def model_from_flat(vec: np.ndarray):
    m = create_model("cifar10", architecture="gnlenet")
    off = 0
    with torch.no_grad():
        for p in m.parameters():
            n = p.numel()
            p.copy_(torch.from_numpy(vec[off:off + n].copy()).view_as(p))
            off += n
    assert off == GNLENET.P
    return m


# ------------------------------------------------------------------------------------------
# scripted dataset stand-in: __getitem__ ignores the sampler's index and returns the
# next entry of the harness-defined order, so the reference consumes the minibatches we script.
# ------------------------------------------------------------------------------------------
class ScriptedPartition:
    def __init__(self, images, labels):
        self.images, self.labels = images, labels
        self.fn = None
        self.script = []
        self.pos = 0
        self.fetched = 0

    def with_transform(self, fn):
        self.fn = fn
        return self

    def set_script(self, order):
        self.script = list(order)
        self.pos = 0

    def __len__(self):
        return len(self.labels)

    def __bool__(self):
        return True

    def __getitem__(self, i):
        sid = self.script[self.pos % len(self.script)]
        self.pos += 1
        self.fetched += 1
        batch = {"img": [self.images[sid]], "label": [int(self.labels[sid])]}
        batch = self.fn(batch)
        return {"img": batch["img"][0], "label": batch["label"][0]}


class SequentialSplit(ScriptedPartition):
    def __getitem__(self, i):
        batch = {"img": [self.images[i]], "label": [int(self.labels[i])]}
        batch = self.fn(batch)
        return {"img": batch["img"][0], "label": batch["label"][0]}


class SyntheticFederatedDataset:
    validationset = []

    def __init__(self, part_images, part_labels, test_images, test_labels):
        self.parts = [ScriptedPartition(part_images[i], part_labels[i]) for i in range(len(part_images))]
        self.test = SequentialSplit(test_images, test_labels)

    def load_partition(self, idx, split):
        assert split == "train"
        return self.parts[idx]

    def load_split(self, split):
        assert split == "test"
        return self.test


def make_settings(n, rounds, k, work_dir, batch=32, local_steps=5, test_interval=0):
    learning = LearningSettings(learning_rate=0.002, momentum=0.9, weight_decay=0, batch_size=batch,
                                local_steps=local_steps)
    return DPSGDSettings(algorithm="dpsgd", seed=SEED, work_dir=work_dir, dataset="cifar10", learning=learning,
                         participants=n, rounds=rounds, dry_run=True, brokers=1, unit_testing=True,
                         model="gnlenet", synchronous=True, topology="kreg", k=k, test_interval=test_interval,
                         duration=10 ** 15, profile=False, torch_threads=1)


def reset_reference_globals(dataset):
    RF.dataset = dataset
    RF.model_trainers = None
    RF.evaluator = None


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def gen_fedavg(work_dir):
    settings = make_settings(16, 1, 4, work_dir)
    out = {}
    for tag, fan, weights in (("f5", 5, None), ("f11", 11, None),
                              ("f2w", 2, [3.0 / 7.0, 4.0 / 7.0]),
                              ("f5w", 5, [0.1, 0.2, 0.3, 0.15, 0.25])):
        X = kat_rows(fan, GNLENET.P, salt=fan + (0 if weights is None else 100))
        models = [model_from_flat(X[i]) for i in range(fan)]
        params = {"models": models, "round": 1, "peer": 0}
        if weights is not None:
            params["weights"] = weights
        res = RF.aggregate(settings, params)
        assert isinstance(res, list) and len(res) == 1
        out[tag + "_y"] = flat(res[0])
        out[tag + "_in_sha"] = np.frombuffer(bytes.fromhex(sha(X)), dtype=np.uint8)
        if weights is not None:
            out[tag + "_w"] = np.array(weights, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "fedavg_kat.npz"), **out)
    print("fedavg_kat.npz written")


def gen_init():
    torch.manual_seed(SEED)                       # functions.py:46-48
    m = create_model("cifar10", architecture="gnlenet")
    v = flat(m)
    np.savez_compressed(os.path.join(HERE, "init_gnlenet.npz"), row=v,
                        torch_version=np.array(torch.__version__))
    print("init_gnlenet.npz written", v.shape)
    return v


def gen_train(work_dir):
    n_nodes, n_i, peer = 4, 80, 3
    imgs, labs = synthetic.make_partitions(n_nodes, n_i, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(64, 32, 3, 10)
    ds = SyntheticFederatedDataset(imgs, labs, timg, tlab)
    reset_reference_globals(ds)
    settings = make_settings(n_nodes, 2, 2, work_dir)
    settings.data_dir = os.path.join(work_dir, "train_kat")
    os.makedirs(settings.data_dir, exist_ok=True)
    rows = []
    model = (None, 0)
    for rnd in (1, 2):
        ds.parts[peer].set_script(synthetic.minibatch_permutation(SEED, peer, rnd, n_i))
        res = RF.train(settings, {"model": model, "round": rnd, "time": 0, "peer": peer})
        model = res[0]
        rows.append(flat(model))
    np.savez_compressed(os.path.join(HERE, "train_kat.npz"), after_round1=rows[0], after_round2=rows[1],
                        meta=np.array(json.dumps({"n_nodes": n_nodes, "n_i": n_i, "peer": peer, "seed": SEED,
                                                  "batch": 32, "local_steps": 5, "lr": 0.002, "momentum": 0.9,
                                                  "wd": 0.0, "torch": torch.__version__, "threads": 1,
                                                  "fetched": ds.parts[peer].fetched})))
    print("train_kat.npz written; __getitem__ calls:", ds.parts[peer].fetched)


def gen_gradient(work_dir):
    """AD-PSGD entry points (functions.py:80-86): `compute_gradient` = forward/backward on the first
    minibatch of a fresh loader, no optimizer step, gradient attached to a copy of the unchanged model;
    `gradient_update` = ONE step of a FRESH SGD optimizer (momentum buffer = g) with a supplied gradient."""
    n_nodes, n_i, peer = 4, 80, 2
    imgs, labs = synthetic.make_partitions(n_nodes, n_i, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(64, 32, 3, 10)
    ds = SyntheticFederatedDataset(imgs, labs, timg, tlab)
    reset_reference_globals(ds)
    settings = make_settings(n_nodes, 2, 2, work_dir)
    settings.data_dir = os.path.join(work_dir, "grad_kat")
    os.makedirs(settings.data_dir, exist_ok=True)
    ds.parts[peer].set_script(synthetic.minibatch_permutation(SEED, peer, 1, n_i))
    gm = RF.compute_gradient(settings, {"model": (None, 0), "round": 1, "time": 0, "peer": peer})[0]
    grad1 = np.concatenate([g.detach().cpu().numpy().reshape(-1) for g in gm.gradient]).astype(np.float32)
    model_after_cg = flat(gm)
    upd1 = flat(RF.gradient_update(settings, {"model": gm, "gradient_model": gm, "round": 1, "time": 0,
                                              "peer": peer})[0])
    # a stale gradient applied to a different (trained) model, then a second gradient at that model
    ds.parts[peer].set_script(synthetic.minibatch_permutation(SEED, peer, 2, n_i))
    trained = RF.train(settings, {"model": model_from_flat(upd1), "round": 2, "time": 0, "peer": peer})[0]
    trained_flat = flat(trained)
    gm_stale = model_from_flat(model_after_cg)
    gm_stale.gradient = [g.clone() for g in gm.gradient]
    upd2 = flat(RF.gradient_update(settings, {"model": trained, "gradient_model": gm_stale, "round": 2, "time": 0,
                                              "peer": peer})[0])
    ds.parts[peer].set_script(synthetic.minibatch_permutation(SEED, peer, 3, n_i))
    gm2 = RF.compute_gradient(settings, {"model": model_from_flat(upd2), "round": 3, "time": 0, "peer": peer})[0]
    grad2 = np.concatenate([g.detach().cpu().numpy().reshape(-1) for g in gm2.gradient]).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "gradient_kat.npz"), grad1=grad1, model_after_cg=model_after_cg, upd1=upd1,
                        trained=trained_flat,
                        upd2=upd2, grad2=grad2,
                        meta=np.array(json.dumps({"n_nodes": n_nodes, "n_i": n_i, "peer": peer, "seed": SEED, "batch": 32,
                                                  "lr": 0.002, "momentum": 0.9, "wd": 0.0, "torch": torch.__version__})))
    print("gradient_kat.npz written; |g1| = %.6f, |g2| = %.6f" % (np.linalg.norm(grad1), np.linalg.norm(grad2)))


def run_des(settings):
    loop = asyncio.new_event_loop()
    asyncio.set_event_loop(loop)
    sim = DPSGDSimulation(settings)
    loop.run_until_complete(sim.run())
    return sim


def solve_sequential(serialized, settings, ds, n_i):
    """In-process '1 broker / 1 worker' solve mirroring Broker.handle_task_result
    (dasklearn/broker.py:277-290) and Worker.start (dasklearn/worker.py:21-38)."""
    dag = WorkflowDAG.unserialize(pickle.loads(pickle.dumps(serialized)))
    queue = list(dag.get_source_tasks())
    results = {}
    while queue:
        task = queue.pop(0)
        if task.func == "train":
            peer, rnd = task.data["peer"], task.data["round"]
            ds.parts[peer].set_script(synthetic.minibatch_permutation(SEED, peer, rnd, n_i))
        res = getattr(RF, task.func)(settings, task.data)
        assert isinstance(res, (list, tuple))
        if not task.outputs:
            results[task.name] = res
        for nxt, idx in task.outputs:
            nxt.set_data((task.name, idx), res[idx])
            if nxt.has_all_inputs():
                queue.append(nxt)
        task.clear_data()
    return results


def gen_dpsgd_mini(work_dir):
    n, k, rounds, n_i, steps, n_test = 8, 2, 3, 64, 2, 256
    imgs, labs = synthetic.make_partitions(n, n_i, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(n_test, 32, 3, 10)
    ds = SyntheticFederatedDataset(imgs, labs, timg, tlab)
    reset_reference_globals(ds)
    settings = make_settings(n, rounds, k, work_dir, local_steps=steps, test_interval=rounds)
    sim = run_des(settings)
    serialized = sim.workflow_dag.serialize()
    os.makedirs(settings.data_dir, exist_ok=True)
    sinks = solve_sequential(serialized, settings, ds, n_i)
    final = np.zeros((n, GNLENET.P), dtype=np.float32)
    for name, res in sinks.items():
        peer = int(name.split("_")[1])
        final[peer] = flat(res[0])
    acc_rows = []
    for peer in range(n):
        with open(os.path.join(settings.data_dir, "accuracies_%d.csv" % peer)) as f:
            acc_rows.append(f.read())
    with open(os.path.join(HERE, "dpsgd_mini_dag.pkl"), "wb") as f:
        pickle.dump(serialized, f, protocol=4)
    np.savez_compressed(os.path.join(HERE, "dpsgd_mini.npz"),
                        final_node0=final[0], final_node5=final[5],
                        final_sum=final.astype(np.float64).sum(axis=1),
                        final_l2=np.sqrt((final.astype(np.float64) ** 2).sum(axis=1)),
                        final_strided=final[:, ::64].copy(),
                        acc_csv=np.array(acc_rows),
                        meta=np.array(json.dumps({"n": n, "k": k, "rounds": rounds, "n_i": n_i, "local_steps": steps,
                                                  "n_test": n_test, "seed": SEED, "batch": 32, "lr": 0.002,
                                                  "momentum": 0.9, "wd": 0.0, "torch": torch.__version__,
                                                  "threads": 1, "n_tasks": len(serialized)})))
    print("dpsgd_mini written: %d tasks" % len(serialized))
    print(acc_rows[0])


def gen_c1(work_dir):
    """BASELINE config C1 (37 s on CPU): D-PSGD, 16 nodes, 4-regular, 5 rounds, test at 5."""
    n, k, rounds, n_i, steps, n_test = 16, 4, 5, 320, 5, 256
    imgs, labs = synthetic.make_partitions(n, n_i, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(n_test, 32, 3, 10)
    ds = SyntheticFederatedDataset(imgs, labs, timg, tlab)
    reset_reference_globals(ds)
    settings = make_settings(n, rounds, k, work_dir, local_steps=steps, test_interval=rounds)
    sim = run_des(settings)
    serialized = sim.workflow_dag.serialize()
    os.makedirs(settings.data_dir, exist_ok=True)
    sinks = solve_sequential(serialized, settings, ds, n_i)
    final = np.zeros((n, GNLENET.P), dtype=np.float32)
    for name, res in sinks.items():
        peer = int(name.split("_")[1])
        final[peer] = flat(res[0])
    acc_rows = []
    for peer in range(n):
        with open(os.path.join(settings.data_dir, "accuracies_%d.csv" % peer)) as f:
            acc_rows.append(f.read())
    with open(os.path.join(HERE, "dpsgd_c1_dag.pkl"), "wb") as f:
        pickle.dump(serialized, f, protocol=4)
    np.savez_compressed(os.path.join(HERE, "dpsgd_c1.npz"),
                        final_node0=final[0], final_node11=final[11],
                        final_sum=final.astype(np.float64).sum(axis=1),
                        final_l2=np.sqrt((final.astype(np.float64) ** 2).sum(axis=1)),
                        final_strided=final[:, ::16].copy(),
                        acc_csv=np.array(acc_rows),
                        meta=np.array(json.dumps({"n": n, "k": k, "rounds": rounds, "n_i": n_i, "local_steps": steps,
                                                  "n_test": n_test, "seed": SEED, "batch": 32, "lr": 0.002,
                                                  "momentum": 0.9, "wd": 0.0, "torch": torch.__version__,
                                                  "threads": 1, "n_tasks": len(serialized)})))
    print("dpsgd_c1 written: %d tasks" % len(serialized))
    print(acc_rows[0])


def gen_femnist():
    """Pin the FEMNIST CNN to the reference's OWN model class.  dasklearn/models/femnist.py:5 imports
    `dasklearn.datasets.femnist`, a module that is not in the reference tree; supplying it as a shim with the
    two constants it is asked for makes `create_model("femnist")` work unmodified.  The
    reference's trainer refuses the dataset (model_trainer.py:77-78), so the loop around the model is
    model_trainer.py:99-126 driven by hand: CrossEntropyLoss (:117-118) + the reference's SGDOptimizer."""
    import types
    from dls_b200.layout import FEMNIST
    shim = types.ModuleType("dasklearn.datasets.femnist")
    shim.FLAT_SIZE, shim.NUM_CLASSES = 28 * 28, 62
    sys.modules["dasklearn.datasets.femnist"] = shim
    from dasklearn.optimizer.sgd import SGDOptimizer
    torch.manual_seed(SEED)
    m = create_model("femnist")
    assert type(m).__module__ == "dasklearn.models.femnist" and type(m).__name__ == "CNN"
    init = flat(m)
    assert init.size == FEMNIST.P
    n_i, B = 96, 32
    imgs, labs = synthetic.make_images(n_i, 28, 1, 62, seed=1234 + 7)
    order = synthetic.minibatch_permutation(SEED, 7, 1, n_i)
    x_all = torch.from_numpy(imgs).permute(0, 3, 1, 2).contiguous().to(torch.float32) / 255.0   # ToTensor
    y_all = torch.from_numpy(labs.astype(np.int64))
    sel = order[:B]
    out = m(x_all[sel])
    loss = torch.nn.CrossEntropyLoss()(out, y_all[sel])
    m.zero_grad()
    loss.backward()
    grad = torch.cat([p.grad.reshape(-1) for p in m.parameters()]).numpy().astype(np.float32)
    # three optimizer steps (lr 0.004, momentum 0: args.py:70-72), last batch partial (96 = 32 + 32 + 32 -> use 32,32,20)
    opt = SGDOptimizer(m, 0.004, 0.0, 0.0)
    counts = [32, 32, 20]
    losses = []
    for s, c in enumerate(counts):
        b = order[s * B: s * B + c]
        m.train()
        opt.optimizer.zero_grad()
        l = torch.nn.CrossEntropyLoss()(m.forward(x_all[b]), y_all[b])
        l.backward()
        opt.optimizer.step()
        losses.append(float(l.detach()))
    after = flat(m)
    with torch.no_grad():
        logits_after = m(x_all[order[:64]]).numpy()
    ent = {e[0]: (e[2], int(np.prod(e[1]))) for e in FEMNIST.entries}
    o1, n1 = ent["fc1.weight"]
    keep = np.ones(FEMNIST.P, dtype=bool)
    keep[o1:o1 + n1] = False
    keep[o1:o1 + n1:16] = True                                   # fc1.weight (1.6 M values) is kept 1 in 16
    np.savez_compressed(os.path.join(HERE, "femnist_kat.npz"),
                        init_sha=np.frombuffer(bytes.fromhex(sha(init)), dtype=np.uint8),
                        init_strided=init[::64].copy(), logits=out.detach().numpy(), loss=np.float32(loss.item()),
                        grad_kept=grad[keep], grad_l2=np.float64(np.linalg.norm(grad.astype(np.float64))),
                        after_kept=after[keep], after_sum=np.float64(after.astype(np.float64).sum()),
                        step_losses=np.array(losses, dtype=np.float64), logits_after=logits_after,
                        meta=np.array(json.dumps({"n_i": n_i, "batch": B, "peer": 7, "round": 1, "seed": SEED,
                                                  "counts": counts, "lr": 0.004, "momentum": 0.0, "wd": 0.0,
                                                  "image_seed": 1234 + 7, "fc1_stride": 16,
                                                  "torch": torch.__version__, "threads": 1})))
    print("femnist_kat.npz written; loss %.6f, |g| %.6f" % (loss.item(), np.linalg.norm(grad)))


def gen_chunks(work_dir):
    """Shatter/Conflux chunk ops (functions.py:136-146 -> ChunkManager, simulation/conflux/chunk_manager.py:12-53):
    `chunk` slices the flat state_dict into n pieces (the last takes the remainder); `reconstruct_from_chunks`
    averages, per chunk index, the pieces of a VARIABLE set of contributors with torch.mean(torch.stack(..), 0)."""
    settings = make_settings(16, 1, 4, work_dir)
    n_chunks = 6
    contributors = [1, 2, 3, 4, 5, 9]                 # contributor count per chunk index
    n_models = max(contributors)
    X = kat_rows(n_models, GNLENET.P, salt=777)
    models = [model_from_flat(X[i]) for i in range(n_models)]
    pieces = [RF.chunk(settings, {"model": mm, "n": n_chunks}) for mm in models]
    sizes = [int(c.numel()) for c in pieces[0]]
    # contributor sets: chunk c is averaged over models (c, c+1, ...) mod n_models, in that order
    sets = [[(c + j) % n_models for j in range(contributors[c])] for c in range(n_chunks)]
    chunks = [[pieces[mi][c] for mi in sets[c]] for c in range(n_chunks)]
    res = RF.reconstruct_from_chunks(settings, {"chunks": chunks})
    assert isinstance(res, list) and len(res) == 1
    np.savez_compressed(os.path.join(HERE, "chunk_kat.npz"), y=flat(res[0]), sizes=np.array(sizes, dtype=np.int64),
                        sets=np.array(json.dumps(sets)), in_sha=np.frombuffer(bytes.fromhex(sha(X)), dtype=np.uint8),
                        piece0_of_model1=pieces[1][0].numpy().copy(),
                        meta=np.array(json.dumps({"n_chunks": n_chunks, "n_models": n_models, "salt": 777,
                                                  "torch": torch.__version__, "threads": 1})))
    print("chunk_kat.npz written; chunk sizes", sizes)


def gen_c2_dag(work_dir):
    """The reference DES's own compute graph for BASELINE config C2 (256 nodes, k = log2 N = 8),
    2 rounds, so the bench can replay a real coordinator-produced DAG."""
    import gzip
    settings = make_settings(256, 2, 0, work_dir, local_steps=5, test_interval=0)
    sim = run_des(settings)
    serialized = sim.workflow_dag.serialize()
    with gzip.open(os.path.join(HERE, "dpsgd_c2_dag.pkl.gz"), "wb") as f:
        pickle.dump(serialized, f, protocol=4)
    print("C2 dag: %d tasks" % len(serialized))


if __name__ == "__main__":
    import tempfile
    wd = tempfile.mkdtemp(prefix="golden_")
    if len(sys.argv) > 1:                                         # only the named fixtures
        for what in sys.argv[1:]:
            {"gradient": lambda: gen_gradient(wd), "c1": lambda: gen_c1(wd), "femnist": gen_femnist,
             "chunks": lambda: gen_chunks(wd), "fedavg": lambda: gen_fedavg(wd), "train": lambda: gen_train(wd),
             "mini": lambda: gen_dpsgd_mini(wd), "c2dag": lambda: gen_c2_dag(wd), "init": gen_init}[what]()
        sys.exit(0)
    gen_fedavg(wd)
    gen_init()
    gen_train(wd)
    gen_gradient(wd)
    gen_dpsgd_mini(wd)
    gen_c2_dag(wd)
    gen_c1(wd)
    gen_femnist()
    gen_chunks(wd)
