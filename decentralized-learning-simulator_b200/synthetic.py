"""Synthetic inputs for the compute-graph solve: images, the minibatch index, topologies and
level-synchronous D-PSGD compute graphs in the reference's serialized task format.

The reference draws real CIFAR-10 partitions through flwr_datasets (network) and a fresh shuffled
`DataLoader` per train call (`dasklearn/model_trainer.py:87`).  Here every node owns a fixed
uint8 HWC partition and each (node, round) gets a scripted permutation, so the CUDA path and
the CPU oracle consume identical minibatches.
"""
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np


def make_images(n: int, hw: int, c: int, num_classes: int, seed: int = 1234, mix: float = 0.6
                ) -> Tuple[np.ndarray, np.ndarray]:
    """uint8 HWC images `[n, hw, hw, c]` and int64 labels; class prototype blended with noise."""
    rs = np.random.RandomState(seed)
    protos = rs.randint(0, 256, size=(num_classes, hw, hw, c)).astype(np.float32)
    labels = rs.randint(0, num_classes, size=(n,)).astype(np.int64)
    noise = rs.randint(0, 256, size=(n, hw, hw, c)).astype(np.float32)
    img = np.clip(mix * protos[labels] + (1.0 - mix) * noise, 0, 255).astype(np.uint8)
    return img, labels


def make_partitions(n_nodes: int, n_i: int, hw: int, c: int, num_classes: int, seed: int = 1234,
                    mix: float = 0.6) -> Tuple[np.ndarray, np.ndarray]:
    """Per-node train partitions: images `[n_nodes, n_i, hw, hw, c]` uint8, labels `[n_nodes, n_i]`.
    Train and test share prototypes because both come from `make_images` with the same seed
    stream prefix (prototypes are drawn first)."""
    img, lab = make_images(n_nodes * n_i, hw, c, num_classes, seed, mix)
    return img.reshape(n_nodes, n_i, hw, hw, c), lab.reshape(n_nodes, n_i)


def make_test_split(n_test: int, hw: int, c: int, num_classes: int, seed: int = 1234,
                    mix: float = 0.6) -> Tuple[np.ndarray, np.ndarray]:
    rs = np.random.RandomState(seed)
    protos = rs.randint(0, 256, size=(num_classes, hw, hw, c)).astype(np.float32)
    rs2 = np.random.RandomState(seed + 7919)
    labels = rs2.randint(0, num_classes, size=(n_test,)).astype(np.int64)
    noise = rs2.randint(0, 256, size=(n_test, hw, hw, c)).astype(np.float32)
    img = np.clip(mix * protos[labels] + (1.0 - mix) * noise, 0, 255).astype(np.uint8)
    return img, labels


_U30, _U27, _U31 = np.uint64(30), np.uint64(27), np.uint64(31)
_M1, _M2 = np.uint64(0xBF58476D1CE4E5B9), np.uint64(0x94D049BB133111EB)


def _mix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser (vectorised, wraps modulo 2^64)."""
    x = (x ^ (x >> _U30)) * _M1
    x = (x ^ (x >> _U27)) * _M2
    return x ^ (x >> _U31)


def minibatch_permutations(seed: int, nodes: Sequence[int], rounds: Sequence[int], n_i: int) -> np.ndarray:
    """`perm[t, :]` = the scripted sample order node `nodes[t]` sees in round `rounds[t]` (stand-in
    for the `shuffle=True` permutation of `dasklearn/model_trainer.py:87`): samples sorted by a
    counter-based hash of (seed, node, round, sample).  Platform independent and vectorised over
    all tasks of a wave (one argsort), so the host can script thousands of nodes per round."""
    with np.errstate(over="ignore"):
        nodes = np.asarray(nodes, dtype=np.uint64).reshape(-1, 1)
        rounds = np.asarray(rounds, dtype=np.uint64).reshape(-1, 1)
        base = _mix64(np.uint64(seed) * np.uint64(0x9E3779B97F4A7C15) + nodes * np.uint64(0xD1B54A32D192ED03)
                      + rounds * np.uint64(0x8CB92BA72F3D8DD7) + np.uint64(0x632BE59BD9B4E019))
        idx = np.arange(n_i, dtype=np.uint64).reshape(1, -1)
        keys = _mix64(base + idx * np.uint64(0x9E3779B97F4A7C15))
        # the sample index in the low 24 bits makes every key unique, so the order does not depend on
        # the sort algorithm, and sorting the keys themselves (cheaper than an argsort) yields the
        # permutation in their low bits
        keys &= np.uint64(0xFFFFFFFFFF000000)
        keys |= idx
        keys.sort(axis=1)
        keys &= np.uint64(0xFFFFFF)
    return keys.astype(np.int32)


def minibatch_permutation(seed: int, node: int, round_nr: int, n_i: int) -> np.ndarray:
    return minibatch_permutations(seed, [node], [round_nr], n_i)[0]


def minibatch_index(seed: int, nodes: Sequence[int], rounds: Sequence[int], n_i: int, batch: int,
                    local_steps: int) -> Tuple[np.ndarray, np.ndarray]:
    """`idx[t, step, slot]` (sample id inside the node's partition, -1 = empty slot) and
    `count[t, step]`.  Steps executed = min(local_steps, ceil(n_i / batch)) and the last batch may
    be partial, exactly like the reference loop (`dasklearn/model_trainer.py:99-102`)."""
    T = len(nodes)
    perm = minibatch_permutations(seed, nodes, rounds, n_i) if T else np.zeros((0, n_i), np.int32)
    padded = -np.ones((T, local_steps * batch), dtype=np.int32)
    m = min(n_i, local_steps * batch)
    padded[:, :m] = perm[:, :m]
    idx = padded.reshape(T, local_steps, batch)
    cnt = (idx >= 0).sum(axis=2).astype(np.int32)
    return idx, cnt


# ---------------------------------------------------------------------------------------------
# Topologies and compute graphs
# ---------------------------------------------------------------------------------------------

def kreg_topology(n: int, k: int, seed: int) -> List[List[int]]:
    """In-neighbour lists of the `kreg` topology (`dasklearn/simulation/dpsgd/simulation.py:35-36`:
    `nx.random_regular_graph(k, N, seed).to_directed()`)."""
    import networkx as nx
    g = nx.random_regular_graph(k, n, seed=seed).to_directed()
    return [sorted(g.predecessors(i)) for i in range(n)]


def ring_topology(n: int) -> List[List[int]]:
    return [[(i - 1) % n, (i + 1) % n] for i in range(n)]


def exponential_topology(n: int) -> List[List[int]]:
    """Directed exponential graph i -> (i + 2^j) mod n; not in the reference, provided
    as the C5 extension.  Returns in-neighbour lists."""
    preds: List[List[int]] = [[] for _ in range(n)]
    j = 1
    while j < n:
        for i in range(n):
            d = (i + j) % n
            if i not in preds[d] and d != i:
                preds[d].append(i)
        j *= 2
    return preds


def dpsgd_dag(n: int, rounds: int, in_neighbours: List[List[int]], test_interval: int = 0,
              name_offset: int = 0, order_seed: Optional[int] = None) -> List[Dict]:
    """Serialized D-PSGD compute graph (list of task dicts, the `WorkflowDAG.serialize()` format of
    `dasklearn/tasks/dag.py:28-32` / `dasklearn/tasks/task.py:59-66`) for a fixed topology with
    every node online: per round N `train` tasks then N `aggregate` tasks whose `models` list is
    in-neighbours (arrival order) followed by the node's own model
    (`dasklearn/simulation/dpsgd/client.py:145`), plus `test_<peer>_<round>` pass-through tasks
    every `test_interval` rounds (:153-159).  Arrival order is the sorted neighbour order unless
    `order_seed` shuffles it per aggregate."""
    rs = np.random.RandomState(order_seed) if order_seed is not None else None
    tasks: Dict[str, Dict] = {}
    counter = {"train": name_offset, "agg": name_offset}
    outputs: Dict[str, List[Tuple[str, int]]] = {}

    def add(name, func, data, inputs):
        tasks[name] = {"name": name, "func": func, "data": data,
                       "inputs": list(inputs), "outputs": []}
        for src, idx in inputs:
            tasks[src]["outputs"].append((name, idx))

    model_of: List[Tuple[Optional[str], int]] = [(None, 0)] * n
    t_us = 0
    for r in range(1, rounds + 1):
        trained: List[Tuple[str, int]] = []
        for peer in range(n):
            counter["train"] += 1
            name = "train_%d" % counter["train"]
            prev = model_of[peer]
            add(name, "train", {"model": prev, "round": r, "time": t_us, "peer": peer},
                [] if prev[0] is None else [prev])
            trained.append((name, 0))
        t_us += 3_000_000
        for peer in range(n):
            nbs = list(in_neighbours[peer])
            if rs is not None:
                rs.shuffle(nbs)
            models = [trained[j] for j in nbs] + [trained[peer]]
            counter["agg"] += 1
            name = "agg_%d" % counter["agg"]
            add(name, "aggregate", {"models": list(models), "round": r, "peer": peer}, models)
            model_of[peer] = (name, 0)
            if test_interval > 0 and r % test_interval == 0:
                tname = "test_%d_%d" % (peer, r)
                add(tname, "test", {"model": (name, 0), "round": r, "time": t_us, "peer": peer},
                    [(name, 0)])
                model_of[peer] = (tname, 0)
        t_us += 1_000_000
    return list(tasks.values())


def gossip_dag(n: int, events: int, seed: int = 0) -> List[Dict]:
    """An asynchronous, gossip-style compute graph in the reference's task format: every event one
    node trains on its current model and pushes the result to a random peer, which averages it with
    its own model using age-proportional `weights` (the shape `AsynchronousClient.aggregate` builds,
    dasklearn/simulation/asynchronous_client.py:53-74).  Wavefronts mix train and aggregate tasks,
    fan-in is 2, and a trained model has two consumers (the peer's aggregate and the node's own
    next train), which exercises the solver's copy-versus-in-place logic."""
    rs = np.random.RandomState(seed)
    tasks: Dict[str, Dict] = {}

    def add(name, func, data, inputs):
        tasks[name] = {"name": name, "func": func, "data": data, "inputs": list(inputs), "outputs": []}
        for src, idx in inputs:
            tasks[src]["outputs"].append((name, idx))

    model_of: List[Tuple[Optional[str], int]] = [(None, 0)] * n
    age = [0] * n
    trains = [0] * n
    n_train = n_agg = 0
    order = list(range(n)) + [int(rs.randint(n)) for _ in range(events)]      # everybody trains once first
    for e, i in enumerate(order):
        n_train += 1
        trains[i] += 1
        name = "train_%d" % n_train
        prev = model_of[i]
        add(name, "train", {"model": prev, "round": trains[i], "time": e * 1000, "peer": i},
            [] if prev[0] is None else [prev])
        model_of[i] = (name, 0)
        age[i] += 1
        if e < n:
            continue
        j = int(rs.randint(n - 1))
        j = j if j < i else j + 1
        n_agg += 1
        aname = "agg_%d" % n_agg
        models = [model_of[i], model_of[j]]
        tot = float(age[i] + age[j])
        add(aname, "aggregate", {"models": list(models), "round": max(age[i], age[j]), "peer": j,
                                 "weights": [age[i] / tot, age[j] / tot]}, models)
        model_of[j] = (aname, 0)
        age[j] = max(age[i], age[j])
    return list(tasks.values())


def adpsgd_dag(n_pairs: int, rounds: int) -> List[Dict]:
    """An AD-PSGD-shaped compute graph (dasklearn/simulation/adpsgd/client.py:32-110) in the reference's
    task format: peers 2i are active, peers 2i+1 passive.  Every round a passive peer starts a
    `compute_gradient` task on its model, exchanges models with its active partner while that gradient is
    in flight (two fan-in-2 `aggregate` tasks), and then applies the now STALE gradient to the averaged
    model with `gradient_update`; the active peer runs a `train` task on its averaged model.  As in the
    reference, `gradient_model` is a bare task name while `model` is a (task name, index) placeholder
    (simulation/client.py:106-115), and a `gradient_update` task has two input edges."""
    tasks: Dict[str, Dict] = {}

    def add(name, func, data, inputs):
        tasks[name] = {"name": name, "func": func, "data": data, "inputs": list(inputs), "outputs": []}
        for src, idx in inputs:
            tasks[src]["outputs"].append((name, idx))

    n = 2 * n_pairs
    own: List[Tuple[str, int]] = []
    cnt = {"train": 0, "cg": 0, "gu": 0, "agg": 0}

    def name_of(kind, prefix):
        cnt[kind] += 1
        return "%s_%d" % (prefix, cnt[kind])

    for i in range(n):                                         # everybody trains once from the initial model
        t = name_of("train", "train")
        add(t, "train", {"model": (None, 0), "round": 1, "time": 0, "peer": i}, [])
        own.append((t, 0))
    for r in range(2, rounds + 2):
        for q in range(n_pairs):
            a, p_ = 2 * q, 2 * q + 1
            cg = name_of("cg", "compute_gradient")
            add(cg, "compute_gradient", {"model": own[p_], "round": r, "time": r * 1000, "peer": p_, "local_steps": 1},
                [own[p_]])
            agg_p, agg_a = name_of("agg", "agg"), name_of("agg", "agg")
            add(agg_p, "aggregate", {"models": [own[a], own[p_]], "round": r, "peer": p_, "weights": [0.5, 0.5]},
                [own[a], own[p_]])
            add(agg_a, "aggregate", {"models": [own[p_], own[a]], "round": r, "peer": a, "weights": [0.25, 0.75]},
                [own[p_], own[a]])
            gu = name_of("gu", "gradient_update")
            add(gu, "gradient_update", {"model": (agg_p, 0), "round": r, "time": r * 1000 + 500, "peer": p_,
                                        "gradient_model": cg}, [(agg_p, 0), (cg, 0)])
            t = name_of("train", "train")
            add(t, "train", {"model": (agg_a, 0), "round": r, "time": r * 1000 + 500, "peer": a, "local_steps": 2},
                [(agg_a, 0)])
            own[p_], own[a] = (gu, 0), (t, 0)
    return list(tasks.values())


def conflux_dag(n: int, rounds: int, sample_size: int, n_chunks: int, seed: int = 0) -> List[Dict]:
    """A Shatter/Conflux-shaped compute graph (dasklearn/simulation/conflux/client.py:93-222) in the reference's
    task format: every round a sample of `sample_size` nodes trains, each splits its model with a `chunk`
    task (n pieces, output idx = piece index) and sends the pieces to the next sample; a node of the next
    sample starts a `reconstruct_from_chunks` task as soon as it holds enough pieces, so the contributor
    set differs per chunk index (between 1 and sample_size pieces, in arrival order)."""
    rs = np.random.RandomState(seed)
    tasks: Dict[str, Dict] = {}

    def add(name, func, data, inputs):
        tasks[name] = {"name": name, "func": func, "data": data, "inputs": list(inputs), "outputs": []}
        for src, idx in inputs:
            tasks[src]["outputs"].append((name, idx))

    cnt = {"train": 0, "chunk": 0, "rec": 0}
    sample = sorted(rs.choice(n, sample_size, replace=False).tolist())
    model_of = {p: (None, 0) for p in sample}
    trains = [0] * n
    for r in range(1, rounds + 1):
        chunk_tasks = []
        for p in sample:
            cnt["train"] += 1
            trains[p] += 1
            tname = "train_%d" % cnt["train"]
            prev = model_of[p]
            add(tname, "train", {"model": prev, "round": trains[p], "time": r * 1000, "peer": p},
                [] if prev[0] is None else [prev])
            cnt["chunk"] += 1
            cname = "chunk_%d" % cnt["chunk"]
            add(cname, "chunk", {"model": (tname, 0), "round": r, "time": r * 1000 + 100, "peer": p, "n": n_chunks},
                [(tname, 0)])
            chunk_tasks.append(cname)
        nxt = sorted(rs.choice(n, sample_size, replace=False).tolist())
        model_of = {}
        for q in nxt:
            received = []
            for ci in range(n_chunks):
                k = int(rs.randint(1, sample_size + 1))
                senders = rs.permutation(sample_size)[:k].tolist()
                received.append([(chunk_tasks[s], ci) for s in senders])
            cnt["rec"] += 1
            rname = "reconstruct_from_chunks_%d" % cnt["rec"]
            add(rname, "reconstruct_from_chunks", {"chunks": received, "round": r + 1, "time": r * 1000 + 500, "peer": q},
                [m for c in received for m in c])
            model_of[q] = (rname, 0)
        sample = nxt
    return list(tasks.values())
