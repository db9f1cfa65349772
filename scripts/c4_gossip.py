#!/usr/bin/env python
"""BASELINE config C4 at size: asynchronous gossip learning, 1024 nodes, heterogeneous speeds and churn.

The reference's fedscale trace / churn files are not in its tree, so the schedule is synthesised
with the schema its loaders read: every node gets a compute speed (ms per batch, log-normal like the fedscale
capacities) and alternating online / offline sessions; a node that is online trains on its current model, pushes
the result to a random online peer, which merges it with its own model using age-proportional weights
(AsynchronousClient.aggregate, simulation/asynchronous_client.py:53-74).  The event order comes from a small
discrete-event loop here; the output is the reference's serialized task format, solved by BatchedSolver.

    python scripts/c4_gossip.py [--nodes 1024] [--events 8192] [--check 4]
prints one JSON line: wavefront statistics, solve time, tasks/s, and the oracle comparison of `--check` sink models.
"""
import argparse, heapq, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def gossip_trace_dag(n, events, seed=0, churn=True):
    rs = np.random.RandomState(seed)
    speed = np.exp(rs.normal(np.log(40.0), 0.6, n))               # ms per batch ("computation" of the fedscale schema)
    session = rs.exponential(20000.0, n) if churn else np.full(n, 1e18)   # ms online before the first offline period
    tasks = {}

    def add(name, func, data, inputs):
        tasks[name] = {"name": name, "func": func, "data": data, "inputs": list(inputs), "outputs": []}
        for src, idx in inputs:
            tasks[src]["outputs"].append((name, idx))

    model_of = [(None, 0)] * n
    age = [0] * n
    trains = [0] * n
    offline_until = np.zeros(n)
    heap = [(float(speed[i] * 5), i) for i in range(n)]           # (finish time of the node's current train, node)
    heapq.heapify(heap)
    n_train = n_agg = skipped = 0
    while n_train < events and heap:
        t, i = heapq.heappop(heap)
        if churn and t > session[i]:                              # the node leaves for a while, its model survives
            offline_until[i] = t + rs.exponential(5000.0)
            session[i] = offline_until[i] + rs.exponential(20000.0)
            heapq.heappush(heap, (float(offline_until[i] + speed[i] * 5), i))
            skipped += 1
            continue
        n_train += 1
        trains[i] += 1
        name = "train_%d" % n_train
        prev = model_of[i]
        add(name, "train", {"model": prev, "round": trains[i], "time": int(t * 1000), "peer": i}, [] if prev[0] is None else [prev])
        model_of[i] = (name, 0)
        age[i] += 1
        j = int(rs.randint(n - 1)); j = j if j < i else j + 1
        if model_of[j][0] is not None and offline_until[j] <= t:  # push to an online peer that has a model
            n_agg += 1
            aname = "agg_%d" % n_agg
            models = [model_of[i], model_of[j]]
            tot = float(age[i] + age[j])
            add(aname, "aggregate", {"models": list(models), "round": max(age[i], age[j]), "peer": j,
                                     "weights": [age[i] / tot, age[j] / tot]}, models)
            model_of[j] = (aname, 0)
            age[j] = max(age[i], age[j])
        heapq.heappush(heap, (float(t + speed[i] * 5), i))
    return list(tasks.values()), {"trains": n_train, "aggregates": n_agg, "offline_periods": skipped}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=1024)
    ap.add_argument("--events", type=int, default=8192)
    ap.add_argument("--check", type=int, default=4)
    ap.add_argument("--precision", type=int, default=1)
    a = ap.parse_args()
    import torch
    import dls_b200
    from dls_b200 import synthetic
    from dls_b200.dataset import DeviceDataset
    from dls_b200.executor import WaveExecutor
    from dls_b200.layout import GNLENET as L
    from dls_b200.settings import SessionSettings, LearningSettings
    from dls_b200.solver import BatchedSolver
    n, n_i, steps = a.nodes, 64, 2
    t0 = time.time()
    dag, info = gossip_trace_dag(n, a.events, seed=1)
    t_gen = time.time() - t0
    dev = torch.device("cuda:0")
    st = SessionSettings(participants=n, seed=42, algorithm="gl", data_dir=os.path.join("gpurun_out", "c4"),
                         learning=LearningSettings(0.002, 0.9, 0.0, 32, steps), precision=a.precision)
    ds = DeviceDataset.synthetic(L, n, n_i, 16, dev, seed=42, per_node_seed=True)
    ex = WaveExecutor(st, ds, device=dev)
    solver = BatchedSolver(ex, dag)
    t0 = time.time()
    solver.plan()
    t_plan = time.time() - t0
    torch.cuda.synchronize()
    t0 = time.time()
    res = solver.run(timing=True)
    torch.cuda.synchronize()
    t_run = time.time() - t0
    widths = {}
    for f, nt, ms in solver.wave_times:
        widths.setdefault(f, []).append(nt)
    out = {"config": "C4: asynchronous gossip, %d nodes, synthetic speeds + churn, %d train events" % (n, a.events),
           "tasks": len(dag), **info, "levels": max(w.level for w in solver.waves) + 1, "waves": len(solver.waves),
           "train_wave_width": {"median": float(np.median(widths.get("train", [0]))), "max": int(max(widths.get("train", [0])))},
           "aggregate_wave_width": {"median": float(np.median(widths.get("aggregate", [0]))), "max": int(max(widths.get("aggregate", [0])))},
           "slots": ex.n_slots, "plan_s": t_plan, "solve_s": t_run, "tasks_per_s": len(dag) / t_run,
           "node_train_steps_per_s": info["trains"] * steps / t_run, "graph_generation_s": t_gen, "sinks": len(res)}
    if a.check:
        from oracle import oracle as O
        # the oracle solves the ancestors of a few sink models only (the full graph would take minutes on the CPU)
        by = {t["name"]: t for t in dag}
        sinks = sorted(res)[:a.check]
        need, stack = set(), list(sinks)
        while stack:
            x = stack.pop()
            if x in need:
                continue
            need.add(x)
            stack.extend(i[0] for i in by[x]["inputs"])
        sub = [dict(t, outputs=[o for o in t["outputs"] if o[0] in need]) for t in dag if t["name"] in need]
        imgs = np.empty((n, n_i, 32, 32, 3), np.uint8); labs = np.empty((n, n_i), np.int64)
        peers = {t["data"]["peer"] for t in sub}
        for p in peers:
            imgs[p], labs[p] = synthetic.make_images(n_i, 32, 3, 10, seed=1234 + p)
        timg, tlab = synthetic.make_test_split(16, 32, 3, 10)
        torch.set_num_threads(8)
        want, _ = O.solve_dag(L, sub, imgs, labs, timg, tlab, seed=42, batch=32, local_steps=steps, lr=0.002, mu=0.9,
                              wd=0.0, n_i=n_i, minibatch_permutation=synthetic.minibatch_permutation)
        errs = []
        for s in sinks:
            got = res[s][0].numpy().astype(np.float64)
            errs.append(float(np.linalg.norm(got - want[s][:L.P]) / np.linalg.norm(want[s][:L.P])))
        out["oracle_check"] = {"sinks": len(sinks), "ancestor_tasks": len(sub), "max_rel_err": max(errs)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
