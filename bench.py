#!/usr/bin/env python
"""bench.py — D-PSGD round throughput of the B200 batched executor.

A "step" is one D-PSGD round over all simulated nodes: one train wave (local_steps minibatches of
local SGD per node, B = 32) followed by one gossip-aggregate wave (k-regular topology, fan-in k+1).
Default workload = the configuration BASELINE.json's metric is quoted on: **1024 nodes**, GNLeNet,
synthetic CIFAR-10-shaped data, k = 10.  With N GPUs the 1024 nodes are partitioned over the GPUs
(strong scaling; `--nodes-per-gpu` switches to weak scaling) and aggregation exchanges the
neighbour rows that cross GPUs.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
                    [--nodes 1024 | --nodes-per-gpu 256] [--model gnlenet|femnist]
                    [--mode round|agg] [--topology kreg|exp|ring] [--transport auto|a2a|push]

Prints ONE JSON line (rank 0).  `value` = node-train-steps/s of the whole job with every input
resident in HBM; `e2e` = the same metric through the public solver API starting from the
host-side task list (planning + descriptor upload + loss read-back inside the timed region).
`--mode agg` is the aggregation-only bandwidth sweep (BASELINE config C5: `--nodes 8192 --topology exp`).
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

NODES = 1024         # the metric's configuration ("... at 1024 nodes, 1-8 B200")
N_I = 320            # samples per node: 10 full batches -> exactly local_steps full steps
BATCH = 32
LOCAL_STEPS = 5
N_TEST = 256
MODELS = {"gnlenet": {"P": 85354, "gflop_per_node_step": 1.9033, "hw": 32, "c": 3, "classes": 10,
                      "lr": 0.002, "mu": 0.9, "label": "GNLeNet", "data": "CIFAR-10-shaped"},
          "femnist": {"P": 1690046, "gflop_per_node_step": 2.321, "hw": 28, "c": 1, "classes": 62,
                      "lr": 0.004, "mu": 0.0, "label": "FEMNIST CNN", "data": "FEMNIST-shaped 28x28"}}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nodes", type=int, default=0, help="total nodes (strong scaling; default 1024)")
    ap.add_argument("--nodes-per-gpu", type=int, default=0, help="weak scaling: nodes per GPU")
    ap.add_argument("--mode", default="round", choices=["round", "agg"],
                    help="round = train + aggregate waves (default); agg = aggregate waves only (C5 sweep)")
    ap.add_argument("--topology", default="kreg", choices=["kreg", "exp", "ring"])
    ap.add_argument("--transport", default="auto", choices=["auto", "a2a", "push"])
    ap.add_argument("--agg-variant", type=int, default=0)
    ap.add_argument("--precision", type=int, default=1,
                    help="1 = bf16 tcgen05 convolutions with fp32 accumulation (default), 0 = fp32 CUDA cores")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the C2 (256-node) secondary line at N=1")
    ap.add_argument("--model", default="gnlenet", choices=["gnlenet", "femnist"],
                    help="gnlenet = the metric's model (default); femnist = the C3 model")
    a = ap.parse_args()
    if a.nodes and a.nodes_per_gpu:
        ap.error("--nodes and --nodes-per-gpu are exclusive")
    return a


def peaks():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"],
                "bf16_tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "source": "measured"}
    # fallback stated in /opt/skills/guides/B200_PROFILING.md
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def traffic_of(key: str):
    """DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of the kernel / kernel group `key`,
    from the ncu captures summarised in profiles/traffic.json (written by scripts/ncu_traffic.py from a
    capture of THIS workload shape); None when no capture of that shape is tracked."""
    path = os.path.join(REPO, "profiles", "traffic.json")
    if not os.path.exists(path):
        return None
    e = json.load(open(path)).get(key)
    return None if e is None else e["bytes"]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.proc = None
        self.lines = []
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu_index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        lines, window = self.lines, "timed region"
        if t0 is not None:
            lines = [(ts, l) for ts, l in self.lines if t0 <= ts <= t1 + 0.2]
            if not lines:                          # a timed region shorter than the sampling period: the samples
                lines = [(ts, l) for ts, l in self.lines if t0 - 0.5 <= ts <= t1 + 0.5]    # around it (warm-up load)
                window = "no sample inside the timed region; +-0.5 s around it"
        for ts, l in lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "window": window}


# ---------------------------------------------------------------------------------------------
# Baseline legs (the ONLY places that touch oracle/): cpu_baseline, gpu_eager_baseline, --impl reference
# ---------------------------------------------------------------------------------------------
def _cpu_worker_init(threads):
    import torch
    torch.set_num_threads(threads)


def _cpu_train_task(args):
    peer, rnd, n_i, seed, model = args
    from oracle import oracle as O
    import dls_b200  # noqa: F401
    from dls_b200 import synthetic
    from dls_b200.layout import GNLENET, FEMNIST
    L = GNLENET if model == "gnlenet" else FEMNIST
    m = MODELS[model]
    imgs, labs = synthetic.make_images(n_i, m["hw"], m["c"], m["classes"], seed=1234 + peer)
    idx, cnt = synthetic.minibatch_index(seed, [peer], [rnd], n_i, BATCH, LOCAL_STEPS)
    cache = _cpu_train_task.__dict__.setdefault("init", {})
    if model not in cache:
        cache[model] = O.initial_row(L, seed)
    t0 = time.perf_counter()
    O.train_task(L, cache[model], None, imgs, labs, idx[0], cnt[0], m["lr"], m["mu"], 0.0)
    return time.perf_counter() - t0, int((cnt[0] > 0).sum())


def _cpu_agg_task(args):
    fan, seed, P = args
    import numpy as np
    from oracle import oracle as O
    rs = np.random.RandomState(seed)
    rows = [rs.standard_normal(P).astype(np.float32) for _ in range(fan)]
    t0 = time.perf_counter()
    O.fedavg_numpy(rows)
    return time.perf_counter() - t0


def cpu_sample(n_tasks: int, fan_in: int, procs: int, threads: int, model: str = "gnlenet"):
    """Oracle port of one D-PSGD round restricted to `n_tasks` nodes: their train tasks then their
    aggregates, on `procs` worker processes x `threads` torch threads (the reference's
    --workers x --torch-threads).  Returns (node-train-steps/s, agg GB/s, seconds, steps, t_train, t_agg)."""
    import concurrent.futures as cf
    import multiprocessing as mp
    P = MODELS[model]["P"]
    t0 = time.perf_counter()
    if procs <= 1:
        _cpu_worker_init(threads)
        tr = [_cpu_train_task((p, 1, N_I, 42, model)) for p in range(n_tasks)]
        t_train = time.perf_counter() - t0
        t1 = time.perf_counter()
        [_cpu_agg_task((fan_in, p, P)) for p in range(n_tasks)]
        t_agg = time.perf_counter() - t1
    else:
        ctx = mp.get_context("spawn")
        with cf.ProcessPoolExecutor(procs, mp_context=ctx, initializer=_cpu_worker_init, initargs=(threads,)) as ex:
            list(ex.map(_cpu_train_task, [(p, 1, 64, 42, model) for p in range(procs)]))     # warm the workers
            t0 = time.perf_counter()
            tr = list(ex.map(_cpu_train_task, [(p, 1, N_I, 42, model) for p in range(n_tasks)]))
            t_train = time.perf_counter() - t0
            t1 = time.perf_counter()
            list(ex.map(_cpu_agg_task, [(fan_in, p, P) for p in range(n_tasks)]))
            t_agg = time.perf_counter() - t1
    steps = sum(s for _, s in tr)
    total = t_train + t_agg
    return steps / total, (2 * n_tasks * P * 4 / 1e9) / t_agg, total, steps, t_train, t_agg


def gpu_eager_sample(n_tasks: int, fan_in: int, model: str, dev):
    """SURVEY §8(d) second baseline ("the real bar"): the reference's per-node loop with the tensors on cuda:0
    in PyTorch eager (reference: device choice main.py:13-19, model_trainer.py:84-85,129) — one node at a time,
    one minibatch at a time, torch autograd + the SGD update, then FedAvg with torch ops, as the oracle port
    does on the CPU.  Returns (node-train-steps/s, aggregate GB/s)."""
    import numpy as np
    import torch
    from oracle import oracle as O
    import dls_b200  # noqa: F401
    from dls_b200 import synthetic
    from dls_b200.layout import GNLENET, FEMNIST
    L = GNLENET if model == "gnlenet" else FEMNIST
    m = MODELS[model]
    init = torch.from_numpy(O.initial_row(L, 42)[:L.P].copy()).to(dev)
    data = []
    for p in range(n_tasks):
        imgs, labs = synthetic.make_images(N_I, m["hw"], m["c"], m["classes"], seed=1234 + p)
        idx, cnt = synthetic.minibatch_index(42, [p], [1], N_I, BATCH, LOCAL_STEPS)
        data.append((imgs, labs, idx[0], cnt[0]))

    def train_one(imgs, labs, idx, cnt):
        w = init.clone()
        buf = None
        for s in range(idx.shape[0]):
            b = int(cnt[s])
            if b <= 0:
                break
            sel = idx[s, :b].astype(np.int64)
            x = O.to_tensor(imgs[sel]).to(dev)                       # the DataLoader batch, moved like :101
            y = torch.from_numpy(labs[sel].astype(np.int64)).to(dev)
            wl = w.clone().requires_grad_(True)
            loss = O.loss_fn(L, O.forward(L, wl, x), y)
            loss.backward()
            g = wl.grad
            with torch.no_grad():
                if m["mu"] != 0:
                    buf = g.clone() if buf is None else buf.mul_(m["mu"]).add_(g)
                    g = buf
                w = w.add(g, alpha=-m["lr"])
        return w

    rows = [train_one(*data[0])]                                      # warm-up (cudnn heuristics, allocator)
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    rows = [train_one(*d) for d in data]
    torch.cuda.synchronize(dev)
    t_train = time.perf_counter() - t0
    steps = sum(int((d[3] > 0).sum()) for d in data)
    t1 = time.perf_counter()
    for p in range(n_tasks):                                          # fedavg.py:12-26 with torch ops on the device
        ms = [rows[(p + j) % n_tasks] for j in range(fan_in)]
        center = ms[0] * 0
        for q in ms:
            center.add_(q * (1.0 / fan_in))
    torch.cuda.synchronize(dev)
    t_agg = time.perf_counter() - t1
    return steps / (t_train + t_agg), (2 * n_tasks * L.P * 4 / 1e9) / t_agg, t_train + t_agg


def resolve_nodes(args, world):
    if args.nodes_per_gpu:
        return args.nodes_per_gpu * world, "weak"
    return (args.nodes or NODES), "strong"


def topology_k(args, n_nodes):
    if args.topology == "kreg":
        k = int(math.floor(math.log2(n_nodes)))
        if (n_nodes * k) % 2:
            k += 1
        return k
    if args.topology == "ring":
        return 2
    return int(math.ceil(math.log2(n_nodes)))


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path.  The reference is pure
    Python and cannot travel to the GPU box (no /root/reference there), so this times the oracle
    port (oracle/oracle.py, pinned bit-exact to the reference) on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_nodes, scaling = resolve_nodes(args, args.gpus)
    k = topology_k(args, n_nodes)
    cores = os.cpu_count() or 1
    threads = 4                                   # reference default --torch-threads (args.py:22)
    procs = max(1, cores // threads)
    sample = max(procs, min(n_nodes, 2 * procs))  # nodes per bounded step
    vals, ms = [], []
    for i in range(args.warmup + args.steps):
        v, agg_gbs, secs, steps, t_train, t_agg = cpu_sample(sample, k + 1, procs, threads, args.model)
        if i >= args.warmup:
            vals.append(v); ms.append(secs * 1e3)
        if sum(ms) > 150e3:
            break
    value = statistics.mean(vals)
    line = {"impl": "reference", "metric": "node-train-steps/sec", "value": value, "unit": "node-train-steps/s",
            "n_gpus": args.gpus, "steps": len(vals), "warmup": args.warmup, "ms_per_step": statistics.mean(ms),
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(n_nodes, k, args, args.gpus),
            "cpu_baseline": {"value": value, "unit": "node-train-steps/s", "cores": procs * threads, "kind": "port",
                             "sample": "%d of %d nodes per step: their train tasks (%d steps, B=%d) + their "
                                       "aggregates (fan-in %d); %d procs x %d threads" %
                                       (sample, n_nodes, LOCAL_STEPS, BATCH, k + 1, procs, threads)},
            "e2e": {"value": value, "unit": "node-train-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def workload_config(n_nodes, k, args, world):
    m = MODELS[args.model]
    per_gpu = n_nodes // max(1, world)
    tag = {1024: "the metric's configuration", 256: "BASELINE configs[1] (C2)"}.get(n_nodes, "")
    if args.model == "femnist" and n_nodes == 1024:
        tag = "BASELINE configs[2] (C3)"
    if args.mode == "agg":
        tag = "BASELINE configs[4] (C5) aggregation sweep" if (n_nodes == 8192 and args.topology == "exp") else "aggregation only"
    return {"workload": "D-PSGD %s: %d nodes (%d per GPU), %s topology (fan-in %d), %s (P=%d), synthetic %s uint8 "
                        "data, n_i=%d, B=%d, local_steps=%d, lr %g, momentum %g%s"
                        % ("round" if args.mode == "round" else "aggregate wave", n_nodes, per_gpu,
                           {"kreg": "%d-regular" % k, "exp": "exponential-graph", "ring": "ring"}[args.topology], k + 1,
                           m["label"], m["P"], m["data"], N_I, BATCH, LOCAL_STEPS, m["lr"], m["mu"],
                           (" — " + tag) if tag else ""),
            "nodes": n_nodes, "k": k, "batch": BATCH, "local_steps": LOCAL_STEPS, "mode": args.mode,
            "l2": "inputs larger than L2: the train wave streams its activation workspace (7.4 GB at 1024 nodes) "
                  "between aggregate launches and X + Y are %d MB, so nothing a wave reads is L2-resident when it starts"
                  % (2 * per_gpu * m["P"] * 4 // 1000000) if args.mode == "round" else
                  "inputs larger than L2: X + Y = %d MB per GPU; the pool alternates between two row sets"
                  % (2 * per_gpu * m["P"] * 4 // 1000000),
            "train_math": "fp32 CUDA cores" if args.precision == 0 else "conv2 on bf16 tcgen05 tensor cores (fp32 accumulate), the rest fp32 CUDA cores" if args.model == "femnist"
            else "bf16 tcgen05 tensor cores, fp32 accumulate",
            "dataset_residency": "node partitions uploaded once (like the reference's cached partitions); "
                                 "e2e uploads per-round task descriptors + minibatch index"}


_REAL_STDOUT = None


def emit(line: dict):
    """Print the ONE JSON line on the real stdout (library chatter such as NCCL's version banner is
    sent to stderr by the fd redirection in main())."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is not None:
        os.write(_REAL_STDOUT, data)
    else:
        sys.stdout.write(data.decode())
        sys.stdout.flush()


class Job:
    """One configured workload on this rank: topology, partition, dataset, executor factory."""

    def __init__(self, args, n_nodes, world, rank, dev, dist):
        import numpy as np
        import dls_b200  # noqa: F401
        from dls_b200 import synthetic
        from dls_b200.dataset import DeviceDataset
        from dls_b200.layout import GNLENET, FEMNIST
        from dls_b200.settings import SessionSettings, LearningSettings
        self.args, self.n_nodes, self.world, self.rank, self.dev, self.dist = args, n_nodes, world, rank, dev, dist
        self.L = GNLENET if args.model == "gnlenet" else FEMNIST
        m = MODELS[args.model]
        self.k = topology_k(args, n_nodes)
        if args.topology == "kreg":
            self.nbrs = synthetic.kreg_topology(n_nodes, self.k, seed=42)
        elif args.topology == "ring":
            self.nbrs = synthetic.ring_topology(n_nodes)
        else:
            self.nbrs = synthetic.exponential_topology(n_nodes)
        self.fan_in = max(len(x) for x in self.nbrs) + 1
        self.st = SessionSettings(participants=n_nodes, seed=42,
                                  dataset="femnist" if args.model == "femnist" else "cifar10",
                                  model=None if args.model == "femnist" else "gnlenet",
                                  data_dir=os.path.join(REPO, "gpurun_out", "bench_data"),
                                  learning=LearningSettings(m["lr"], m["mu"], 0.0, BATCH, LOCAL_STEPS), k=self.k,
                                  precision=args.precision)
        if world > 1:
            from dls_b200.partition import block_partition, partition_nodes
            # exponential graph: contiguous blocks (edges i -> i + 2^j); else grow parts inside the topology
            self.owner = block_partition(n_nodes, world) if args.topology == "exp" else partition_nodes(self.nbrs, world)
            self.owned = [int(x) for x in np.nonzero(self.owner == rank)[0]]
        else:
            self.owner, self.owned = None, list(range(n_nodes))
        self.n_local = len(self.owned)
        n_i = N_I if args.mode == "round" else 1
        self.ds = DeviceDataset.synthetic(self.L, n_nodes, n_i, N_TEST, dev, seed=42, node_ids=self.owned,
                                          per_node_seed=True)

    def executor(self):
        from dls_b200.executor import WaveExecutor
        return WaveExecutor(self.st, self.ds, device=self.dev, node_ids=self.owned if self.world > 1 else None)

    def solver(self, dag, ex):
        from dls_b200.solver import BatchedSolver
        if self.world > 1:
            from dls_b200.partition import DistributedSolver
            return DistributedSolver(ex, dag, self.rank, self.world, self.owner, agg_variant=self.args.agg_variant,
                                     transport=self.args.transport)
        return BatchedSolver(ex, dag, agg_variant=self.args.agg_variant)

    def barrier(self):
        import torch
        if self.dist is not None:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        import torch
        if self.dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def min_over_ranks(self, x: float) -> float:
        return -self.max_over_ranks(-x)


def device_leg(job, W, Ksteps, sampler=None):
    """Plan everything first (inputs resident in HBM), then time K rounds with CUDA events on the launching
    stream.  Returns a dict of timings; leaves `solver`, `ex` in it for the parity check."""
    import torch
    from dls_b200 import _lib, synthetic
    lib = _lib.lib()
    rounds = W + Ksteps
    if job.args.mode == "agg":
        dag = agg_only_dag(job, rounds)
    else:
        dag = synthetic.dpsgd_dag(job.n_nodes, rounds, job.nbrs)
    ex = job.executor()
    solver = job.solver(dag, ex)
    solver.plan()
    if job.args.mode == "agg":
        waves = [w for w in solver.waves if w.func in ("aggregate", "exchange")]
        init_waves = [w for w in solver.waves if w.func == "train"]
        for w in init_waves:                      # "train" placeholders: give every node a distinct row, no SGD
            ex.init_rows(w.init_slots)
            ex.X[w.out_slots] += 1e-3 * torch.randn((len(w.out_slots), ex.X.shape[1]), device=job.dev)
    else:
        waves = solver.waves
    per_round = len(waves) // rounds
    assert per_round * rounds == len(waves), (per_round, rounds, len(waves))

    def run_waves(ws, events=None):
        for w in ws:
            if events is not None:
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
            solver.run_wave(w)
            if events is not None:
                e1.record()
                events.append((w.func, e0, e1))

    run_waves(waves[:W * per_round])
    job.barrier()
    if sampler is not None:
        sampler.start()
        time.sleep(0.3)
    launches0 = lib.dls_launch_count()
    events = []
    t_wall0 = time.time()
    s0 = torch.cuda.Event(enable_timing=True); s1 = torch.cuda.Event(enable_timing=True)
    job.barrier()
    s0.record()
    run_waves(waves[W * per_round:], events)
    s1.record()
    job.barrier()
    t_wall1 = time.time()
    out = {"launches": lib.dls_launch_count() - launches0,
           "clocks": sampler.stop(t_wall0, t_wall1) if sampler is not None else None,
           "total_ms": job.max_over_ranks(s0.elapsed_time(s1)), "solver": solver, "ex": ex, "waves": waves,
           "per_round": per_round}
    for f in ("train", "aggregate"):
        ms = [e0.elapsed_time(e1) for fn, e0, e1 in events if fn == f]
        out[f + "_ms"] = job.max_over_ranks(statistics.mean(ms)) if ms else None
    return out


def agg_only_dag(job, rounds):
    """Aggregation-only graph for the bandwidth sweep: one placeholder `train` source per node (the executor only
    initialises its row) followed by `rounds` aggregate levels over the topology."""
    tasks = {}

    def add(name, func, data, inputs):
        tasks[name] = {"name": name, "func": func, "data": data, "inputs": list(inputs), "outputs": []}
        for src, idx in inputs:
            tasks[src]["outputs"].append((name, idx))

    cur = []
    for p in range(job.n_nodes):
        add("train_%d" % p, "train", {"model": (None, 0), "round": 0, "time": 0, "peer": p, "local_steps": 0}, [])
        cur.append(("train_%d" % p, 0))
    c = 0
    for r in range(1, rounds + 1):
        nxt = []
        for p in range(job.n_nodes):
            models = [cur[j] for j in job.nbrs[p]] + [cur[p]]
            c += 1
            add("agg_%d" % c, "aggregate", {"models": models, "round": r, "peer": p}, models)
            nxt.append(("agg_%d" % c, 0))
        cur = nxt
    return list(tasks.values())


def parity_check(job, leg):
    """N > 1: evidence that the multi-GPU aggregate path (plan -> halo exchange -> SpMM on pool + halo rows)
    computes what a single address space would.  (a) the LAST aggregate wave of the timed run is re-evaluated
    from an all_gather of every rank's source rows with the gather kernel (variant 1) on global row indices and
    must match bit for bit; (b) a 64-node, 2-round graph is solved distributed and alone on rank 0: the
    checksums of all final rows must be equal.  Raises on any mismatch."""
    import numpy as np
    import torch
    from dls_b200 import kernels as K, synthetic
    from dls_b200.solver import BatchedSolver
    dist, world, rank, dev = job.dist, job.world, job.rank, job.dev
    solver, ex, L = leg["solver"], leg["ex"], job.L
    last = [w for w in leg["waves"] if w.func == "aggregate"][-1]
    got = ex.X[torch.as_tensor(last.out_slots, device=dev).long()].clone()          # the real path's result
    # every rank's source rows of that level = the outputs of its tasks one level below, in wave order
    lvl = last.level
    prod = [w for w in solver.waves if w.level == lvl - 1 and w.func in ("train", "aggregate")]
    names = [n for w in prod for n in w.names]
    slots = [s for w in prod for s in w.out_slots]
    n_loc = torch.tensor([len(names)], device=dev)
    n_all = [torch.zeros_like(n_loc) for _ in range(world)]
    dist.all_gather(n_all, n_loc)
    n_max = max(int(x.item()) for x in n_all)
    mine = torch.zeros((n_max, ex.X.shape[1]), device=dev)
    mine[:len(slots)] = ex.X[torch.as_tensor(slots, device=dev).long()]
    allrows = torch.empty((world * n_max, ex.X.shape[1]), device=dev)
    dist.all_gather_into_tensor(allrows, mine)
    name_lists = [None] * world
    dist.all_gather_object(name_lists, names)
    where = {n: r * n_max + i for r in range(world) for i, n in enumerate(name_lists[r])}
    rowptr, col, wv = [0], [], []
    for t_name in last.names:
        d = solver.tasks[t_name]["data"]
        ms = [tuple(m) for m in d["models"]]
        ws = d.get("weights") or [float(1.0 / len(ms))] * len(ms)
        col += [where[m[0]] for m in ms]
        wv += list(ws)
        rowptr.append(len(col))
    Y = torch.zeros((len(last.names), ex.X.shape[1]), device=dev)
    K.agg_spmm(allrows, Y, torch.tensor(rowptr, dtype=torch.int32, device=dev),
               torch.tensor(col, dtype=torch.int32, device=dev),
               torch.tensor(np.asarray(wv, dtype=np.float64).astype(np.float32), device=dev), L.P, variant=1)
    torch.cuda.synchronize()
    ok_a = bool(torch.equal(Y[:, :L.P].view(torch.int32), got[:, :L.P].view(torch.int32)))
    # (b) small graph: distributed vs alone
    n_small = 8 * world
    nb = synthetic.kreg_topology(n_small, 4, seed=7)
    dag = synthetic.dpsgd_dag(n_small, 2, nb)
    from dls_b200.dataset import DeviceDataset
    from dls_b200.executor import WaveExecutor
    from dls_b200.partition import DistributedSolver, partition_nodes
    from dls_b200.settings import SessionSettings, LearningSettings
    m = MODELS[job.args.model]
    st = SessionSettings(participants=n_small, seed=42, dataset=job.st.dataset, model=job.st.model,
                         data_dir=job.st.data_dir, learning=LearningSettings(m["lr"], m["mu"], 0.0, BATCH, 2),
                         precision=job.st.precision)
    owner = partition_nodes(nb, world)
    owned = [int(x) for x in np.nonzero(owner == rank)[0]]
    ds = DeviceDataset.synthetic(L, n_small, 64, 8, dev, seed=42, node_ids=owned, per_node_seed=True)
    ex2 = WaveExecutor(st, ds, device=dev, node_ids=owned)
    s2 = DistributedSolver(ex2, dag, rank, world, owner, transport=job.args.transport)
    res = s2.plan().run()
    sums = {n: int(ex2.X[h[0].slot, :L.P].view(torch.int32).to(torch.int64).sum().item()) for n, h in res.items()}
    all_sums = [None] * world
    dist.all_gather_object(all_sums, sums)
    ok_b, n_rows_b = True, 0
    if rank == 0:
        ds0 = DeviceDataset.synthetic(L, n_small, 64, 8, dev, seed=42, per_node_seed=True)
        ex0 = WaveExecutor(st, ds0, device=dev)
        res0 = BatchedSolver(ex0, dag).plan().run()
        want = {n: int(ex0.X[h[0].slot, :L.P].view(torch.int32).to(torch.int64).sum().item()) for n, h in res0.items()}
        merged = {}
        for d_ in all_sums:
            merged.update(d_)
        ok_b = merged == want
        n_rows_b = len(want)
    flag = torch.tensor([1 if (ok_a and ok_b) else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    n_rows = torch.tensor([len(last.names)], device=dev)
    dist.all_reduce(n_rows)
    result = {"agg_bit_exact": bool(flag.item() == 1), "rows_checked": int(n_rows.item()),
              "small_graph_rows_checked": n_rows_b, "transport": solver.transport}
    if flag.item() != 1:
        raise SystemExit("parity_check FAILED on rank %d: aggregate wave bit-exact %s, small graph checksums %s"
                         % (rank, ok_a, ok_b))
    return result


def e2e_leg(job, W, Ksteps):
    """Host task list -> plan -> run -> loss read-back.  Every round: the solver is built from the round's
    serialized task list (host objects), plans and uploads its descriptors + minibatch index (pinned staging),
    runs both waves and copies the train losses back into pinned host memory.  Two timings of the same call
    sequence: sync (one round at a time, host waits for the loss) and pipelined (the host builds/queues round
    i+1 while round i runs; all loss copies complete inside the timed region) — the throughput figure."""
    import torch
    from dls_b200 import synthetic
    n_sync = min(Ksteps, 5)
    n_e2e = W + n_sync + Ksteps
    round_dags = [synthetic.dpsgd_dag(job.n_nodes, 1, job.nbrs, name_offset=i * job.n_nodes) for i in range(n_e2e)]
    host_losses = [torch.zeros((LOCAL_STEPS, job.n_local), dtype=torch.float32).pin_memory() for _ in range(n_e2e)]
    state = {"ex": job.executor(), "prev": None, "h2d": 0}

    def e2e_round(i):
        s2 = job.solver(round_dags[i], state["ex"])                  # host-side task list -> solver
        if state["prev"] is not None:
            s2.seed_models(state["prev"])                            # models stay in HBM between rounds
        loss = s2.run_with_loss()                                    # streams: plans level k+1 while level k runs
        host_losses[i].copy_(loss, non_blocking=True)                # D2H result read into pinned memory
        state["prev"] = s2.final_slots()
        s2.finish()
        state["h2d"] = s2.h2d_bytes

    prof = None
    if os.environ.get("DLS_PROFILE_E2E") and job.rank == 0:          # host-side profile of the e2e loop (stderr)
        import cProfile
        prof = cProfile.Profile()
    sync_ms = []
    for i in range(W + n_sync):
        job.barrier()
        t0 = time.perf_counter()
        e2e_round(i)
        job.barrier()                                                # loss is on the host
        if i >= W:
            sync_ms.append(job.max_over_ranks((time.perf_counter() - t0) * 1e3))
    job.barrier()
    if prof is not None:
        prof.enable()
    t0 = time.perf_counter()
    for i in range(W + n_sync, n_e2e):
        e2e_round(i)
    job.barrier()                                                    # every round's loss is on the host
    ms_per_step = job.max_over_ranks((time.perf_counter() - t0) * 1e3) / Ksteps
    if prof is not None:
        prof.disable()
        import pstats
        pstats.Stats(prof, stream=sys.stderr).sort_stats("tottime").print_stats(30)
    assert all(bool(torch.isfinite(h).all()) and float(h.abs().sum()) > 0 for h in host_losses), "e2e losses"
    return {"ms_per_step": ms_per_step, "sync_ms": statistics.mean(sync_ms), "h2d": int(state["h2d"]),
            "d2h": host_losses[0].numel() * 4}


def test_wave_leg(job, leg):
    """The test wave (`functions.test` for every local node; the reference's dominant cost)."""
    import torch
    ex = leg["ex"]
    slots = [w for w in leg["waves"] if w.func == "aggregate"][-1].out_slots
    ex.test_wave(slots)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    e0.record()
    for _ in range(reps):
        ex.test_wave(slots)
    e1.record()
    torch.cuda.synchronize()
    ms = job.max_over_ranks(e0.elapsed_time(e1) / reps)
    return {"ms_per_wave": ms, "models": len(slots) * job.world, "test_images": N_TEST,
            "model_test_images_per_s": len(slots) * job.world * N_TEST / (ms / 1e3),
            "note": "every node's model evaluated on the %d-image test split (host read-back of hits / loss included)" % N_TEST}


def main():
    global _REAL_STDOUT
    args = parse()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)                       # everything else that writes to fd 1 now goes to stderr
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import dls_b200  # noqa: F401
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node == --gpus"

    n_nodes, scaling = resolve_nodes(args, world)
    assert n_nodes % world == 0, "nodes must divide over the GPUs"
    m = MODELS[args.model]
    P, W, Ksteps = m["P"], args.warmup, args.steps
    job = Job(args, n_nodes, world, rank, dev, dist)
    pk = peaks()
    leg = device_leg(job, W, Ksteps, ClockSampler(local_rank) if job.rank == 0 else None)
    total_ms = leg["total_ms"]
    per_gpu = job.n_local
    agg_bytes = 2 * per_gpu * P * 4                                   # per GPU per launch
    agg_t, train_t = leg["aggregate_ms"], leg["train_ms"]
    agg_gbs = agg_bytes / 1e9 / (agg_t / 1e3)
    roof_agg = {"kernel": "aggregate wave (agg_spmm%s)" % (" + halo exchange" if world > 1 else ""), "bound": "hbm",
                "achieved": agg_gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": agg_gbs / pk["hbm_gbs"],
                "traffic": traffic_of("agg_%s_%d" % (args.model, per_gpu)),
                "algorithmic_bytes": agg_bytes, "ms_per_wave": agg_t,
                "peak_source": pk["source"] + " (burst copy bandwidth, MEASURED_PEAKS.json)"}
    parity = parity_check(job, leg) if world > 1 else None

    if args.mode == "agg":
        value = agg_bytes * world / 1e9 / (total_ms / Ksteps / 1e3)
        line = {"metric": "gossip-aggregate HBM GB/s", "value": value, "unit": "GB/s", "n_gpus": world, "steps": Ksteps,
                "warmup": W, "ms_per_step": total_ms / Ksteps, "higher_is_better": True, "scaling": scaling,
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(n_nodes, job.k, args, world),
                "roofline": roof_agg, "gpu_launches": int(leg["launches"]), "clocks": leg["clocks"],
                "e2e": None}
        if world > 1:
            s = leg["solver"]
            line["halo"] = halo_info(s, W + Ksteps, agg_t)
            line["parity_check"] = parity
        if rank == 0:
            emit(line)
        if dist is not None:
            dist.destroy_process_group()
        return

    node_steps_per_round = n_nodes * LOCAL_STEPS
    value = node_steps_per_round * Ksteps / (total_ms / 1e3)
    train_tflops = per_gpu * LOCAL_STEPS * m["gflop_per_node_step"] / 1e3 / (train_t / 1e3)
    tw = test_wave_leg(job, leg)
    leg["solver"].finish()
    e2e = e2e_leg(job, W, Ksteps)
    e2e_value = node_steps_per_round / (e2e["ms_per_step"] / 1e3)
    math_label = "fp32 CUDA cores" if args.precision == 0 else "bf16 tcgen05"

    if rank == 0:
        line = {
            "metric": "node-train-steps/sec", "value": value, "unit": "node-train-steps/s", "n_gpus": world,
            "steps": Ksteps, "warmup": W, "ms_per_step": total_ms / Ksteps, "higher_is_better": True,
            "scaling": scaling, "vs_baseline": None,
            "dtype": "f32" if args.precision == 0 else "bf16",
            "data": "synthetic", "config": workload_config(n_nodes, job.k, args, world),
            # the dominant kernel group: all kernels of the train wave (99 % of the step)
            "roofline": {"kernel": "train wave (all kernels of %d local steps x %d nodes per GPU)" % (LOCAL_STEPS, per_gpu),
                         "bound": "tensor", "achieved": train_tflops, "peak": pk["bf16_tflops_sustained"],
                         "unit": "TFLOP/s", "frac": train_tflops / pk["bf16_tflops_sustained"],
                         "traffic": traffic_of("train_%s_%d" % (args.model, per_gpu)),
                         "algorithmic_flops": per_gpu * LOCAL_STEPS * m["gflop_per_node_step"] * 1e9,
                         "ms_per_wave": train_t,
                         "peak_source": pk["source"] + " (sustained bf16 dense, MEASURED_PEAKS.json; kernel math: " + math_label + ")"},
            "roofline_agg": roof_agg,
            "aggregate": {"metric": "gossip-aggregate HBM GB/s", "value": agg_gbs * world, "unit": "GB/s",
                          "ms_per_wave": agg_t, "bytes_per_wave_per_gpu": agg_bytes,
                          "frac_of_hbm_peak": agg_gbs / pk["hbm_gbs"]},
            "train": {"ms_per_wave": train_t, "tflops_gemm": train_tflops * world,
                      "node_train_steps_per_s_train_only": node_steps_per_round / (train_t / 1e3)},
            "test_wave": tw,
            "e2e": {"value": e2e_value, "unit": "node-train-steps/s", "h2d_bytes_per_step": e2e["h2d"],
                    "d2h_bytes_per_step": e2e["d2h"], "ms_per_step": e2e["ms_per_step"],
                    "mode": "pipelined: %d rounds back to back through the solver's run_with_loss(); the host "
                            "builds and queues round i+1 while round i runs; every round's descriptors go H2D from "
                            "pinned staging and its losses come back D2H inside the timed region" % Ksteps,
                    "sync_ms_per_step": e2e["sync_ms"],
                    "sync_value": node_steps_per_round / (e2e["sync_ms"] / 1e3)},
            "gpu_launches": int(leg["launches"]), "clocks": leg["clocks"],
        }
        if world > 1:
            line["halo"] = halo_info(leg["solver"], W + Ksteps, agg_t)
            line["parity_check"] = parity
    # baselines and the secondary (C2) line: rank 0 at N = 1 only
    if world == 1 and rank == 0:
        if not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            n_cpu = 256 if args.model == "gnlenet" else 32
            v, cpu_agg_gbs, secs, steps, t_train, t_agg = cpu_sample(n_cpu, job.fan_in, 1, 4, args.model)
            line["cpu_baseline"] = {"value": v, "unit": "node-train-steps/s", "cores": 4, "kind": "port",
                                    "agg_gbs": cpu_agg_gbs,
                                    "sample": "%d of %d nodes: their train tasks (%d steps, B=%d) + aggregates "
                                              "(fan-in %d), 1 worker x 4 torch threads (reference defaults), "
                                              "%.1f s; host has %d cores" % (n_cpu, n_nodes, LOCAL_STEPS, BATCH,
                                                                            job.fan_in, secs, cores)}
            n_eager = 256 if args.model == "gnlenet" else 32
            ev, eg, esecs = gpu_eager_sample(n_eager, job.fan_in, args.model, dev)
            line["gpu_eager_baseline"] = {"value": ev, "unit": "node-train-steps/s", "agg_gbs": eg,
                                          "kind": "port on cuda:0, PyTorch eager, one node at a time",
                                          "sample": "%d of %d nodes, %.1f s" % (n_eager, n_nodes, esecs)}
        if not args.no_secondary and n_nodes != 256 and args.model == "gnlenet":
            del job, leg
            torch.cuda.empty_cache()
            job2 = Job(args, 256, 1, 0, dev, None)
            leg2 = device_leg(job2, W, Ksteps)
            t2, a2 = leg2["train_ms"], leg2["aggregate_ms"]
            line["secondary"] = {"workload": "BASELINE configs[1] (C2): 256 nodes, 8-regular, one GPU",
                                 "value": 256 * LOCAL_STEPS * Ksteps / (leg2["total_ms"] / 1e3),
                                 "unit": "node-train-steps/s", "ms_per_step": leg2["total_ms"] / Ksteps,
                                 "train_ms_per_wave": t2, "agg_ms_per_wave": a2,
                                 "train_frac_of_tensor_peak": 256 * LOCAL_STEPS * m["gflop_per_node_step"] / 1e3 / (t2 / 1e3) / pk["bf16_tflops_sustained"],
                                 "agg_frac_of_hbm_peak": 2 * 256 * P * 4 / 1e9 / (a2 / 1e3) / pk["hbm_gbs"]}
    if rank == 0:
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


def halo_info(solver, rounds, agg_ms):
    b = int(solver.halo_bytes // max(1, rounds))
    return {"bytes_per_gpu_per_wave": b, "rows_per_gpu_per_wave": int(solver.max_halo),
            "transport": {"push": "one-sided push into peers' halo rows over NVLink (dls_push_rows, symmetric memory) + "
                                  "one cross-GPU barrier",
                          "a2a": "pack + NCCL all_to_all_single over NVLink, rows land after the pool"}[solver.transport],
            "inbound_gbs_per_gpu_over_whole_wave": b / 1e9 / (agg_ms / 1e3) if agg_ms else None,
            "note": "aggregate ms_per_wave includes the exchange"}


if __name__ == "__main__":
    main()
