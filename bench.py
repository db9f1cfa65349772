#!/usr/bin/env python
"""bench.py — D-PSGD round throughput of the B200 batched executor.

A "step" is one D-PSGD round over all simulated nodes: one train wave (local_steps minibatches
of local SGD per node, GNLeNet, B=32) followed by one gossip-aggregate wave (k-regular topology,
fan-in k+1).  Workload at 1 GPU = BASELINE.json configs[1]: 256 nodes, GNLeNet, synthetic
CIFAR-10-shaped data (k = log2 N = 8).  With N GPUs the nodes are sharded 256 per GPU (weak
scaling) and aggregation exchanges the neighbour rows that cross GPUs.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Prints ONE JSON line (rank 0).  `value` = node-train-steps/s of the whole job with every input
resident in HBM; `e2e` = the same metric through the public solver API starting from the
host-side task list (planning + descriptor upload + loss read-back inside the timed region).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

NODES_PER_GPU = 256
N_I = 320            # samples per node: 10 full batches -> exactly local_steps full steps
BATCH = 32
LOCAL_STEPS = 5
GNLENET_P = 85354
GFLOP_PER_NODE_STEP = 1.9033          # GEMM FLOPs only, fwd + wgrad + dgrad


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nodes-per-gpu", type=int, default=NODES_PER_GPU)
    ap.add_argument("--agg-variant", type=int, default=0)
    ap.add_argument("--precision", type=int, default=1,
                    help="1 = bf16 tcgen05 convolutions with fp32 accumulation (default), 0 = fp32 CUDA cores")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--model", default="gnlenet", choices=["gnlenet", "femnist"],
                    help="gnlenet = BASELINE configs[1] (default); femnist = the C3 model (fp32 path)")
    return ap.parse_args()


def peaks():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"],
                "bf16_tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "source": "measured"}
    # fallback stated in /opt/skills/guides/B200_PROFILING.md
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


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
                                          "--format=csv,noheader,nounits", "-lms", "100"],
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
        for ts, l in self.lines:
            if t0 is not None and not (t0 <= ts <= (t1 + 0.2)):
                continue
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
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU legs (the ONLY places that touch oracle/): cpu_baseline and --impl reference
# ---------------------------------------------------------------------------------------------
def _cpu_worker_init(threads):
    import torch
    torch.set_num_threads(threads)


def _cpu_train_task(args):
    peer, rnd, n_i, seed = args
    import numpy as np
    from oracle import oracle as O
    import dls_b200  # noqa: F401
    from dls_b200 import synthetic
    from dls_b200.layout import GNLENET as L
    rs = np.random.RandomState(peer)
    imgs, labs = synthetic.make_images(n_i, 32, 3, 10, seed=1234 + peer)
    idx, cnt = synthetic.minibatch_index(seed, [peer], [rnd], n_i, BATCH, LOCAL_STEPS)
    row = _cpu_train_task.init if hasattr(_cpu_train_task, "init") else O.initial_row(L, seed)
    _cpu_train_task.init = row
    t0 = time.perf_counter()
    new_row, mom, _ = O.train_task(L, row, None, imgs, labs, idx[0], cnt[0], 0.002, 0.9, 0.0)
    return time.perf_counter() - t0, int((cnt[0] > 0).sum())


def _cpu_agg_task(args):
    fan, seed = args
    import numpy as np
    from oracle import oracle as O
    rs = np.random.RandomState(seed)
    rows = [rs.standard_normal(GNLENET_P).astype(np.float32) for _ in range(fan)]
    t0 = time.perf_counter()
    O.fedavg_numpy(rows)
    return time.perf_counter() - t0


def cpu_sample(n_tasks: int, fan_in: int, procs: int, threads: int):
    """Oracle port of one D-PSGD round restricted to `n_tasks` nodes: their train tasks then their
    aggregates, on `procs` worker processes x `threads` torch threads (the reference's
    --workers x --torch-threads).  Returns (node-train-steps/s, agg GB/s, seconds, steps)."""
    import concurrent.futures as cf
    import multiprocessing as mp
    t0 = time.perf_counter()
    if procs <= 1:
        _cpu_worker_init(threads)
        tr = [_cpu_train_task((p, 1, N_I, 42)) for p in range(n_tasks)]
        t_train = time.perf_counter() - t0
        t1 = time.perf_counter()
        ag = [_cpu_agg_task((fan_in, p)) for p in range(n_tasks)]
        t_agg = time.perf_counter() - t1
    else:
        ctx = mp.get_context("spawn")
        with cf.ProcessPoolExecutor(procs, mp_context=ctx, initializer=_cpu_worker_init, initargs=(threads,)) as ex:
            list(ex.map(_cpu_train_task, [(p, 1, 64, 42) for p in range(procs)]))     # warm the workers
            t0 = time.perf_counter()
            tr = list(ex.map(_cpu_train_task, [(p, 1, N_I, 42) for p in range(n_tasks)]))
            t_train = time.perf_counter() - t0
            t1 = time.perf_counter()
            ag = list(ex.map(_cpu_agg_task, [(fan_in, p) for p in range(n_tasks)]))
            t_agg = time.perf_counter() - t1
    steps = sum(s for _, s in tr)
    total = t_train + t_agg
    return steps / total, (2 * n_tasks * GNLENET_P * 4 / 1e9) / t_agg, total, steps, t_train, t_agg


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path.  The reference is pure
    Python and cannot travel to the GPU box (no /root/reference there), so this times the oracle
    port (oracle/oracle.py, pinned bit-exact to the reference) on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import math
    n_nodes = args.nodes_per_gpu * args.gpus
    k = int(math.floor(math.log2(n_nodes)))
    cores = os.cpu_count() or 1
    threads = 4                                   # reference default --torch-threads (args.py:22)
    procs = max(1, cores // threads)
    sample = max(procs, min(n_nodes, 2 * procs))  # nodes per bounded step
    vals, ms = [], []
    for i in range(args.warmup + args.steps):
        v, agg_gbs, secs, steps, t_train, t_agg = cpu_sample(sample, k + 1, procs, threads)
        if i >= args.warmup:
            vals.append(v); ms.append(secs * 1e3)
        if sum(ms) > 150e3:
            break
    value = statistics.mean(vals)
    line = {"impl": "reference", "metric": "node-train-steps/sec", "value": value, "unit": "node-train-steps/s",
            "n_gpus": args.gpus, "steps": len(vals), "warmup": args.warmup, "ms_per_step": statistics.mean(ms),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(n_nodes, k, args),
            "cpu_baseline": {"value": value, "unit": "node-train-steps/s", "cores": procs * threads, "kind": "port",
                             "sample": "%d of %d nodes per step: their train tasks (%d steps, B=%d) + their "
                                       "aggregates (fan-in %d); %d procs x %d threads" %
                                       (sample, n_nodes, LOCAL_STEPS, BATCH, k + 1, procs, threads)},
            "e2e": {"value": value, "unit": "node-train-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def workload_config(n_nodes, k, args):
    return {"workload": "D-PSGD round: %d nodes (%d per GPU), %d-regular topology, %s (P=%d), synthetic "
                        "%s uint8 data, n_i=%d, B=%d, local_steps=%d, %s"
                        % (n_nodes, args.nodes_per_gpu, k, "GNLeNet" if args.model == "gnlenet" else "FEMNIST CNN",
                           GNLENET_P, "CIFAR-10-shaped" if args.model == "gnlenet" else "FEMNIST-shaped 28x28",
                           N_I, BATCH, LOCAL_STEPS,
                           "lr 0.002, momentum 0.9" if args.model == "gnlenet" else "lr 0.004, momentum 0"),
            "nodes": n_nodes, "k": k, "batch": BATCH, "local_steps": LOCAL_STEPS,
            "l2": "inputs larger than L2: the train wave streams its ~1.9 GB activation workspace between "
                  "aggregate launches, so X/Y (87 MB) are not L2-resident when a wave starts",
            "train_math": "fp32 CUDA cores" if (args.precision == 0 or args.model == "femnist")
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


def main():
    global _REAL_STDOUT
    args = parse()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)                       # everything else that writes to fd 1 now goes to stderr
    if args.impl == "reference":
        return run_reference(args)

    import math
    import numpy as np
    import torch
    import dls_b200  # noqa: F401
    from dls_b200 import _lib, synthetic
    from dls_b200.dataset import DeviceDataset
    from dls_b200.executor import WaveExecutor
    from dls_b200.layout import GNLENET, FEMNIST
    from dls_b200.settings import SessionSettings, LearningSettings
    from dls_b200.solver import BatchedSolver
    L = GNLENET if args.model == "gnlenet" else FEMNIST
    global GNLENET_P, GFLOP_PER_NODE_STEP
    if args.model == "femnist":
        GNLENET_P, GFLOP_PER_NODE_STEP = FEMNIST.P, 2.321

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

    n_nodes = args.nodes_per_gpu * world
    k = int(math.floor(math.log2(n_nodes)))
    if (n_nodes * k) % 2:
        k += 1
    W, Ksteps = args.warmup, args.steps
    rounds = W + Ksteps
    if args.model == "femnist":      # reference defaults for femnist: lr 0.004, momentum 0 (args.py:70-72)
        st = SessionSettings(participants=n_nodes, seed=42, dataset="femnist", model=None,
                             data_dir=os.path.join(REPO, "gpurun_out", "bench_data"),
                             learning=LearningSettings(0.004, 0.0, 0.0, BATCH, LOCAL_STEPS), k=k, precision=0)
    else:
        st = SessionSettings(participants=n_nodes, seed=42, data_dir=os.path.join(REPO, "gpurun_out", "bench_data"),
                             learning=LearningSettings(0.002, 0.9, 0.0, BATCH, LOCAL_STEPS), k=k,
                             precision=args.precision)
    nbrs = synthetic.kreg_topology(n_nodes, k, seed=42)
    if world > 1:
        from dls_b200.partition import DistributedSolver, partition_nodes
        owner = partition_nodes(nbrs, world)
        owned = [int(x) for x in np.nonzero(owner == rank)[0]]
        assert len(owned) == args.nodes_per_gpu
    else:
        owner, owned = None, list(range(n_nodes))
    ds = DeviceDataset.synthetic(L, n_nodes, N_I, 256, dev, seed=42, node_ids=owned if world > 1 else None)

    def make_solver(n_rounds, name_offset=0, ex=None, dag=None):
        if dag is None:
            dag = synthetic.dpsgd_dag(n_nodes, n_rounds, nbrs, name_offset=name_offset)
        if ex is None:
            ex = WaveExecutor(st, ds, device=dev, node_ids=owned if world > 1 else None)
        if world > 1:
            return DistributedSolver(ex, dag, rank, world, owner, agg_variant=args.agg_variant), ex
        return BatchedSolver(ex, dag, agg_variant=args.agg_variant), ex

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- device-resident leg: plan everything first, then time K rounds -----------
    solver, ex = make_solver(rounds)
    solver.plan()
    waves = solver.waves
    per_round = len(waves) // rounds
    assert per_round * rounds == len(waves)
    sampler = ClockSampler(local_rank)
    lib = _lib.lib()

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
    barrier()
    sampler.start()
    time.sleep(0.3)
    launches0 = lib.dls_launch_count()
    events = []
    t_wall0 = time.time()
    s0 = torch.cuda.Event(enable_timing=True); s1 = torch.cuda.Event(enable_timing=True)
    barrier()
    s0.record()
    run_waves(waves[W * per_round:], events)
    s1.record()
    barrier()
    t_wall1 = time.time()
    launches = lib.dls_launch_count() - launches0
    clocks = sampler.stop(t_wall0, t_wall1)
    total_ms = max_over_ranks(s0.elapsed_time(s1))
    train_ms = [e0.elapsed_time(e1) for f, e0, e1 in events if f == "train"]
    agg_ms = [e0.elapsed_time(e1) for f, e0, e1 in events if f == "aggregate"]
    node_steps_per_round = n_nodes * LOCAL_STEPS
    value = node_steps_per_round * Ksteps / (total_ms / 1e3)
    pk = peaks()
    agg_bytes = 2 * args.nodes_per_gpu * GNLENET_P * 4            # per GPU per launch
    agg_t = max_over_ranks(statistics.mean(agg_ms))
    train_t = max_over_ranks(statistics.mean(train_ms))
    agg_gbs = agg_bytes / 1e9 / (agg_t / 1e3)
    train_tflops = args.nodes_per_gpu * LOCAL_STEPS * GFLOP_PER_NODE_STEP / 1e3 / (train_t / 1e3)
    solver.finish()

    # ---------------- end-to-end leg: host task list -> plan -> run -> loss read-back -----------
    # Every round: the solver is built from the round's serialized task list (host objects), plans
    # and uploads its descriptors + minibatch index (pinned staging), runs both waves and copies the
    # train losses back into pinned host memory.  Two timings of the same call sequence:
    #   sync      — one round at a time, host waits for the loss before it looks at the next list
    #   pipelined — the host builds/queues round i+1 while round i runs (loss copies are asynchronous,
    #               all of them complete inside the timed region); this is the throughput figure
    n_sync = min(Ksteps, 5)
    n_e2e = W + n_sync + Ksteps
    # the per-round task lists are the INPUT (what the coordinator sends); they are built beforehand
    round_dags = [synthetic.dpsgd_dag(n_nodes, 1, nbrs, name_offset=i * n_nodes) for i in range(n_e2e)]
    host_losses = [torch.zeros((LOCAL_STEPS, args.nodes_per_gpu), dtype=torch.float32).pin_memory()
                   for _ in range(n_e2e)]
    state = {"ex": None, "prev": None, "h2d": 0}

    def e2e_round(i):
        s2, state["ex"] = make_solver(1, ex=state["ex"], dag=round_dags[i])   # host-side task list -> solver
        if state["prev"] is not None:
            s2.seed_models(state["prev"])                            # models stay in HBM between rounds
        loss = s2.run_with_loss()                                    # streams: plans level k+1 while level k runs
        host_losses[i].copy_(loss, non_blocking=True)                # D2H result read into pinned memory
        state["prev"] = s2.final_slots()
        s2.finish()
        state["h2d"] = s2.h2d_bytes

    prof = None
    if os.environ.get("DLS_PROFILE_E2E") and rank == 0:              # host-side profile of the e2e loop (stderr)
        import cProfile
        prof = cProfile.Profile()
    sync_ms = []
    for i in range(W + n_sync):
        barrier()
        t0 = time.perf_counter()
        e2e_round(i)
        barrier()                                                    # loss is on the host
        if i >= W:
            sync_ms.append(max_over_ranks((time.perf_counter() - t0) * 1e3))
    barrier()
    if prof is not None:
        prof.enable()
    t0 = time.perf_counter()
    for i in range(W + n_sync, n_e2e):
        e2e_round(i)
    barrier()                                                        # every round's loss is on the host
    e2e_ms_per_step = max_over_ranks((time.perf_counter() - t0) * 1e3) / Ksteps
    if prof is not None:
        prof.disable()
        import pstats
        pstats.Stats(prof, stream=sys.stderr).sort_stats("tottime").print_stats(30)
    assert all(bool(torch.isfinite(h).all()) and float(h.abs().sum()) > 0 for h in host_losses), "e2e losses"
    h2d, d2h = state["h2d"], host_losses[0].numel() * 4
    e2e_value = node_steps_per_round / (e2e_ms_per_step / 1e3)

    if rank == 0:
        line = {
            "metric": "node-train-steps/sec", "value": value, "unit": "node-train-steps/s", "n_gpus": world,
            "steps": Ksteps, "warmup": W, "ms_per_step": total_ms / Ksteps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32" if (args.precision == 0 or args.model == "femnist") else "bf16",
            "data": "synthetic", "config": workload_config(n_nodes, k, args),
            "aggregate": {"metric": "gossip-aggregate HBM GB/s", "value": agg_gbs * world, "unit": "GB/s",
                          "ms_per_wave": agg_t, "bytes_per_wave_per_gpu": agg_bytes,
                          "frac_of_hbm_peak": agg_gbs / pk["hbm_gbs"]},
            "train": {"ms_per_wave": train_t, "tflops_gemm": train_tflops * world,
                      "node_train_steps_per_s_train_only": node_steps_per_round / (train_t / 1e3)},
            "roofline": {"kernel": "agg_spmm (gossip aggregate wave)", "bound": "hbm", "achieved": agg_gbs,
                         "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": agg_gbs / pk["hbm_gbs"],
                         # dram__bytes_read.sum + dram__bytes_write.sum of agg_staged_uniform_kernel<16,9>, one launch,
                         # ncu --set full (profiles/r01_ncu_summary_final.md): 87.46 MB read + 36.01 MB written
                         # (N = 256 GNLeNet only; most of Y is still in the 126 MB L2 when the kernel ends, so
                         # DRAM traffic is BELOW the 174.8 MB algorithmic bytes: no wasted re-reads)
                         "traffic": 123472128 if (args.model == "gnlenet" and args.nodes_per_gpu == 256) else None,
                         "peak_source": pk["source"] + " (burst copy bandwidth, MEASURED_PEAKS.json)"},
            "roofline_train": {"kernel": "train wave (all kernels of %d local steps)" % LOCAL_STEPS, "bound": "tensor",
                               "achieved": train_tflops, "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
                               "frac": train_tflops / pk["bf16_tflops_sustained"], "traffic": None,
                               "peak_source": pk["source"] + " (sustained bf16 dense; kernel math: " +
                                              ("fp32 CUDA cores" if args.precision == 0 else "bf16 tcgen05") + ")"},
            "e2e": {"value": e2e_value, "unit": "node-train-steps/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms_per_step,
                    "mode": "pipelined: %d rounds back to back through BatchedSolver.run_with_loss(); the host "
                            "builds and queues round i+1 while round i runs; every round's descriptors go H2D from "
                            "pinned staging and its losses come back D2H inside the timed region" % Ksteps,
                    "sync_ms_per_step": statistics.mean(sync_ms),
                    "sync_value": node_steps_per_round / (statistics.mean(sync_ms) / 1e3)},
            "gpu_launches": int(launches), "clocks": clocks,
        }
        if world > 1:
            line["halo"] = {"bytes_per_gpu_per_wave": int(solver.halo_bytes // max(1, rounds)),
                            "rows_per_gpu_per_wave": int(solver.max_halo),
                            "transport": "NCCL all_to_all_single over NVLink, rows land in the tail of X",
                            "note": "aggregate ms_per_wave includes the exchange"}
        if not args.no_cpu_baseline and world == 1:
            cores = os.cpu_count() or 1
            v, cpu_agg_gbs, secs, steps, t_train, t_agg = cpu_sample(24, k + 1, 1, 4)
            line["cpu_baseline"] = {"value": v, "unit": "node-train-steps/s", "cores": 4, "kind": "port",
                                    "agg_gbs": cpu_agg_gbs,
                                    "sample": "24 of %d nodes: their train tasks (%d steps, B=%d) + aggregates "
                                              "(fan-in %d), 1 worker x 4 torch threads (reference defaults), "
                                              "%.1f s; host has %d cores" % (n_nodes, LOCAL_STEPS, BATCH, k + 1,
                                                                            secs, cores)}
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
