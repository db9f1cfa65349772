"""Batched compute-graph solver: the drop-in replacement for the broker's ready-queue loop +
worker processes (`Broker.handle_task_result` dasklearn/broker.py:277-290, `Worker.start`
dasklearn/worker.py:21-38).

Input is exactly what a broker receives in the coordinator's `config` message: the serialized
task list (`WorkflowDAG.serialize()`, dasklearn/tasks/dag.py:28-32; task dict format
dasklearn/tasks/task.py:59-66).  Instead of dispatching one ready task at a time, the solver
levelises the DAG into wavefronts, groups every wavefront by task function and runs each group
as ONE batched GPU wave on rows of the stacked parameter tensor.

`plan()` does all host-side work once (levelisation, slot assignment with reference counts,
CSR construction, minibatch index); `run()` then only launches.  Without a prior `plan()`,
`run()` / `run_with_loss()` stream the graph: level k+1 is planned on the host while level k
executes on the GPU, so only the first level's planning is exposed.
"""
from collections import Counter, defaultdict
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ._lib import DlsError
from .executor import RowHandle, WaveExecutor

TRAIN_FUNCS = ("train",)
SUPPORTED = ("train", "aggregate", "test", "compute_gradient", "gradient_update", "chunk", "reconstruct_from_chunks")


class Wave:
    __slots__ = ("level", "func", "names", "peers", "rounds", "times", "steps", "in_slots", "out_slots",
                 "model_slots", "weights", "init_slots", "copy_src", "copy_dst", "release_after", "prepared",
                 "xchg", "grad_slots", "segments")

    def __init__(self, level: int, func: str):
        self.level, self.func = level, func
        self.names: List[str] = []
        self.peers: List[int] = []
        self.rounds: List[int] = []
        self.times: List[int] = []
        self.steps: List[int] = []
        self.in_slots: List[int] = []
        self.out_slots: List[int] = []
        self.model_slots: List[List[int]] = []
        self.weights: List[Optional[List[float]]] = []
        self.init_slots: List[int] = []
        self.copy_src: List[int] = []
        self.copy_dst: List[int] = []
        self.release_after: List[int] = []
        self.prepared = None
        self.xchg = None            # multi-GPU: halo exchange that precedes this wave
        self.grad_slots: List[int] = []   # compute_gradient: row receiving the gradient; gradient_update: row read
        self.segments: List[Tuple[int, int, int, List[int]]] = []   # reconstruct_from_chunks: (out slot, col0, len, contributor slots)


class BatchedSolver:
    def __init__(self, executor: WaveExecutor, serialized_tasks: Sequence[Dict],
                 on_sink: Optional[Callable[[str, List[RowHandle]], None]] = None,
                 owned_peers: Optional[Sequence[int]] = None, agg_variant: int = 0):
        self.ex = executor
        self.tasks: Dict[str, Dict] = {}
        for t in serialized_tasks:
            if t["func"] not in SUPPORTED:
                # same failure mode as the worker: unknown function name (dasklearn/worker.py:26-27)
                raise RuntimeError("Task function %s not found!" % t["func"])
            self.tasks[t["name"]] = t
        self.on_sink = on_sink
        self.agg_variant = agg_variant
        self.owned = set(int(p) for p in owned_peers) if owned_peers is not None else None
        self.waves: List[Wave] = []
        self.sink_slots: Dict[str, int] = {}
        self.wave_times: List[Tuple[str, int, float]] = []   # (func, n_tasks, ms) when timing is on
        self.planned = False
        self.initial_models: Dict[int, int] = {}              # peer -> slot holding its current model
        self.h2d_bytes = 0
        self.last_train_loss = None

    def seed_models(self, slots_by_peer: Dict[int, int]):
        """Continue from models already in HBM: a `train` task whose model is `(None, 0)` starts
        from `slots_by_peer[peer]` instead of the initial model (used to solve a long run one
        round-graph at a time)."""
        self.initial_models = dict(slots_by_peer)
        return self

    def final_slots(self) -> Dict[int, int]:
        return {int(self.tasks[name]["data"]["peer"]): slot for name, slot in self.sink_slots.items()}

    # -- planning ---------------------------------------------------------------------------
    def levelise(self) -> List[List[str]]:
        """Wavefronts: level(t) = 1 + max(level of inputs); source tasks are level 0
        (`get_source_tasks`, dasklearn/tasks/dag.py:15-19).  The serialized list of the reference is in creation
        order, which is topological (a consumer names producers that already exist), so one pass in list order
        settles every level; any other order falls back to Kahn's algorithm."""
        level: Dict[str, int] = {}
        get = level.__getitem__
        try:
            for n, t in self.tasks.items():
                ins = t["inputs"]
                level[n] = 1 + max(map(get, next(zip(*ins)))) if ins else 0
        except KeyError:
            return self._levelise_kahn()
        self._level_of = level
        levels: List[List[str]] = [[] for _ in range(max(level.values()) + 1)] if level else []
        for n, l in level.items():
            levels[l].append(n)
        return levels

    def _levelise_kahn(self) -> List[List[str]]:
        tasks = self.tasks
        indeg = {n: len(t["inputs"]) for n, t in tasks.items()}
        frontier = [n for n, d in indeg.items() if d == 0]
        levels: List[List[str]] = []
        seen = 0
        while frontier:
            levels.append(frontier)
            seen += len(frontier)
            nxt: List[str] = []
            push = nxt.append
            for n in frontier:
                for out in tasks[n]["outputs"]:
                    m = out[0]
                    d = indeg[m] - 1
                    indeg[m] = d
                    if d == 0:
                        push(m)
            frontier = nxt
        if seen != len(self.tasks):
            raise RuntimeError("compute graph has a cycle or dangling inputs (%d of %d tasks reachable)"
                               % (seen, len(self.tasks)))
        self._level_of = {n: li for li, names in enumerate(levels) for n in names}
        return levels

    def _plan_levels(self):
        """Generator: assigns slots for one wavefront at a time and yields that level's waves.
        Nothing touches the device here; `_commit` does (pool growth, descriptor upload)."""
        ex = self.ex
        ls = ex.settings.learning
        value_slot: Dict[Tuple[str, int], int] = {}
        sim_refs: Dict[int, int] = defaultdict(int)     # simulated reference counts during planning
        live = np.nonzero(ex.refs)[0]
        sim_refs.update(zip(live.tolist(), ex.refs[live].tolist()))
        free: List[int] = list(ex.free)
        self._planned_slots = ex.n_slots
        self._plan_free, self._plan_refs = free, sim_refs
        tasks, owned, initial = self.tasks, self.owned, self.initial_models
        default_steps = int(ls.local_steps)

        def take() -> int:
            if not free:
                free.append(self._planned_slots)
                self._planned_slots += 1
            return free.pop()

        self._value_slot = value_slot
        grad_slot: Dict[Tuple[str, int], int] = {}        # outputs of compute_gradient tasks also carry a gradient row
        self._grad_slot = grad_slot
        chunk_range: Dict[Tuple[str, int], Tuple[int, int]] = {}   # outputs of chunk tasks: column range of the row

        def chunk_bounds(n: int):
            """ChunkManager.chunk_model (simulation/conflux/chunk_manager.py:13-25): n slices of P // n elements,
            the last one takes the remainder."""
            P = ex.layout.P
            size = P // n
            b = [(i * size, (i + 1) * size) for i in range(n)]
            b[-1] = (b[-1][0], P)
            return b

        def ref_of(r):
            """Placeholders are (task name, output index) tuples; the AD-PSGD client also emits bare task
            names (dasklearn/simulation/adpsgd/client.py:55-57, simulation/client.py:112-114)."""
            if r is None:
                return None
            if isinstance(r, str):
                return (r, 0)
            r = tuple(r)
            return None if r[0] is None else r

        def consume(ref, pending):
            """One consuming edge of `ref`: its model row and, if it has one, its gradient row."""
            pending.append(value_slot[ref])
            if ref in grad_slot:
                pending.append(grad_slot[ref])

        # single-GPU solve: every model reference is a pool slot (no halo rows) -> plain dictionary lookups
        local_only = type(self)._model_slot is BatchedSolver._model_slot
        model_slot = value_slot.__getitem__ if local_only else value_slot.get
        for li, names in enumerate(self.levelise()):
            by_func: Dict[str, Wave] = {}
            # a level may hold two train tasks of one node only in malformed graphs; detect in executor
            pending_release: List[int] = []
            self._begin_level(li, names, pending_release)
            for name in names:
                t = tasks[name]
                d = t["data"]
                peer = int(d["peer"])
                if owned is not None and peer not in owned:
                    continue
                func = t["func"]
                w = by_func.get(func)
                if w is None:
                    w = by_func[func] = Wave(li, func)
                n_cons = len(t["outputs"])
                keep = n_cons if n_cons > 0 else 1          # sinks keep their row as the result
                w.names.append(name)
                w.peers.append(peer)
                w.rounds.append(int(d.get("round", 0)))
                w.times.append(int(d.get("time", 0)))
                if func == "train":
                    src = ref_of(d["model"])
                    w.steps.append(int(d.get("local_steps", default_steps)))
                    if src is None and peer in initial:
                        slot = initial[peer]                 # resume: train in place
                        sim_refs[slot] -= 1
                    elif src is None:                        # first round: (None, 0)
                        slot = take()
                        w.init_slots.append(slot)
                    else:
                        s_in = value_slot[src]
                        if sim_refs[s_in] == 1:              # sole consumer: train in place
                            slot = s_in
                            sim_refs[s_in] = 0
                            if src in grad_slot:
                                pending_release.append(grad_slot[src])
                        else:
                            slot = take()
                            w.copy_src.append(s_in)
                            w.copy_dst.append(slot)
                            consume(src, pending_release)
                    w.in_slots.append(slot)
                    w.out_slots.append(slot)
                    sim_refs[slot] += keep
                    value_slot[(name, 0)] = slot
                elif func == "aggregate":
                    refs = d["models"]
                    try:
                        ms = list(map(model_slot, refs))             # placeholders are tuples after pickle
                    except TypeError:                                # lists (e.g. a JSON round trip)
                        refs = [tuple(m) for m in refs]
                        ms = list(map(model_slot, refs))
                    if not local_only and None in ms:                # rows of other GPUs: halo references (negative)
                        halo = self._halo_of
                        ms = [-1 - halo[m] if x is None else x for x, m in zip(ms, refs)]
                    slot = take()
                    w.model_slots.append(ms)
                    wts = d.get("weights")
                    w.weights.append(list(wts) if wts else None)
                    w.out_slots.append(slot)
                    sim_refs[slot] += keep
                    value_slot[(name, 0)] = slot
                    if local_only:
                        pending_release.extend(ms)
                    else:
                        pending_release.extend([x for x in ms if x >= 0])
                    if grad_slot:
                        pending_release.extend([grad_slot[m] for m in map(tuple, refs) if m in grad_slot])
                elif func == "compute_gradient":
                    # functions.py:80-82: forward/backward on the first minibatch, no optimizer step; the
                    # result is the UNCHANGED model (an alias of the input row) plus a gradient row
                    src = ref_of(d["model"])
                    if src is None and peer in initial:
                        s_in = initial[peer]
                        sim_refs[s_in] -= 1
                    elif src is None:
                        s_in = take()
                        w.init_slots.append(s_in)
                    else:
                        s_in = value_slot[src]
                        consume(src, pending_release)        # after the alias below has taken its references
                    g = take()
                    w.in_slots.append(s_in)
                    w.out_slots.append(s_in)
                    w.grad_slots.append(g)
                    sim_refs[s_in] += keep
                    sim_refs[g] += keep
                    value_slot[(name, 0)] = s_in
                    grad_slot[(name, 0)] = g
                elif func == "gradient_update":
                    # functions.py:84-86: one step of a fresh SGD optimizer on `model` with the gradient
                    # carried by `gradient_model` (possibly computed on an older model)
                    src, gsrc = ref_of(d["model"]), ref_of(d["gradient_model"])
                    if gsrc not in grad_slot:
                        raise RuntimeError("gradient_update task %s: %r carries no gradient" % (name, gsrc))
                    s_in, g = value_slot[src], grad_slot[gsrc]
                    if sim_refs[s_in] == 1 and gsrc != src:  # sole consumer of the model: update in place
                        slot = s_in
                        sim_refs[s_in] = 0
                        if src in grad_slot:
                            pending_release.append(grad_slot[src])
                    else:
                        slot = take()
                        consume(src, pending_release)
                    if gsrc != src:                          # a second edge (simulation/client.py:112-115)
                        consume(gsrc, pending_release)
                    w.in_slots.append(s_in)
                    w.out_slots.append(slot)
                    w.grad_slots.append(g)
                    sim_refs[slot] += keep
                    value_slot[(name, 0)] = slot
                elif func == "chunk":
                    # functions.py:136-140: the n pieces are views of the flat model = column ranges of its row;
                    # output idx i aliases the input row, one reference per consuming edge
                    src = ref_of(d["model"])
                    s_in = value_slot[src]
                    bounds = chunk_bounds(int(d["n"]))
                    for ci, b in enumerate(bounds):
                        value_slot[(name, ci)] = s_in
                        chunk_range[(name, ci)] = b
                    w.in_slots.append(s_in)
                    w.out_slots.append(s_in)
                    sim_refs[s_in] += keep
                    consume(src, pending_release)
                elif func == "reconstruct_from_chunks":
                    # functions.py:142-146: per chunk index the mean over the pieces received for it
                    lists = [[tuple(m) for m in c] for c in d["chunks"]]
                    bounds = chunk_bounds(len(lists))
                    slot = take()
                    for ci, refs in enumerate(lists):
                        if not refs:
                            raise AssertionError("No chunks received at index %d!" % ci)     # chunk_manager.py:36
                        for m in refs:
                            if chunk_range.get(m) != bounds[ci]:
                                raise RuntimeError("task %s: piece %r is not chunk %d of %d" % (name, m, ci, len(lists)))
                        w.segments.append((slot, bounds[ci][0], bounds[ci][1] - bounds[ci][0],
                                           [value_slot[m] for m in refs]))
                        pending_release.extend(value_slot[m] for m in refs)
                    w.out_slots.append(slot)
                    sim_refs[slot] += keep
                    value_slot[(name, 0)] = slot
                else:                                        # test: identity on the model data path
                    src = ref_of(d["model"])
                    s_in = value_slot[src]
                    w.in_slots.append(s_in)
                    w.out_slots.append(s_in)
                    sim_refs[s_in] += keep                   # alias first, then consume the input edge
                    value_slot[(name, 0)] = s_in
                    if src in grad_slot:                     # the alias keeps the gradient reachable too
                        grad_slot[(name, 0)] = grad_slot[src]
                        sim_refs[grad_slot[src]] += keep
                    consume(src, pending_release)
                if n_cons == 0:
                    self.sink_slots[name] = value_slot[(name, 0)]
            # inputs are released only after the whole level ran (their rows are read by this level)
            for s, c in Counter(pending_release).items():      # one visit per distinct row, not per edge
                n = sim_refs[s] - c
                assert n >= 0
                sim_refs[s] = n
                if n == 0:
                    free.append(s)
            level_waves = [by_func[f] for f in ("train", "compute_gradient", "gradient_update", "aggregate",
                                                "reconstruct_from_chunks", "chunk", "test") if f in by_func]
            for w in level_waves:
                self._end_level_wave(li, w)
            yield self._level_waves(li, level_waves)

    def plan(self):
        """Whole-graph planning: every level's slots and descriptors before the first launch."""
        for level_waves in self._plan_levels():
            self.waves.extend(level_waves)
        self._commit(self.waves)
        self.planned = True
        return self

    def _commit(self, waves: List[Wave]):
        """Device side of planning for `waves`: grow the pool to the planned size, then upload
        the wave descriptors (pinned staging, asynchronous)."""
        self._grow_pool(self._planned_slots - self.ex.n_slots)
        for w in waves:
            self._fixup_wave(w)
            self._prepare(w)

    def _prepare(self, w: Wave):
        ex = self.ex
        if w.func == "train":
            w.prepared = ex.prepare_train(w.peers, w.rounds, w.steps)
            self.h2d_bytes += sum(t.numel() * t.element_size() for t in w.prepared[:2] if hasattr(t, 'numel'))
        elif w.func == "aggregate":
            w.prepared = ex.prepare_aggregate(w.model_slots, w.out_slots, w.weights)
            self.h2d_bytes += sum(t.numel() * t.element_size() for t in w.prepared[:5] if hasattr(t, 'numel'))

    def _stream(self):
        """Waves in execution order.  Unplanned solvers plan level k+1 on the host while level k
        runs on the GPU (launches and descriptor uploads are asynchronous)."""
        if self.planned:
            yield from self.waves
            return
        for level_waves in self._plan_levels():
            self._commit(level_waves)
            for w in level_waves:
                self.waves.append(w)
                yield w
        self.planned = True

    # hooks overridden by the multi-GPU solver (partition.DistributedSolver)
    def _begin_level(self, level: int, names: List[str], pending_release: List[int]):
        pass

    def _model_slot(self, ref: Tuple[str, int]) -> int:
        return self._value_slot[ref]

    def _end_level_wave(self, level: int, wave: Wave):
        pass

    def _level_waves(self, level: int, waves: List[Wave]) -> List[Wave]:
        return waves

    def _grow_pool(self, extra_slots: int):
        if extra_slots:
            self.ex._grow(extra_slots)

    def _fixup_wave(self, wave: Wave):
        pass

    # -- execution --------------------------------------------------------------------------
    def run_wave(self, w: Wave, loss_out=None):
        ex = self.ex
        if w.func == "train":
            ex.init_rows(w.init_slots)
            ex.copy_rows(w.copy_src, w.copy_dst)
            ex.train_wave(w.out_slots, w.peers, w.rounds, w.steps, prepared=w.prepared, loss_out=loss_out)
        elif w.func == "aggregate":
            ex.aggregate_wave(w.model_slots, w.out_slots, w.weights, variant=self.agg_variant,
                              prepared=w.prepared)
        elif w.func == "compute_gradient":
            ex.init_rows(w.init_slots)
            ex.grad_wave(w.in_slots, w.grad_slots, w.peers, w.rounds)
        elif w.func == "gradient_update":
            ex.gradient_update_wave(w.in_slots, w.grad_slots, w.out_slots)
        elif w.func == "reconstruct_from_chunks":
            ex.reconstruct_wave(w.segments)
        elif w.func == "chunk":
            pass                                  # pieces are column ranges of the model's row: nothing to compute
        else:
            acc, loss = ex.test_wave(w.out_slots)
            for i, name in enumerate(w.names):
                ex.write_accuracy(w.peers[i], w.rounds[i], w.times[i], float(acc[i]), float(loss[i]))

    def finish(self):
        """Commit the planned pool state (free list, reference counts) to the executor."""
        ex = self.ex
        ex.free = list(self._plan_free)
        ex.refs[:] = 0
        for s, n in self._plan_refs.items():
            if n > 0:
                ex.refs[s] = n

    def run(self, timing: bool = False):
        """Solve the graph.  `timing`: per-wave durations in `wave_times` as (func, n_tasks, ms) — CUDA events on the
        launching stream (host clock for a CPU stand-in executor)."""
        import time as _time
        ex = self.ex
        on_gpu = getattr(ex.device, "type", "cpu") == "cuda"
        results: Dict[str, List[RowHandle]] = {}
        for w in self._stream():
            if timing:
                if on_gpu:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                else:
                    e0 = _time.perf_counter()
            self.run_wave(w)
            if timing:
                if on_gpu:
                    e1.record()
                else:
                    e1 = _time.perf_counter()
                self.wave_times.append((w.func, len(w.names), e0, e1))
            for name in w.names:
                if name in self.sink_slots:
                    res = [RowHandle(ex, self.sink_slots[name])]
                    results[name] = res
                    if self.on_sink:
                        self.on_sink(name, res)
        self.finish()
        if timing:
            if on_gpu:
                torch.cuda.synchronize()
                self.wave_times = [(f, n, e0.elapsed_time(e1)) for f, n, e0, e1 in self.wave_times]
            else:
                self.wave_times = [(f, n, (e1 - e0) * 1e3) for f, n, e0, e1 in self.wave_times]
        return results

    def run_with_loss(self):
        """Run all waves; returns the device tensor `[steps, n_tasks]` of mean minibatch losses of
        the last train wave (the step result a caller reads back)."""
        loss = None
        for w in self._stream():
            if w.func == "train":
                loss = torch.zeros((max(w.steps), len(w.names)), dtype=torch.float32, device=self.ex.device)
                self.run_wave(w, loss_out=loss)
            else:
                self.run_wave(w)
        return loss


def solve(settings, serialized_tasks, dataset, on_sink=None, device=None, agg_variant: int = 0):
    """One-call form: build the executor, plan and run the graph, return {sink task: [RowHandle]}."""
    ex = WaveExecutor(settings, dataset, device=device)
    solver = BatchedSolver(ex, serialized_tasks, on_sink=on_sink, agg_variant=agg_variant)
    return solver.plan().run(), ex
