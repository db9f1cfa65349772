"""Multi-GPU solve: nodes are partitioned over the ranks of one box (one process per GPU); train
and test waves need no communication; an aggregate wave first receives the neighbour rows that
live on other GPUs ("halo") and then runs the same SpMM on pool + halo rows.

The reference's analogue is the modulo client->broker map plus pickled `nn.Module` results sent
broker to broker over ZeroMQ (`distribute_brokers` dasklearn/simulation/simulation.py:97-111,
result routing dasklearn/broker.py:207-219); here rows move GPU to GPU over NVLink through one
`all_to_all_single` per aggregate wave (NCCL; gloo in the CPU tests).

Every rank holds the whole serialized DAG (as every broker does), so send and receive lists are
derived locally and identically on all ranks from the `models` lists — the DAG is authoritative
for topology, arrival order and churn.
"""
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .solver import BatchedSolver, Wave


def partition_nodes(in_neighbours: Sequence[Sequence[int]], world: int) -> np.ndarray:
    """owner[node] in [0, world): greedy graph growing — every part is grown breadth-first inside
    the still unassigned nodes, seeded next to the parts already built.  Keeps neighbourhoods
    together: exact for rings (every part is one arc: 2 remote rows per GPU even after the
    reference's random relabelling, dpsgd/simulation.py:37-41), a mild gain for sparse graphs;
    random regular graphs are expanders — about (world-1)/world of their edges cross whatever we do."""
    n = len(in_neighbours)
    adj: List[set] = [set() for _ in range(n)]
    for i, nb in enumerate(in_neighbours):
        for j in nb:
            adj[i].add(int(j)); adj[int(j)].add(i)
    owner = -np.ones(n, dtype=np.int32)
    per = (n + world - 1) // world
    assigned_order: List[int] = []
    for rank in range(world):
        target = per if rank < world - 1 else n - len(assigned_order)
        target = min(target, n - len(assigned_order))
        size = 0
        frontier: List[int] = []
        while size < target:
            if not frontier:
                # seed: an unassigned node touching an assigned one (latest parts first), else lowest id
                seed = -1
                for u in reversed(assigned_order):
                    cand = sorted(v for v in adj[u] if owner[v] < 0)
                    if cand:
                        seed = cand[0]
                        break
                if seed < 0:
                    seed = int(np.nonzero(owner < 0)[0][0])
                owner[seed] = rank
                assigned_order.append(seed)
                size += 1
                frontier = [seed]
                continue
            nxt: List[int] = []
            for u in frontier:
                for v in sorted(adj[u]):
                    if owner[v] < 0 and size < target:
                        owner[v] = rank
                        assigned_order.append(v)
                        size += 1
                        nxt.append(v)
            frontier = nxt
    return owner


def block_partition(n: int, world: int) -> np.ndarray:
    per = (n + world - 1) // world
    return np.minimum(np.arange(n) // per, world - 1).astype(np.int32)


def topology_from_dag(serialized_tasks: Sequence[Dict], n: int) -> List[List[int]]:
    """In-neighbour lists read off the first aggregate task of every peer."""
    peer_of = {t["name"]: int(t["data"]["peer"]) for t in serialized_tasks}
    nb: List[Optional[List[int]]] = [None] * n
    for t in serialized_tasks:
        if t["func"] == "aggregate":
            p = int(t["data"]["peer"])
            if nb[p] is None:
                nb[p] = [peer_of[m[0]] for m in t["data"]["models"] if peer_of[m[0]] != p]
    return [x or [] for x in nb]


class DistributedSolver(BatchedSolver):
    def __init__(self, executor, serialized_tasks, rank: int, world: int, owner_of: Sequence[int],
                 on_sink=None, agg_variant: int = 0, group=None):
        owner_of = np.asarray(owner_of, dtype=np.int32)
        owned = [int(p) for p in np.nonzero(owner_of == rank)[0]]
        super().__init__(executor, serialized_tasks, on_sink=on_sink, owned_peers=owned, agg_variant=agg_variant)
        self.rank, self.world, self.owner_of, self.group = rank, world, owner_of, group
        self._halo_of: Dict[Tuple[str, int], int] = {}
        self._level_xchg: Dict[int, dict] = {}
        self.max_halo = 0
        self.max_send = 0
        self.halo_bytes = 0

    # -- planning hooks ---------------------------------------------------------------------
    # Every rank holds the whole task list, but its planning work is O(own tasks + cut edges): task owners
    # and levels are looked up in dictionaries built once, and the rows this rank must SEND are found from
    # the `outputs` lists of its own tasks instead of scanning every other rank's aggregate tasks.
    def _owner_of_task(self, name: str) -> int:
        return self._task_owner[name]

    def levelise(self):
        levels = super().levelise()
        tasks, me = self.tasks, self.rank
        owner_list = self.owner_of.tolist()       # plain ints: numpy scalar indexing is ~10x slower per lookup
        self._task_owner = {name: owner_list[t["data"]["peer"]] for name, t in tasks.items()}
        own = self._task_owner
        level_of = {}
        for li, names in enumerate(levels):
            for n in names:
                level_of[n] = li
        remote = {}                               # level -> [(consumer rank, my value consumed there)], one per edge
        # Levels at which ANY rank has a cut edge: the exchange is a collective over the whole group, so every
        # rank must enter it at exactly these levels, also when it neither sends nor receives there itself
        # (irregular graphs: async gossip / AD-PSGD levels, world > 2).
        xchg_levels = set()
        for name, t in tasks.items():
            mine = own[name]
            if mine != me:
                for cons, _ in t["outputs"]:
                    if own[cons] != mine:
                        xchg_levels.add(level_of[cons])
                continue
            for src, _ in t["inputs"]:
                if own[src] != me and t["func"] != "aggregate":
                    raise RuntimeError("task %s consumes a model of another GPU's node outside an aggregate" % name)
            for cons, idx in t["outputs"]:
                q = own[cons]
                if q != me:
                    if tasks[cons]["func"] != "aggregate":
                        raise RuntimeError("task %s consumes a model of another GPU's node outside an aggregate" % cons)
                    remote.setdefault(level_of[cons], []).append((q, (name, idx)))
                    xchg_levels.add(level_of[cons])
        self._remote = remote
        self._xchg_levels = xchg_levels
        # the base planner only looks at this rank's tasks
        return [[n for n in names if own[n] == me] for names in levels]

    def _halo_key(self, m):
        return (int(self.tasks[m[0]]["data"]["peer"]), m[0], m[1])

    def _begin_level(self, level: int, names: List[str], pending_release: List[int]):
        me, own, tasks = self.rank, self._task_owner, self.tasks
        recv = [set() for _ in range(self.world)]
        send = [set() for _ in range(self.world)]
        for name in names:                                    # this rank's tasks of the level
            t = tasks[name]
            if t["func"] != "aggregate":
                continue
            for m in t["data"]["models"]:
                m = tuple(m)
                r = own[m[0]]
                if r != me:
                    recv[r].add(m)
        for q, m in self._remote.get(level, ()):              # one reference per consuming remote task
            pending_release.append(self._value_slot[m])
            send[q].add(m)
        # both ends order a (source rank -> destination rank) row list the same way: by producer peer, task, index
        recv_l = [sorted(x, key=self._halo_key) for x in recv]
        send_l = [sorted(x, key=self._halo_key) for x in send]
        self._halo_of = {}
        h = 0
        for r in range(self.world):                           # all_to_all output is ordered by source rank
            for m in recv_l[r]:
                self._halo_of[m] = h
                h += 1
        n_send = sum(len(x) for x in send_l)
        if h or n_send or level in self._xchg_levels:
            self._level_xchg[level] = {"send_slots": [self._value_slot[m] for x in send_l for m in x],
                                       "send_counts": [len(x) for x in send_l],
                                       "recv_counts": [len(x) for x in recv_l]}
            self.max_halo = max(self.max_halo, h)
            self.max_send = max(self.max_send, n_send)

    def _model_slot(self, ref):
        if ref in self._value_slot:
            return self._value_slot[ref]
        return -1 - self._halo_of[ref]                        # halo row, fixed up in _fixup_wave

    def _end_level_wave(self, level: int, wave: Wave):
        if wave.func == "aggregate" and level in self._level_xchg:
            wave.xchg = self._level_xchg[level]

    def _level_waves(self, level: int, waves: List[Wave]) -> List[Wave]:
        # a level where this rank only sends (it owns no aggregate there) still takes part in the exchange
        x = self._level_xchg.get(level)
        if x is not None and not any(getattr(w, "xchg", None) for w in waves):
            w = Wave(level, "exchange")
            w.xchg = x
            waves = waves + [w]
        return waves

    def _grow_pool(self, extra_slots: int):
        self.ex._grow(extra_slots, halo_rows=self.max_halo, send_rows=self.max_send)

    def _fixup_wave(self, w: Wave):
        """Halo references become rows after the pool (valid once the pool has its planned size)."""
        ex = self.ex
        base = ex.halo_base
        if w.func == "aggregate":
            w.model_slots = [[s if s >= 0 else base + (-1 - s) for s in ms] for ms in w.model_slots]
        x = getattr(w, "xchg", None)
        if x and "send_idx" not in x:
            x["send_idx"] = ex._dev_i32(x["send_slots"]) if x["send_slots"] else None
            self.halo_bytes += sum(x["recv_counts"]) * ex.layout.P * 4

    # -- execution --------------------------------------------------------------------------
    def exchange(self, x: dict):
        import torch.distributed as dist
        ex = self.ex
        ld = ex.layout.ld
        n_send, n_recv = sum(x["send_counts"]), sum(x["recv_counts"])
        if n_send:
            sbuf = ex.gather_rows(x["send_idx"], n_send)
        else:
            sbuf = ex.X[0:0]
        rbuf = ex.X[ex.halo_base: ex.halo_base + n_recv]
        dist.all_to_all_single(rbuf.reshape(-1), sbuf.reshape(-1),
                               output_split_sizes=[c * ld for c in x["recv_counts"]],
                               input_split_sizes=[c * ld for c in x["send_counts"]], group=self.group)

    def run_wave(self, w: Wave, loss_out=None):
        x = getattr(w, "xchg", None)
        if x:
            self.exchange(x)
        if w.func != "exchange":
            super().run_wave(w, loss_out=loss_out)
