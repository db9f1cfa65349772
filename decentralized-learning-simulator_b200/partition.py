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
    """`transport`: how rows cross GPUs at an aggregate level.
      "a2a"  — pack (`dls_copy_rows`) + one `all_to_all_single` (NCCL over NVLink; gloo in the CPU tests) into the
               halo rows after the pool;
      "push" — one-sided: every rank writes the rows its neighbours need straight into THEIR halo rows over
               NVLink (`dls_push_rows`, peer-mapped symmetric memory), then one cross-GPU barrier; no pack copy,
               no receive side.  Needs the executor's X in symmetric memory (`WaveExecutor.enable_symmetric`);
      "auto" — "push" on NCCL process groups when symmetric memory can be set up on every rank, else "a2a"."""

    HALO_VERSIONS = 2        # push: rows of one remote node that may be live in a halo at one level

    def __init__(self, executor, serialized_tasks, rank: int, world: int, owner_of: Sequence[int],
                 on_sink=None, agg_variant: int = 0, group=None, transport: str = "auto"):
        owner_of = np.asarray(owner_of, dtype=np.int32)
        owned = [int(p) for p in np.nonzero(owner_of == rank)[0]]
        super().__init__(executor, serialized_tasks, on_sink=on_sink, owned_peers=owned, agg_variant=agg_variant)
        self.rank, self.world, self.owner_of, self.group = rank, world, owner_of, group
        self._halo_of: Dict[Tuple[str, int], int] = {}
        self._level_xchg: Dict[int, dict] = {}
        self.max_halo = 0
        self.max_send = 0
        self.halo_bytes = 0
        self.transport = self._choose_transport(transport)

    def _choose_transport(self, want: str) -> str:
        ex = self.ex
        have = getattr(ex, "transport", None)
        if have is not None:                          # settled once per executor (a collective decision)
            if want not in ("auto", have):
                raise RuntimeError("executor already uses transport %r" % have)
            return have
        if want == "a2a" or not hasattr(ex, "enable_symmetric"):
            got = "a2a"
        else:
            import torch.distributed as dist
            ok = dist.is_initialized() and dist.get_backend(self.group) == "nccl"
            if ok:
                counts = np.bincount(self.owner_of, minlength=self.world)
                n_local_max = int(counts.max())
                halo_cap = self.HALO_VERSIONS * int(len(self.owner_of) - counts.min())
                ok = ex.enable_symmetric(self.group, pool_cap=3 * n_local_max + 8, halo_cap=halo_cap)
            if want == "push" and not ok:
                raise RuntimeError("transport 'push' needs NCCL and symmetric memory on every rank")
            got = "push" if ok else "a2a"
        ex.transport = got
        return got

    # -- planning hooks ---------------------------------------------------------------------
    # Every rank holds the whole task list.  One cheap pass over it settles the levels and the levels at which
    # ANY rank has a cut edge (the exchange is entered by every rank at exactly those levels, also by a rank
    # that neither sends nor receives there: irregular graphs, world > 2); everything else a rank plans is
    # O(own tasks + cut edges): the rows to SEND are found from the `outputs` lists of its own tasks.
    def _owner_of_task(self, name: str) -> int:
        return self._task_owner[name]

    def levelise(self):
        tasks, me = self.tasks, self.rank
        owner_list = self.owner_of.tolist()       # plain ints: numpy scalar indexing is ~10x slower per lookup
        own: Dict[str, int] = {}
        level_of: Dict[str, int] = {}
        own_get, lvl_get = own.__getitem__, level_of.__getitem__
        xchg_levels = set()
        mine: List[str] = []
        try:
            # ONE pass in list order (creation order is topological, see BatchedSolver.levelise): owner, level,
            # and whether the task has an input produced on another GPU
            for name, t in tasks.items():
                o = owner_list[t["data"]["peer"]]
                own[name] = o
                if o == me:
                    mine.append(name)
                ins = t["inputs"]
                if not ins:
                    level_of[name] = 0
                    continue
                if len(ins) == 1:
                    src = ins[0][0]
                    l = level_of[src] + 1
                    cut = own[src] != o
                else:
                    names = next(zip(*ins))
                    l = 1 + max(map(lvl_get, names))
                    # the owners of the inputs only matter until the level is known to be an exchange level
                    if l in xchg_levels and t["func"] == "aggregate":
                        cut = False
                    else:
                        owners = set(map(own_get, names))
                        cut = len(owners) > 1 or o not in owners
                level_of[name] = l
                if cut:
                    if t["func"] != "aggregate":
                        raise RuntimeError("task %s consumes a model of another GPU's node outside an aggregate" % name)
                    xchg_levels.add(l)
            self._level_of = level_of
            n_levels = max(level_of.values()) + 1 if level_of else 0
        except KeyError:                          # not in topological order: generic levelisation, then the cut scan
            own = {name: owner_list[t["data"]["peer"]] for name, t in tasks.items()}
            n_levels = len(super()._levelise_kahn())
            level_of = self._level_of
            mine = [n for n in tasks if own[n] == me]
            xchg_levels = set()
            for name, t in tasks.items():
                if any(own[src] != own[name] for src, _ in t["inputs"]):
                    if t["func"] != "aggregate":
                        raise RuntimeError("task %s consumes a model of another GPU's node outside an aggregate" % name)
                    xchg_levels.add(level_of[name])
        self._task_owner = own
        remote = {}                               # level -> [(consumer rank, my value consumed there)], one per edge
        for name in mine:
            for cons, idx in tasks[name]["outputs"]:
                q = own[cons]
                if q != me:
                    remote.setdefault(level_of[cons], []).append((q, (name, idx)))
        self._remote = remote
        self._xchg_levels = xchg_levels
        self._xchg_index = {l: i for i, l in enumerate(sorted(xchg_levels))}
        if self.transport == "push":
            # static halo numbering: node p's row lands in slot (index of p among the nodes rank q does not own)
            self._remote_index = [None] * self.world
            counts = np.bincount(self.owner_of, minlength=self.world)
            self._n_remote = [int(len(self.owner_of) - c) for c in counts]
        # the base planner only looks at this rank's tasks
        out = [[] for _ in range(n_levels)]
        for n in mine:
            out[level_of[n]].append(n)
        return out

    def _halo_key(self, m):
        return (int(self.tasks[m[0]]["data"]["peer"]), m[0], m[1])

    def _remote_idx(self, q: int) -> np.ndarray:
        """index of node p among the nodes NOT owned by rank q (push transport: halo slot numbering of rank q)."""
        r = self._remote_index[q]
        if r is None:
            not_q = (self.owner_of != q)
            r = (np.cumsum(not_q) - 1).astype(np.int64)
            self._remote_index[q] = r
        return r

    def _push_slots(self, q: int, rows):
        """Halo slots (without the parity offset) on rank q for `rows` = [(peer, task, idx)], sorted by (task, idx),
        coming from ONE source rank: slot = remote index of the peer + version * n_remote, the version counting rows
        of the same node within this level in list order.  Receiver and sender evaluate this on the same list, so
        they agree."""
        ridx = self._remote_idx(q)
        n_remote = self._n_remote[q]
        out, seen = [], {}
        for peer, _, _ in rows:
            v = seen.get(peer, 0)
            seen[peer] = v + 1
            if v >= self.HALO_VERSIONS:
                raise RuntimeError("push transport: more than %d live rows of node %d needed by rank %d at one level; "
                                   "use transport='a2a'" % (self.HALO_VERSIONS, peer, q))
            out.append(int(ridx[peer]) + v * n_remote)
        return out

    def _begin_level(self, level: int, names: List[str], pending_release: List[int]):
        me, own, tasks = self.rank, self._task_owner, self.tasks
        recv = [set() for _ in range(self.world)]
        send = [set() for _ in range(self.world)]
        for name in names:                                    # this rank's tasks of the level
            t = tasks[name]
            if t["func"] != "aggregate":
                continue
            for m in t["data"]["models"]:
                r = own[m[0]]
                if r != me:
                    recv[r].add(m if type(m) is tuple else tuple(m))
        vs = self._value_slot
        for q, m in self._remote.get(level, ()):              # one reference per consuming remote task
            pending_release.append(vs[m])
            send[q].add(m)
        # both ends order a (source rank -> destination rank) row list the same way: by (task name, index) — plain
        # tuple comparison — and attach the producer's peer
        peer_of = lambda m: tasks[m[0]]["data"]["peer"]
        recv_l = [[(peer_of(m), m[0], m[1]) for m in sorted(x)] for x in recv]
        send_l = [[(peer_of(m), m[0], m[1]) for m in sorted(x)] for x in send]
        self._halo_of = {}
        n_recv = sum(len(x) for x in recv_l)
        n_send = sum(len(x) for x in send_l)
        if not (n_recv or n_send or level in self._xchg_levels):
            return
        if self.transport == "push":
            ex = self.ex
            parity = self._xchg_index[level] & 1
            for r in range(self.world):
                if recv_l[r]:
                    for (peer, tn, idx), slot in zip(recv_l[r], self._push_slots(me, recv_l[r])):
                        self._halo_of[(tn, idx)] = parity * ex.halo_cap + slot
            # Transfer order = what the GPUs write at the same time (the push kernel sweeps the list front to back):
            # destinations are taken round-robin starting from rank + 1, so at any moment every GPU is written by all
            # its peers evenly.  Listing them by ascending rank instead makes every GPU write to rank 0 first, then
            # rank 1, ...: measured on 8 B200s 345 GB/s per GPU against 700 GB/s in this order
            # (scripts/push_bw.py, profiles/r02_halo_bandwidth_8gpu_*.json).
            row_bytes = ex.layout.ld * 4
            per_dst = []
            for k in range(1, self.world):
                q = (me + k) % self.world
                if send_l[q]:
                    base = ex.peer_x_ptrs[q] + (ex.halo_base + parity * ex.halo_cap) * row_bytes
                    per_dst.append([(vs[(tn, idx)], base + slot * row_bytes)
                                    for (peer, tn, idx), slot in zip(send_l[q], self._push_slots(q, send_l[q]))])
            send_slots, dst = [], []
            for i in range(max((len(x) for x in per_dst), default=0)):
                for x in per_dst:
                    if i < len(x):
                        send_slots.append(x[i][0])
                        dst.append(x[i][1])
            self._level_xchg[level] = {"send_slots": send_slots, "dst_ptrs": dst, "n_recv": n_recv}
            self.max_halo = max(self.max_halo, n_recv)
            self.max_send = max(self.max_send, n_send)
            return
        h = 0
        for r in range(self.world):                           # all_to_all output is ordered by source rank
            for peer, tn, idx in recv_l[r]:
                self._halo_of[(tn, idx)] = h
                h += 1
        self._level_xchg[level] = {"send_slots": [vs[(tn, idx)] for x in send_l for _, tn, idx in x],
                                   "send_counts": [len(x) for x in send_l],
                                   "recv_counts": [len(x) for x in recv_l]}
        self.max_halo = max(self.max_halo, h)
        self.max_send = max(self.max_send, n_send)

    def _model_slot(self, ref):
        s = self._value_slot.get(ref)
        if s is not None:
            return s
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
        if self.transport == "push":
            self.ex._grow(extra_slots)                        # inside the fixed symmetric allocation
        else:
            self.ex._grow(extra_slots, halo_rows=self.max_halo, send_rows=self.max_send)

    def _fixup_wave(self, w: Wave):
        """Halo references become rows after the pool (valid once the pool has its planned size)."""
        ex = self.ex
        base = ex.halo_base
        if w.func == "aggregate":
            w.model_slots = [[s if s >= 0 else base + (-1 - s) for s in ms] for ms in w.model_slots]
        x = getattr(w, "xchg", None)
        if x and "send_idx" not in x:
            if self.transport == "push":
                if x["send_slots"]:
                    x["send_idx"], x["dst_dev"] = ex._upload(np.asarray(x["send_slots"], dtype=np.int32),
                                                             np.asarray(x["dst_ptrs"], dtype=np.int64))
                else:
                    x["send_idx"] = None
                self.halo_bytes += x["n_recv"] * ex.layout.P * 4
            else:
                x["send_idx"] = ex._dev_i32(x["send_slots"]) if x["send_slots"] else None
                self.halo_bytes += sum(x["recv_counts"]) * ex.layout.P * 4

    # -- execution --------------------------------------------------------------------------
    def exchange(self, x: dict):
        ex = self.ex
        if self.transport == "push":
            if x["send_idx"] is not None:
                ex.push_rows(x["send_idx"], x["dst_dev"])
            ex.symm_barrier()                                 # every rank, every exchange level
            return
        import torch.distributed as dist
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

    # Rows of at least PIPELINE_ROW_BYTES (FEMNIST: 6.8 MB) are exchanged and aggregated in PIPELINE_CHUNKS column
    # ranges: the aggregate is elementwise over columns, so range i can be averaged (main stream) while range i + 1 is
    # still being pushed (side stream) — the push of 4 GB per GPU and the SpMM over it overlap instead of adding up.
    # The chunk count depends on the model alone, so every rank runs the same number of pushes and barriers.
    PIPELINE_ROW_BYTES = 2 << 20
    PIPELINE_CHUNKS = 4

    def _column_chunks(self):
        P = self.ex.layout.P
        if self.transport != "push" or P * 4 < self.PIPELINE_ROW_BYTES or self.PIPELINE_CHUNKS < 2:
            return None
        C, unit = self.PIPELINE_CHUNKS, 4096
        cuts = [min(P, (i * P // C + unit - 1) // unit * unit) for i in range(C)] + [P]
        return [(a, b - a) for a, b in zip(cuts[:-1], cuts[1:]) if b > a]

    def _run_pipelined(self, w: Wave, x: dict, chunks):
        import torch
        ex = self.ex
        main = torch.cuda.current_stream(ex.device)
        side = getattr(ex, "_xchg_stream", None)
        if side is None:
            side = ex._xchg_stream = torch.cuda.Stream(device=ex.device)
        side.wait_stream(main)                                # the rows to push are final on the main stream
        do_agg = w.func == "aggregate" and len(w.out_slots) > 0
        for cols in chunks:
            with torch.cuda.stream(side):
                if x["send_idx"] is not None:
                    ex.push_rows(x["send_idx"], x["dst_dev"], cols=cols)
                ex.symm_barrier()                             # every rank, every chunk of every exchange level
                ev = side.record_event()
            main.wait_event(ev)
            if do_agg:
                ex.aggregate_wave(w.model_slots, w.out_slots, w.weights, variant=self.agg_variant,
                                  prepared=w.prepared, cols=cols)

    def run_wave(self, w: Wave, loss_out=None):
        x = getattr(w, "xchg", None)
        if x:
            chunks = self._column_chunks()
            if chunks and w.func in ("aggregate", "exchange"):
                self._run_pipelined(w, x, chunks)
                return
            self.exchange(x)
        if w.func != "exchange":
            super().run_wave(w, loss_out=loss_out)
