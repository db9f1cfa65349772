"""Drop-in broker for the batched B200 executor.

The reference broker (`dasklearn/broker.py`) receives the coordinator's `config` message (settings + the
serialized compute graph), starts worker processes and feeds them one task at a time
(`Broker.on_message` broker.py:292-331, `schedule_task` :265-275, `handle_task_result` :277-290); per task it
records `(name, func, worker, transfer_to, execute, transfer_from, total)` and writes `tasks_<broker>.csv`
(`write_statistics` broker.py:71-77).

`BatchedBrokerMixin` keeps that message protocol and those output files but solves the graph with
`dls_b200.solver.BatchedSolver`: no worker processes, one GPU wave per wavefront and task function.
Use it in the reference tree as

    from dasklearn.broker import Broker
    from integration.broker_b200 import BatchedBrokerMixin
    class B200Broker(BatchedBrokerMixin, Broker):
        pass

(the mixin only overrides `on_message`, `write_statistics` and `shutdown`; `start()`, the ZeroMQ
`Communication` and the monitor stay the reference's).  Standalone — without the reference package — the
mixin works with any object that has `identity`, `logger` and a `communication` offering
`send_message_to_coordinator(bytes)` (see tests/test_broker_integration.py).

Differences the protocol hides: a sink task's `result` is a small dict (`{"peer", "slot"}`) instead of a pickled
nn.Module — the coordinator only counts results (simulation.py:120-138); settings classes the reference broker
cannot rebuild (GL / AD-PSGD / ..., broker.py:305-312) fall back to `dls_b200.settings.SessionSettings.from_dict`.
"""
import logging
import os
import pickle
import time
from typing import Callable, Dict, List, Optional, Tuple

TASKS_HEADER = "broker,task_name,function,worker,transfer_to_time,execute_time,transfer_from_time,total_time\n"


def settings_from_message(msg: Dict):
    """Settings object of the `config` message.  The reference maps `settings_class` through a dict of five
    classes (broker.py:305-312); anything else — and the case where the reference package is absent — becomes
    `dls_b200.settings.SessionSettings`, which carries every field the executor reads."""
    name = msg.get("settings_class", "SessionSettings")
    try:                                                        # reference classes, when importable
        import importlib
        mod = {"SessionSettings": "dasklearn.session_settings",
               "DPSGDSettings": "dasklearn.simulation.dpsgd.settings",
               "ConfluxSettings": "dasklearn.simulation.conflux.settings",
               "ShatterSettings": "dasklearn.simulation.shatter.settings",
               "TeleportationSettings": "dasklearn.simulation.teleportation.settings",
               "GLSettings": "dasklearn.simulation.gl.settings",
               "ADPSGDSettings": "dasklearn.simulation.adpsgd.settings",
               "SuperGossipSettings": "dasklearn.simulation.super_gossip.settings",
               "LuborSettings": "dasklearn.simulation.lubor.settings",
               "FLSettings": "dasklearn.simulation.fl.settings"}.get(name)
        if mod:
            cls = getattr(importlib.import_module(mod), name)
            return cls.from_dict(msg["settings"])
    except Exception:                                           # noqa: BLE001 — no reference package / unknown class
        pass
    from dls_b200.settings import SessionSettings
    return SessionSettings.from_dict(msg["settings"])


class BatchedBrokerMixin:
    """Message handling of a broker that solves its compute graph in GPU waves."""

    #: `executor_factory(settings, owned_peers) -> WaveExecutor`; default: synthetic partitions on cuda:0.  A real
    #: run passes a factory that opens the CIFAR-10 partition cache (dls_b200.cifar_cache).
    executor_factory: Optional[Callable] = None

    def _b200_init(self):
        if getattr(self, "_b200_ready", False):
            return
        self._b200_ready = True
        self.task_statistics: List[Tuple] = []
        self.serialized_dag: Optional[List[Dict]] = None
        self.executor = None
        self.sinks_sent = 0
        if not hasattr(self, "logger"):
            self.logger = logging.getLogger(self.__class__.__name__)
        if not hasattr(self, "identity"):
            self.identity = "broker_b200"

    # -- protocol ---------------------------------------------------------------------------
    def on_message(self, identity: str, msg: Dict):
        self._b200_init()
        if identity == "coordinator" and msg["type"] == "config":               # broker.py:293-317
            self.start_time = time.time()
            self.broker_addresses = msg.get("brokers", {})
            self.brokers_to_clients = msg.get("brokers_to_clients", {})
            self.clients_to_brokers = {c: b for b, cs in self.brokers_to_clients.items() for c in cs}
            self.settings = settings_from_message(msg)
            os.makedirs(self.settings.data_dir, exist_ok=True)
            self.serialized_dag = msg["dag"]                                      # WorkflowDAG.serialize(), dag.py:28-32
            owned = self.brokers_to_clients.get(self.identity)
            self.executor = self._make_executor(self.settings, owned)
        elif identity == "coordinator" and msg["type"] == "tasks":              # broker.py:318-322
            self.solve(msg["tasks"])
        elif msg["type"] == "shutdown":                                          # broker.py:323-325
            self.shutdown()
        elif msg["type"] == "result":
            raise RuntimeError("the batched broker solves its whole partition itself; cross-broker results travel "
                               "GPU to GPU (dls_b200.partition), not as `result` messages")
        else:
            raise RuntimeError("Unknown message with type %s" % msg["type"])     # broker.py:330-331

    def _make_executor(self, settings, owned):
        if self.executor_factory is not None:
            return self.executor_factory(settings, owned)
        import torch
        from dls_b200.dataset import DeviceDataset
        from dls_b200.executor import WaveExecutor
        from dls_b200.layout import layout_for
        L = layout_for(settings.dataset, settings.model)
        ds = DeviceDataset.synthetic(L, settings.participants, 320, 2048, torch.device("cuda:0"), seed=settings.seed)
        return WaveExecutor(settings, ds)

    def solve(self, start_tasks):
        """The `tasks` message names the source tasks (`get_source_tasks`, dag.py:15-19); the batched solver
        derives them itself from the graph, so the list is only checked."""
        from dls_b200.solver import BatchedSolver
        sources = {t["name"] for t in self.serialized_dag if not t["inputs"]}
        unknown = set(start_tasks) - sources
        if unknown:
            raise RuntimeError("tasks message names non-source tasks: %s" % sorted(unknown)[:3])
        t0 = time.time()
        try:
            solver = BatchedSolver(self.executor, self.serialized_dag, on_sink=self._on_sink)
            solver.run(timing=True)
        except Exception as exc:                                  # worker.py:34-38 + broker.py:171-181: error -> shutdown everyone
            self.logger.exception(exc)
            self.shutdown_everyone()
            raise
        # tasks_<broker>.csv rows (broker.py:195-201): a wave's time is shared equally by its tasks; there is no
        # transfer to or from a worker process
        elapsed = 0.0
        w_iter = iter(solver.waves)
        for func, n_tasks, ms in solver.wave_times:
            w = next(w_iter)
            while w.func != func or len(w.names) != n_tasks:
                w = next(w_iter)
            per_task = ms / 1e3 / max(1, n_tasks)
            elapsed += ms / 1e3
            for name in w.names:
                self.task_statistics.append((name, func, 0, 0.0, per_task, 0.0, elapsed))
        self.solve_seconds = time.time() - t0

    def _on_sink(self, name: str, res):
        # broker.py:203-206: sink tasks report to the coordinator
        h = res[0]
        msg = pickle.dumps({"type": "result", "task": name, "result": {"slot": int(h.slot)}})
        self.communication.send_message_to_coordinator(msg)
        self.sinks_sent += 1

    # -- outputs ----------------------------------------------------------------------------
    def write_statistics(self):
        """`tasks_<broker>.csv`, the reference's columns (broker.py:71-77)."""
        self._b200_init()
        broker_id = self.identity.split("_")[1] if "_" in self.identity else self.identity
        with open(os.path.join(self.settings.data_dir, "tasks_%s.csv" % self.identity), "w") as f:
            f.write(TASKS_HEADER)
            for st in self.task_statistics:
                f.write("%s,%s\n" % (broker_id, "%s,%s,%d,%f,%f,%f,%f" % st))

    def shutdown(self):
        self._b200_init()
        if getattr(self, "settings", None) is not None:
            self.write_statistics()
        sup = getattr(super(), "shutdown", None)
        if sup is not None and getattr(self, "workers", None) is not None:
            try:
                sup()
            except Exception:                                     # noqa: BLE001 — no event loop in standalone use
                pass

    def shutdown_everyone(self):
        msg = pickle.dumps({"type": "shutdown"})
        comm = getattr(self, "communication", None)
        if comm is not None:
            if hasattr(comm, "send_message_to_all_brokers"):
                comm.send_message_to_all_brokers(msg)
            comm.send_message_to_coordinator(msg)
        self.shutdown()
