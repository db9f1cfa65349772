"""The batched broker (integration/broker_b200.py) driven with the coordinator's message protocol
(reference: dasklearn/broker.py:292-331, dasklearn/simulation/simulation.py:448-457) on CPU: rows and arithmetic
are the oracle's (tests/cpu_standin.py), the protocol, the sink results and `tasks_<broker>.csv` are what is tested."""
import csv
import os
import pickle
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import dls_b200
from dls_b200 import synthetic
from dls_b200.layout import GNLENET as L
from dls_b200.settings import SessionSettings, LearningSettings
from integration.broker_b200 import BatchedBrokerMixin, TASKS_HEADER, settings_from_message


class FakeComm:
    def __init__(self):
        self.to_coordinator = []

    def send_message_to_coordinator(self, msg):
        self.to_coordinator.append(pickle.loads(msg))

    def send_message_to_all_brokers(self, msg):
        pass


def test_config_tasks_shutdown_round_trip(tmp_path):
    from cpu_standin import OracleExecutor
    from oracle import oracle as O
    n, n_i = 6, 32
    dag = synthetic.dpsgd_dag(n, 2, synthetic.ring_topology(n), test_interval=2)
    st = SessionSettings(participants=n, seed=42, data_dir=str(tmp_path), learning=LearningSettings(0.002, 0.9, 0.0, 32, 1))
    imgs, labs = synthetic.make_partitions(n, n_i, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(16, 32, 3, 10)

    class Broker(BatchedBrokerMixin):
        identity = "broker_abc123"
        executor_factory = staticmethod(lambda s, owned: OracleExecutor(s, imgs, labs, timg, tlab, list(range(n)), n_i))

    torch.set_num_threads(1)
    b = Broker()
    b.communication = FakeComm()
    # the coordinator's config message (simulation.py:448-457): settings as a dict, the graph serialized
    b.on_message("coordinator", {"type": "config", "brokers": {"broker_abc123": "tcp://127.0.0.1:1"},
                                 "brokers_to_clients": {"broker_abc123": list(range(n))},
                                 "settings_class": "DPSGDSettings", "settings": st.to_dict(), "dag": dag})
    assert b.settings.participants == n and b.settings.learning.local_steps == 1
    sources = [t["name"] for t in dag if not t["inputs"]]
    b.on_message("coordinator", {"type": "tasks", "tasks": sources})
    sinks = [t["name"] for t in dag if not t["outputs"]]
    got = [m for m in b.communication.to_coordinator if m["type"] == "result"]
    assert sorted(m["task"] for m in got) == sorted(sinks)                      # one result per sink task
    b.on_message("coordinator", {"type": "shutdown"})
    path = os.path.join(str(tmp_path), "tasks_broker_abc123.csv")
    text = open(path).read()
    assert text.startswith(TASKS_HEADER)
    rows = list(csv.reader(text.strip().split("\n")))[1:]
    assert len(rows) == len(dag) and all(len(r) == 8 for r in rows)
    assert {r[0] for r in rows} == {"abc123"} and {r[2] for r in rows} == {"train", "aggregate", "test"}
    assert all(float(r[5]) >= 0 for r in rows)
    # the models are what the sequential oracle solve produces, bit for bit
    want, _ = O.solve_dag(L, dag, imgs, labs, timg, tlab, seed=42, batch=32, local_steps=1, lr=0.002, mu=0.9, wd=0.0,
                          n_i=n_i, minibatch_permutation=synthetic.minibatch_permutation)
    for m in got:
        row = b.executor.X[m["result"]["slot"], :L.P].numpy()
        assert np.array_equal(row, want[m["task"]][:L.P])
    # accuracy files exist, one per peer, in the reference's schema
    for peer in range(n):
        head = open(os.path.join(str(tmp_path), "accuracies_%d.csv" % peer)).readline().strip()
        assert head == "algorithm,dataset,partitioner,alpha,peer,round,time,accuracy,loss"


def test_protocol_errors(tmp_path):
    class Broker(BatchedBrokerMixin):
        identity = "broker_x"
    b = Broker()
    b.communication = FakeComm()
    with pytest.raises(RuntimeError, match="Unknown message"):
        b.on_message("coordinator", {"type": "frobnicate"})
    st = settings_from_message({"settings_class": "GLSettings", "settings": {"participants": 4, "dataset": "cifar10",
                                                                           "learning": {"batch_size": 16}}})
    assert st.participants == 4 and st.learning.batch_size == 16              # classes the reference broker cannot rebuild
