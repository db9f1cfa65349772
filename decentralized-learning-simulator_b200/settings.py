"""Settings objects with the reference's field names (dasklearn/session_settings.py:9-63,
dasklearn/simulation/dpsgd/settings.py:7-12).  The executor only READS settings; any object
with these attributes works (the reference's own dataclasses included).  `from_dict` accepts
the `settings` dict of the coordinator's `config` message (dasklearn/simulation/simulation.py:448-457)."""
from dataclasses import dataclass, field, fields, asdict
from typing import Optional


@dataclass
class LearningSettings:
    learning_rate: float = 0.002
    momentum: float = 0.9
    weight_decay: float = 0.0
    batch_size: int = 32
    local_steps: int = 5


@dataclass
class SessionSettings:
    algorithm: str = "dpsgd"
    seed: int = 42
    work_dir: str = ""
    dataset: str = "cifar10"
    learning: LearningSettings = field(default_factory=LearningSettings)
    participants: int = 16
    synchronous: bool = False
    model: Optional[str] = "gnlenet"
    alpha: float = 1
    partitioner: str = "iid"
    torch_device_name: str = "cuda:0"
    test_interval: int = 0
    brokers: Optional[int] = 1
    rounds: int = 10
    data_dir: str = ""
    log_level: str = "INFO"
    # D-PSGD extras
    topology: str = "kreg"
    k: int = 0
    el: str = "oracle"
    # executor extras (not in the reference): math mode of the train/test kernels
    precision: int = 0

    @classmethod
    def from_dict(cls, d: dict) -> "SessionSettings":
        known = {f.name for f in fields(cls)}
        kw = {k: v for k, v in d.items() if k in known and k != "learning"}
        lr = d.get("learning")
        if isinstance(lr, dict):
            lk = {f.name for f in fields(LearningSettings)}
            kw["learning"] = LearningSettings(**{k: v for k, v in lr.items() if k in lk})
        elif lr is not None:
            kw["learning"] = lr
        return cls(**kw)

    def to_dict(self) -> dict:
        return asdict(self)
