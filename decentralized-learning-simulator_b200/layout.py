"""Flat parameter layouts of the simulator's models as rows of the stacked `[N, ld]` tensor.

A node model is one fp32 row in `parameters()` order (reference: `GNLeNet.__init__`
`dasklearn/models/cifar10.py:111-136`, FEMNIST `CNN.__init__` `dasklearn/models/femnist.py:59-63`).
Neither model has buffers, so `parameters()` order == `state_dict()` order.  Rows are padded to
`ld` (a multiple of 64 floats = 256 B) so every row starts 128-bit aligned for vector loads.
"""
from dataclasses import dataclass
from typing import Dict, List, Tuple

ROW_ALIGN_FLOATS = 64


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


@dataclass(frozen=True)
class ModelLayout:
    name: str            # kernel family id ("gnlenet" | "femnist")
    arch_id: int         # id understood by the C-ABI
    entries: Tuple[Tuple[str, Tuple[int, ...], int], ...]  # (state_dict key, shape, offset)
    P: int
    ld: int
    img_hw: int
    img_c: int
    num_classes: int
    log_softmax_output: bool   # True: model ends in log_softmax and the loss is NLL

    def offset(self, key: str) -> int:
        for k, _, off in self.entries:
            if k == key:
                return off
        raise KeyError(key)

    @property
    def img_bytes(self) -> int:
        return self.img_hw * self.img_hw * self.img_c


def _mk(name, arch_id, shapes: List[Tuple[str, Tuple[int, ...]]], img_hw, img_c, ncls, lsm) -> ModelLayout:
    entries = []
    off = 0
    for key, shp in shapes:
        n = 1
        for s in shp:
            n *= s
        entries.append((key, tuple(shp), off))
        off += n
    return ModelLayout(name, arch_id, tuple(entries), off, _round_up(off, ROW_ALIGN_FLOATS),
                       img_hw, img_c, ncls, lsm)


# create_model("cifar10", "gnlenet") -> GNLeNet(input_channel=3, output=10, model_input=(32, 32))
# (`dasklearn/models/__init__.py:24-26`); classifier input = 64*3*3 = 576.
GNLENET = _mk("gnlenet", 0, [
    ("features.0.weight", (32, 3, 5, 5)), ("features.0.bias", (32,)),
    ("features.2.weight", (32,)), ("features.2.bias", (32,)),
    ("features.4.weight", (32, 32, 5, 5)), ("features.4.bias", (32,)),
    ("features.5.weight", (32,)), ("features.5.bias", (32,)),
    ("features.8.weight", (64, 32, 5, 5)), ("features.8.bias", (64,)),
    ("features.9.weight", (64,)), ("features.9.bias", (64,)),
    ("classifier.0.weight", (10, 576)), ("classifier.0.bias", (10,)),
], 32, 3, 10, True)

# FEMNIST CNN (`dasklearn/models/femnist.py:45-85`); NUM_CLASSES=62 from the docstrings (:19, :54)
FEMNIST = _mk("femnist", 1, [
    ("conv1.weight", (32, 1, 5, 5)), ("conv1.bias", (32,)),
    ("conv2.weight", (64, 32, 5, 5)), ("conv2.bias", (64,)),
    ("fc1.weight", (512, 3136)), ("fc1.bias", (512,)),
    ("fc2.weight", (62, 512)), ("fc2.bias", (62,)),
], 28, 1, 62, False)

assert GNLENET.P == 85354 and FEMNIST.P == 1690046

LAYOUTS: Dict[str, ModelLayout] = {"gnlenet": GNLENET, "femnist": FEMNIST}


def layout_for(dataset: str, model: str = None) -> ModelLayout:
    """Same selection rule as `create_model` (`dasklearn/models/__init__.py:19-42`) for the models
    this executor implements; anything else is an error (no fallback path)."""
    if dataset == "cifar10" and model == "gnlenet":
        return GNLENET
    if dataset == "femnist":
        return FEMNIST
    raise RuntimeError("Unknown or unsupported model for the B200 executor: dataset=%s model=%s" % (dataset, model))


def flatten_state_dict(layout: ModelLayout, state_dict) -> "torch.Tensor":
    import torch
    row = torch.zeros(layout.ld, dtype=torch.float32)
    for key, shp, off in layout.entries:
        t = state_dict[key].detach().to(torch.float32).reshape(-1)
        row[off:off + t.numel()] = t
    return row


def unflatten_row(layout: ModelLayout, row) -> Dict[str, "torch.Tensor"]:
    out = {}
    for key, shp, off in layout.entries:
        n = 1
        for s in shp:
            n *= s
        out[key] = row[off:off + n].reshape(shp).clone()
    return out
