"""Initial model rows.  Every node starts from the identical model created under
`torch.manual_seed(settings.seed)` (reference: dasklearn/functions.py:46-48, create_model
dasklearn/models/__init__.py:19-42).  The init is produced on the CPU with torch's default
layer initialisers, constructed in the reference's module order so the RNG stream matches,
then flattened into the row layout and broadcast on the device."""
import torch
import torch.nn as nn

from .layout import ModelLayout


def initial_row(layout: ModelLayout, seed: int) -> torch.Tensor:
    gen_state = torch.random.get_rng_state()
    try:
        torch.manual_seed(seed)
        if layout.name == "gnlenet":
            # GNLeNet.__init__ (dasklearn/models/cifar10.py:111-136): features then classifier
            convs = [nn.Conv2d(3, 32, 5, 1, 2), nn.GroupNorm(2, 32), nn.Conv2d(32, 32, 5, 1, 2),
                     nn.GroupNorm(2, 32), nn.Conv2d(32, 64, 5, 1, 2), nn.GroupNorm(2, 64), nn.Linear(576, 10)]
        elif layout.name == "femnist":
            # CNN.__init__ (dasklearn/models/femnist.py:59-63)
            convs = [nn.Conv2d(1, 32, 5, padding=2), nn.Conv2d(32, 64, 5, padding=2),
                     nn.Linear(7 * 7 * 64, 512), nn.Linear(512, 62)]
        else:
            raise RuntimeError("unknown layout %s" % layout.name)
    finally:
        torch.random.set_rng_state(gen_state)
    row = torch.zeros(layout.ld, dtype=torch.float32)
    params = [p for m in convs for p in (m.weight, m.bias)]
    assert len(params) == len(layout.entries)
    for p, (key, shp, off) in zip(params, layout.entries):
        assert tuple(p.shape) == tuple(shp), (key, p.shape, shp)
        row[off:off + p.numel()] = p.detach().reshape(-1)
    return row
