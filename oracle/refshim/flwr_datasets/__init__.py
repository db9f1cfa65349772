"""Stand-in for flwr_datasets (needs network + is not installed). Test infrastructure only."""


class FederatedDataset:
    def __init__(self, *a, **kw):
        raise RuntimeError("flwr_datasets is a shim: inject a synthetic dataset instead")
