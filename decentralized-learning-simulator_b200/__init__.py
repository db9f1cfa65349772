"""B200-native batched executor for the decentralized-learning simulator's compute-graph solve.

Importable as `dls_b200` (see dls_b200.py at the repo root).
"""
__version__ = "0.1.0"
