"""Data plane: flat arenas, FedAvg, broadcast, replication, codec, peer memory."""
from .arena import Arena, Manifest  # noqa: F401
