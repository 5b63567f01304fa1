"""``from models.model import Network`` -- the reference's import path (models/model.py:9-27).

The class lives in the package (distributed_neural_network_b200/models/network.py); this module
keeps the reference's module name working for scripts and checkpoints written against it.
"""
from distributed_neural_network_b200.models.network import Network  # noqa: F401

__all__ = ["Network"]
