"""Tensor-parallel plumbing: shard plan, Communicator, symmetric NVLink workspace."""
from .comm import Communicator, comm_for  # noqa: F401
