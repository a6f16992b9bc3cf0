"""A stand-in for the `mpi4py` package, sufficient to run the reference scripts unmodified.

mpi4py / an MPI library cannot be installed on the build or GPU boxes (no network, not in the
offline wheelhouse), so the reference's `from mpi4py import MPI` is served by this module.  It
implements exactly the lowercase (pickle) API the three scripts use — COMM_WORLD.Get_rank,
Get_size, gather, bcast, Barrier, Abort — with the same semantics, on top of
torch.distributed's gloo backend (one process per rank, started by torchrun or by
baseline/run_reference.py).  With WORLD_SIZE unset or 1 it is a single-rank communicator.
None of the framework's own code is involved.
"""
from . import MPI  # noqa: F401
