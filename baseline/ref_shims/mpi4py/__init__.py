"""Single-node stand-in for the ``mpi4py`` package (which, like ``mpirun``, does not exist in this image).

Only what the reference's scripts call is provided (``from mpi4py import MPI``): see ``MPI.py``. The reference
files themselves are run byte-identical from ``baseline/_ref``; this package merely answers their import.
"""
from . import MPI  # noqa: F401

__all__ = ["MPI"]
