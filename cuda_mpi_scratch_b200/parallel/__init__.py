"""Worlds, symmetric memory, Cartesian topology, MPI-like comm, datatypes, halo plans."""
