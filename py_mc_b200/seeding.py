"""Seed of the counter-based random stream.

The reference seeds nothing itself and draws from Python's global ``random`` module
(``py_mc/mc.py:2``, ``examples/ising2D.py:2``).  The kernels use Philox keyed by a 64-bit seed:
``set_seed(n)`` fixes it; otherwise it is drawn from ``random.getrandbits(64)`` the first time a
lattice goes to the device, so ``random.seed(k)`` in a user script still makes a run reproducible.
"""
import random as _random

_seed = None


def set_seed(seed):
    global _seed
    _seed = None if seed is None else int(seed) & 0xFFFFFFFFFFFFFFFF


def get_seed():
    """The seed for the next device lattice (drawn once from ``random`` if not set)."""
    global _seed
    if _seed is None:
        _seed = _random.getrandbits(64)
    return _seed
