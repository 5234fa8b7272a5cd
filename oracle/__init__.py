"""CPU checker for the exact-attention hot path (TEST INFRASTRUCTURE, not product).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package; ``cuda_flashattention_b200`` never does.
"""
from .oracle_lib import *  # noqa: F401,F403
