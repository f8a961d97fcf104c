"""TEST INFRASTRUCTURE — the CPU oracle.  NOT product code.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs
may import this package; the product (``plssvm_sparse_fp22_b200``) never does.

Two checkers live here:

* ``oracle.port``  — ctypes binding of ``oracle/lssvm_oracle.c``, the plain-C restatement of the reference's
  hot path (each C function cites the reference file:line it follows).
* ``oracle.ref``   — ctypes binding of ``oracle/_ref/liboracle_ref_*.so``: the UNMODIFIED reference OpenMP
  backend compiled from ``/root/reference`` by ``oracle/Makefile`` (sources are never copied into this repo).

Parity pin: the port is checked against ``ref`` and against the reference's golden predict fixture
(``tests/test_oracle.py``, ``tests/golden/``).
"""
from . import port, ref  # noqa: F401
