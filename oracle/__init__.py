"""TEST INFRASTRUCTURE -- the CPU oracle for the sdrx per-channel DSP chain.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product (``sdrx_b200``) never does.

Two back ends with the same Python surface (see ``oracle.bindings``):

* ``"oracle"``  -- ``oracle/liboracle.so``: the plain-C restatement in ``sdrx_oracle.c``.
* ``"ref"``     -- ``oracle/_ref/libsdrx_ref.so``: the reference's own headers compiled from
  ``/root/reference/src`` with the pinned portable flags (built by ``oracle/Makefile`` in the
  container that has the reference; the .so travels to the GPU box).
* ``"ref_avx2"`` -- same harness, the reference's Release flags (timed CPU baseline).
"""
from .bindings import *  # noqa: F401,F403
from . import bindings  # noqa: F401
