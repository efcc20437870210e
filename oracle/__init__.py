"""CPU oracle for the Neural-Random-Walker `--run_mode pc` path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and there only as the checker (or, for
the CPU-baseline legs, as the thing timed *beside* the GPU path).  The product
path (``neural-random-walker_b200/``) never imports this package and fails
loudly when its CUDA library is missing.

Parity pinning: the reference ships no tests and no golden vectors for this
path, so the restatement in ``restated.py`` is pinned against
outputs of the *reference itself*, generated in the build container by
``oracle/make_golden.py`` (which imports ``/root/reference`` unmodified through
``ref_import.py``) and committed under ``tests/golden/``.
"""
