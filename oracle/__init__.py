"""CPU oracle for the hydrogym/jax Kolmogorov-flow environment step.

TEST INFRASTRUCTURE ONLY.  Nothing under ``hydrogym_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs use it, and only as the checker.

PARITY PINNED (see ``kolmogorov_oracle.py``'s header): the reference ships no test,
golden vector or fixture for the JAX Kolmogorov path and JAX is not
installable here, so the pin is the reference's own unmodified source executed over
NumPy stand-ins for its third-party imports (``tests/golden/jax_numpy_shim.py``); its
outputs are committed as ``tests/golden/reference_shim_64_re100.npz`` and both this
oracle (CPU tests) and the CUDA path (``tests/test_gpu_reference_golden.py``) are
checked against them, next to the analytic known-answer tests.
"""
