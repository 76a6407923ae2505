"""CPU oracle for the hydrogym/jax Kolmogorov-flow environment step.

TEST INFRASTRUCTURE ONLY.  Nothing under ``hydrogym_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs use it, and only as the checker.

PARITY UNPINNED: the reference ships no test, golden vector or fixture for the
JAX Kolmogorov path and JAX is not installable here, so this
restatement is anchored on the reference's source lines (cited per function)
and on analytic known-answer tests, not on reference-produced outputs.
"""
