"""oracle/ -- TEST INFRASTRUCTURE ONLY.

CPU checker for the CUDA VQ path: `vq_oracle.c` (our FP64 restatement of the
reference algorithm) and `_ref/` (the reference itself, compiled from its own
sources by `oracle/Makefile`).  Only `tests/`, `__graft_entry__.smoke()` and
`bench.py`'s cpu_baseline / `--impl reference` legs may import this package.
The product package `ecoz2rs_b200` never does.
"""
