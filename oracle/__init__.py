"""CPU oracle for the SuperPoint-open + LightGlue forward path.

TEST INFRASTRUCTURE ONLY.  This package is a plain PyTorch-CPU fp32 restatement of the
reference algorithm, function by function with reference file:line citations.  It may be
imported only by `tests/`, `__graft_entry__.smoke()` and the baseline legs of `bench.py`
(`cpu_baseline`, `--impl reference`, and `gpu_library_baseline`, which feeds the same functions CUDA
tensors to time the reference algorithm through PyTorch eager), and only as the checker or the timed
baseline -- never by the product package `glue_factory_b200`, which has no CPU fallback.

Parity pin: the restatement is checked against outputs of the UNMODIFIED reference, imported
from /root/reference in the authoring container by `tests/golden/make_golden.py`; the vectors
are committed under `tests/golden/` and replayed by `tests/test_oracle_golden.py`.
(The reference itself ships no golden vectors or known-answer tests for this path)
"""
