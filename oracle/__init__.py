"""CPU oracle for the fallback placement path: test infrastructure only.

Importable from tests/, __graft_entry__.smoke() and bench.py's CPU legs - never from
the product package.  See py_oracle.py (literal CPython restatement, pinned to the
reference's golden vectors) and c_oracle.c / c_oracle.py (fast C restatement over the
int64/float64 wire tables, pinned to py_oracle and the same vectors).
"""
