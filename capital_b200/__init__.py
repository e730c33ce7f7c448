"""capital_b200: B200-native CholeskyQR2 / cholinv / SUMMA hot path behind the reference's engine interface.

The product is the C-ABI library (include/capital_b200.h, csrc/); this package is the thin Python host
used by tests and bench.py.  Nothing here computes on the CPU."""
from ._lib import CapitalError, load  # noqa: F401
