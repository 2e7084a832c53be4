"""h2ok_hist — host-side mirror of the C ABI in include/h2ok_hist.h (B200 GBM histogram path).

Importing this package never touches oracle/: the product path is the CUDA library only and raises if
libh2ok_hist.so is missing (h2ok_hist.capi.load()).
"""
from . import abi  # noqa: F401
