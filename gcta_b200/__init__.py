"""gcta_b200 - B200-native (sm_100a) drop-in for GCTA's `--bfile ... --make-grm` hot path.

The product is the C-ABI shared library ``gcta_b200/libgcta_b200.so`` (declared in
``include/gcta_b200.h``) and the C++ host front-end ``gcta_b200/gcta64_b200``.  The Python in
this package is harness only: a ctypes binding used by tests/ and bench.py, the synthetic
.bed generator and GRM file readers.
"""
__version__ = "0.1.0"
