"""CHOLMOD/B200: a B200-native (sm_100a) supernodal numeric Cholesky for CHOLMOD.

`b200`      ctypes binding of the C ABI (include/cholmod_b200.h): Factor, panel_factor_noalloc, ...
`matrices`  the synthetic SPD matrices of BASELINE.json's configs.

The product is the shared library suitesparse_b200/lib/libcholmod_b200.so (sources in csrc/); there is no
CPU fallback -- without it, or without an sm_100 device, everything that computes raises b200.B200Error.
"""
