"""ctypes binding of libmst_b200.so (the C ABI in include/mst_b200.h) — see lib.py."""
