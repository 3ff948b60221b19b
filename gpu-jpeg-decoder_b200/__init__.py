"""gjd — B200-native baseline-JPEG decode path.  The product is the C ABI (include/gjd.h,
build/libgjd.so); `gjd.py` beside this file is a ctypes shim used by the tests and bench.py."""
