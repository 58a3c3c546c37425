"""tokp: B200-native key-point FD derivatives + iLQR (C ABI in include/tokp.h, kernels in csrc/)."""
