"""Hand-written sm_100a kernels (ctypes-bound) and their CPU/PyTorch reference restatements."""
