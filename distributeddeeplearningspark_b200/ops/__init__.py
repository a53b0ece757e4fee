"""Python wrappers (ctypes + autograd) around the hand-written sm_100a kernels in ``csrc/``."""
