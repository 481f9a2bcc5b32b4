"""Physical constants (reference: multi_constants.py:4-5). The CUDA kernels compile the same values in."""
g = 9.81
rho = 1023
