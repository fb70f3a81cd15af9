"""CUDA kernel wrappers (sm_100a).  See include/dryad_b200.h for the C ABI."""
