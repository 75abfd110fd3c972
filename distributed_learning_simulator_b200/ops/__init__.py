"""Python front-ends of the hand-written sm_100a kernels (csrc/*.cu -> lib/libdls_b200.so,
loaded through the C ABI declared in include/dls_b200.h).  Every op has a plain PyTorch
oracle next to it; the CUDA path is mandatory on a GPU box (a missing library raises)."""
