"""lime-b200: LiME's classification hot path (ClusterLCP -> ClusterBWT_DA -> Classify) on B200.

The product is the C-ABI shared library `liblime_b200.so` (include/lime_cuda.h) and the C++ host
executables in `lime_b200/host/`; the Python modules here are plumbing around the ABI:

  _lib      ctypes prototypes of every exported symbol
  api       LimeContext: one context per process and GPU
  pipeline  the strand loop of LiME_paired.sh / single-end runs on one GPU
  multi     position-sharded multi-GPU driver (torch.distributed)
  datagen   inputs in the reference file formats (exact tiny generator, large synthesiser)
"""
__version__ = "0.1"
