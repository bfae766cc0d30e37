"""B200-native (sm_100a) implementation of SGLang's per-layer inference hot path.

Layout:
  csrc/     hand-written CUDA kernels + the extern "C" ABI (include/b200_hotpath.h)
  _cabi.py  ctypes binding (fails loudly if the library is missing; no CPU fallback)
  kernels.py  op wrappers named after the sgl_kernel ops they replace
  srt/      host-side mirror of the reference plugin interfaces for this path
            (AttentionBackend / LinearMethodBase / memory pools / ForwardBatch)
"""
__version__ = "0.1.0"
