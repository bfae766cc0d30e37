"""TEST INFRASTRUCTURE ONLY — never imported by the product path.

Import shims that let the UNMODIFIED reference attention stack under
/root/reference/python be imported in the build container (no GPU, torch 2.11,
transformers 5.5) so that it can be run on CPU to (a) validate the oracle
restatement in oracle/attention_oracle.py and (b) generate the golden vectors
committed under tests/golden/ (see tests/golden/make_golden.py).

/root/reference does not exist on the GPU box, so nothing here may be used by
`-m gpu` tests, smoke() or bench.py.

Shims:
  1. stub `IPython` / `IPython.display` (python/sglang/utils.py:25)
  2. define `torch._C._cpu._is_amx_tile_supported` (python/sglang/srt/utils.py:216-217)
  3. pre-register a stub `sglang.srt.model_executor.cuda_graph_runner` exposing
     `get_is_capture_mode()` (lazy import at mem_cache/memory_pool.py:379)
"""
import os
import sys
import types

REFERENCE_ROOT = "/root/reference"
REFERENCE_PY = os.path.join(REFERENCE_ROOT, "python")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_PY, "sglang", "srt"))


def install():
    """Make `import sglang.srt...` resolve to the read-only reference tree."""
    if not reference_available():
        raise RuntimeError("reference tree not present (expected on the GPU box)")
    sys.dont_write_bytecode = True
    os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
    if REFERENCE_PY not in sys.path:
        sys.path.insert(0, REFERENCE_PY)

    import torch

    if "IPython" not in sys.modules:
        ip = types.ModuleType("IPython")
        ipd = types.ModuleType("IPython.display")
        ipd.HTML = lambda *a, **k: None
        ipd.display = lambda *a, **k: None
        ip.display = ipd
        sys.modules["IPython"] = ip
        sys.modules["IPython.display"] = ipd

    if not hasattr(torch._C._cpu, "_is_amx_tile_supported"):
        torch._C._cpu._is_amx_tile_supported = lambda: False

    name = "sglang.srt.model_executor.cuda_graph_runner"
    if name not in sys.modules:
        stub = types.ModuleType(name)
        stub.get_is_capture_mode = lambda: False
        sys.modules[name] = stub


def load_reference_attention():
    """Return the reference classes that make up the attention oracle."""
    install()
    from sglang.srt.layers.attention.torch_native_backend import TorchNativeAttnBackend
    from sglang.srt.layers.radix_attention import RadixAttention
    from sglang.srt.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
    from sglang.srt.model_executor.forward_batch_info import ForwardBatch, ForwardMode

    return dict(
        TorchNativeAttnBackend=TorchNativeAttnBackend,
        RadixAttention=RadixAttention,
        MHATokenToKVPool=MHATokenToKVPool,
        ReqToTokenPool=ReqToTokenPool,
        ForwardBatch=ForwardBatch,
        ForwardMode=ForwardMode,
    )
