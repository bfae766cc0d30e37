"""Drop-in proof: B200AttnBackend driven by the REFERENCE's own objects.

The unmodified reference classes — ForwardBatch, ForwardMode, RadixAttention, MHATokenToKVPool, ReqToTokenPool
(python/sglang/srt/model_executor/forward_batch_info.py:164-300, layers/radix_attention.py:84-110,
mem_cache/memory_pool.py:47-94, :215-407) — are imported from /root/reference through oracle/ref_shims.py, a batch is
built the way the reference's backend unit test builds one (python/sglang/test/attention/test_flashattn_backend.py:
15-69, :163-200), and `RadixAttention.forward(q, k, v, forward_batch)` is called with `forward_batch.attn_backend` set
to OUR backend.  Every attribute the backend reads must therefore exist on the reference objects with the dtype /
layout the op wrappers assert.

There is no GPU here, so the C-ABI library is replaced by a recording fake: the REAL wrappers in
sglang_npu_b200/kernels.py still run (their stride / dtype assertions and argument marshalling see the reference's
tensors), the fake records which C entry point was called with how many arguments, and the result is then computed by
the oracle from the very arguments the wrapper received.  The output and the KV pool contents are compared with the
reference's own TorchNativeAttnBackend run on an identical second set of reference objects.

Runs only where /root/reference exists (this container); skipped on the GPU box."""
import types

import pytest
import torch

from oracle import attention_oracle as ao
from oracle import ref_shims
from sglang_npu_b200 import _cabi, kernels, synth
from sglang_npu_b200.srt.layers.attention import b200_backend

pytestmark = pytest.mark.skipif(not ref_shims.reference_available(), reason="needs /root/reference (build container only)")


class _FakeLib:
    """Stands in for libb200_hotpath.so: records (symbol, argc), returns B200_OK."""

    def __init__(self):
        self.calls = []

    def __getattr__(self, name):
        if not name.startswith("b200_"):
            raise AttributeError(name)

        def fn(*args):
            self.calls.append((name, len(args)))
            if name == "b200_decode_attn_workspace_bytes":
                batch, hq, d, splits, out = args
                out._obj.value = batch * hq * splits * (d + 1) * 4
            return 0

        return fn


class _KernelShim:
    """kernels.* with the three attention entry points followed by an oracle evaluation of the same arguments."""

    def __init__(self, real):
        self._real = real
        self.ops = []

    def __getattr__(self, name):
        return getattr(self._real, name)

    def set_kv_buffer(self, k_cache, v_cache, loc, cache_k, cache_v, k_scale=None, v_scale=None):
        self._real.set_kv_buffer(k_cache, v_cache, loc, cache_k, cache_v, k_scale, v_scale)
        self.ops.append("set_kv_buffer")
        ao.set_kv_buffer(k_cache, v_cache, loc, cache_k.view(loc.numel(), *k_cache.shape[1:]),
                         cache_v.view(loc.numel(), *v_cache.shape[1:]))

    def decode_attention_fwd(self, q, k_buffer, v_buffer, o, req_to_token, req_pool_indices, seq_lens, workspace,
                             kv_chunk, max_splits, sm_scale, logit_cap=0.0, int8_out=False, **kw):
        self._real.decode_attention_fwd(q, k_buffer, v_buffer, o, req_to_token, req_pool_indices, seq_lens, workspace,
                                        kv_chunk, max_splits, sm_scale, logit_cap, int8_out, **kw)
        self.ops.append("decode_attention_fwd")
        assert kv_chunk % 16 == 0 and max_splits >= 1
        assert max_splits == 1 or workspace is not None
        o.copy_(ao.decode_attention(q, k_buffer, v_buffer, req_to_token, req_pool_indices, seq_lens, sm_scale))

    def extend_attention_fwd(self, q, k_new, v_new, o, k_buffer, v_buffer, req_to_token, req_pool_indices, qo_indptr,
                             prefix_lens, max_extend_len, sm_scale, logit_cap=0.0, causal=True):
        self._real.extend_attention_fwd(q, k_new, v_new, o, k_buffer, v_buffer, req_to_token, req_pool_indices,
                                        qo_indptr, prefix_lens, max_extend_len, sm_scale, logit_cap, causal)
        self.ops.append("extend_attention_fwd")
        ext = (qo_indptr[1:] - qo_indptr[:-1])
        assert int(ext.max()) == max_extend_len and int(qo_indptr[-1]) == q.shape[0]
        seq = prefix_lens.long() + ext.long()
        o.copy_(ao.extend_attention(q, k_buffer, v_buffer, req_to_token, req_pool_indices, seq, prefix_lens, ext,
                                    sm_scale, causal))


@pytest.fixture()
def fake_device(monkeypatch):
    lib = _FakeLib()
    shim = _KernelShim(kernels)
    monkeypatch.setattr(_cabi, "load", lambda: lib)
    monkeypatch.setattr(kernels, "require_cuda", lambda *t: None)
    monkeypatch.setattr(kernels, "stream_ptr", lambda device=None: 0)
    monkeypatch.setattr(b200_backend, "kernels", shim)
    monkeypatch.setattr(b200_backend.B200AttnBackend, "_require_device", staticmethod(lambda device: None))
    monkeypatch.setattr(b200_backend.B200AttnBackend, "_device_sm_count", staticmethod(lambda device: 148))
    return lib, shim


def _reference_objects(ref, c, dtype):
    pb = synth.make_paged_batch(c["lens"], c["page"], seed=77)
    r2t_pool = ref["ReqToTokenPool"](size=pb.req_to_token.shape[0], max_context_len=pb.req_to_token.shape[1],
                                     device="cpu", enable_memory_saver=False)
    r2t_pool.req_to_token.copy_(pb.req_to_token)
    kv_pool = ref["MHATokenToKVPool"](size=pb.num_slots, page_size=c["page"], dtype=dtype, head_num=c["hkv"],
                                      head_dim=c["d"], layer_num=2, device="cpu", enable_memory_saver=False)
    k0, v0 = synth.randn_kv(pb.num_slots, c["page"], c["hkv"], c["d"], dtype, seed=7)
    for l in range(2):
        kv_pool.k_buffer[l].copy_(k0)
        kv_pool.v_buffer[l].copy_(v0)
    layer = ref["RadixAttention"](num_heads=c["hq"], head_dim=c["d"], scaling=c["d"] ** -0.5, num_kv_heads=c["hkv"],
                                  layer_id=1)
    return pb, r2t_pool, kv_pool, layer


def _forward_batch(ref, c, pb, backend, r2t_pool, kv_pool):
    FM = ref["ForwardMode"]
    bs = len(c["lens"])
    seq_lens = pb.seq_lens
    if c["mode"] == "decode":
        T = bs
        out_loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], seq_lens[i] - 1] for i in range(bs)]).long()
        fb = ref["ForwardBatch"](forward_mode=FM.DECODE, batch_size=bs, input_ids=torch.zeros(T, dtype=torch.long),
                                 req_pool_indices=pb.req_pool_indices, seq_lens=seq_lens, out_cache_loc=out_loc,
                                 seq_lens_sum=int(seq_lens.sum()), seq_lens_cpu=seq_lens.clone(), attn_backend=backend)
    else:
        prefix = torch.tensor(c["prefix"], dtype=torch.int32)
        ext = (seq_lens - prefix).to(torch.int32)
        T = int(ext.sum())
        out_loc = torch.cat([pb.req_to_token[pb.req_pool_indices[i], int(prefix[i]): int(seq_lens[i])]
                             for i in range(bs)]).long()
        fb = ref["ForwardBatch"](forward_mode=FM.EXTEND, batch_size=bs, input_ids=torch.zeros(T, dtype=torch.long),
                                 req_pool_indices=pb.req_pool_indices, seq_lens=seq_lens, out_cache_loc=out_loc,
                                 seq_lens_sum=int(seq_lens.sum()), attn_backend=backend, extend_num_tokens=T,
                                 extend_seq_lens=ext, extend_prefix_lens=prefix,
                                 extend_prefix_lens_cpu=prefix.tolist(), extend_seq_lens_cpu=ext.tolist())
    fb.req_to_token_pool = r2t_pool
    fb.token_to_kv_pool = kv_pool
    return fb, T


CASES = [
    dict(mode="decode", hq=8, hkv=2, d=128, page=16, lens=[37, 16, 1, 130, 700]),
    dict(mode="decode", hq=14, hkv=2, d=64, page=1, lens=[5, 64, 33]),
    dict(mode="extend", hq=8, hkv=2, d=128, page=16, lens=[70, 33, 16], prefix=[32, 0, 15]),
    dict(mode="extend", hq=4, hkv=4, d=128, page=16, lens=[9, 65], prefix=[0, 0]),
]


@pytest.mark.parametrize("c", CASES, ids=lambda c: f"{c['mode']}-hq{c['hq']}-p{c['page']}")
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_backend_runs_on_reference_objects(fake_device, c, dtype):
    lib, shim = fake_device
    ref = ref_shims.load_reference_attention()
    # --- side A: the reference's own backend on reference objects
    pb, r2t_a, kv_a, layer_a = _reference_objects(ref, c, dtype)
    native = ref["TorchNativeAttnBackend"](types.SimpleNamespace(device="cpu"))
    fb_a, T = _forward_batch(ref, c, pb, native, r2t_a, kv_a)
    g = torch.Generator().manual_seed(5)
    q = torch.randn((T, c["hq"] * c["d"]), generator=g).to(dtype)
    k = torch.randn((T, c["hkv"] * c["d"]), generator=g).to(dtype)
    v = torch.randn((T, c["hkv"] * c["d"]), generator=g).to(dtype)
    native.init_forward_metadata(fb_a)
    o_ref = layer_a(q.clone(), k.clone(), v.clone(), fb_a)

    # --- side B: OUR backend on a second, identical set of reference objects
    pb_b, r2t_b, kv_b, layer_b = _reference_objects(ref, c, dtype)
    model_config = types.SimpleNamespace(num_attention_heads=c["hq"], context_len=pb_b.req_to_token.shape[1],
                                         get_num_kv_heads=lambda tp: c["hkv"] // tp)
    runner = types.SimpleNamespace(device="cpu", req_to_token_pool=r2t_b, token_to_kv_pool=kv_b, tp_size=1,
                                   model_config=model_config, sliding_window_size=None)
    ours = b200_backend.B200AttnBackend(runner)
    assert isinstance(ours.forward_metadata, type(None))
    fb_b, _ = _forward_batch(ref, c, pb_b, ours, r2t_b, kv_b)
    ours.init_forward_metadata(fb_b)
    o = layer_b(q.clone(), k.clone(), v.clone(), fb_b)  # reference RadixAttention.forward -> backend.forward dispatch

    # the dispatch of the reference's base class reached the right entry point and C symbols
    want_op = "decode_attention_fwd" if c["mode"] == "decode" else "extend_attention_fwd"
    assert shim.ops == ["set_kv_buffer", want_op]
    syms = [s for s, _ in lib.calls if s != "b200_decode_attn_workspace_bytes"]
    assert syms == ["b200_set_kv_buffer", "b200_decode_attn" if c["mode"] == "decode" else "b200_extend_attn"]
    for name, argc in lib.calls:  # marshalled argument count == the ctypes signature declared for the symbol
        assert argc == len(_cabi._SIGNATURES[name]), name

    # same output shape / dtype as the reference backend returns, same numbers (oracle == torch_native up to bf16)
    assert o.shape == o_ref.shape and o.dtype == o_ref.dtype
    assert torch.allclose(o.float(), o_ref.float(), atol=2e-2, rtol=0)
    # set_kv_buffer wrote exactly the slots the reference wrote, in layer 1 only
    assert torch.equal(kv_b.k_buffer[1], kv_a.k_buffer[1]) and torch.equal(kv_b.v_buffer[1], kv_a.v_buffer[1])
    assert torch.equal(kv_b.k_buffer[0], kv_a.k_buffer[0])


def test_backend_graph_hooks_accept_reference_forward_mode(fake_device):
    """cuda_graph_runner.py:283-288, :642, :773 call the three graph hooks with the reference's ForwardMode."""
    ref = ref_shims.load_reference_attention()
    c = CASES[0]
    pb, r2t, kv, _ = _reference_objects(ref, c, torch.bfloat16)
    model_config = types.SimpleNamespace(num_attention_heads=c["hq"], context_len=4096,
                                         get_num_kv_heads=lambda tp: c["hkv"] // tp)
    runner = types.SimpleNamespace(device="cpu", req_to_token_pool=r2t, token_to_kv_pool=kv, tp_size=1,
                                   model_config=model_config, sliding_window_size=None)
    ours = b200_backend.B200AttnBackend(runner)
    FM = ref["ForwardMode"]
    ours.init_cuda_graph_state(8, 8)
    ours.init_forward_metadata_capture_cuda_graph(8, 8, pb.req_pool_indices, pb.seq_lens, None, FM.DECODE, None)
    md = ours.forward_metadata
    assert md.kv_chunk % 128 == 0 and md.max_splits == (4096 + md.kv_chunk - 1) // md.kv_chunk
    assert md.max_splits == 1 or md.workspace is not None
    ours.init_forward_metadata_replay_cuda_graph(8, pb.req_pool_indices, pb.seq_lens, int(pb.seq_lens.sum()), None,
                                                 FM.DECODE, None, pb.seq_lens)
    assert ours.get_cuda_graph_seq_len_fill_value() == 1
    with pytest.raises(NotImplementedError):
        ours.init_forward_metadata_capture_cuda_graph(8, 8, pb.req_pool_indices, pb.seq_lens, None, FM.EXTEND, None)


def test_reference_objects_expose_every_attribute_the_backend_reads():
    """Static check: each `forward_batch.X` / `layer.X` / `pool.X` the backend source touches exists on the reference
    class (dataclass field, attribute set in __init__, or method)."""
    import inspect
    import re

    ref = ref_shims.load_reference_attention()
    src = inspect.getsource(b200_backend)
    fb_attrs = set(re.findall(r"forward_batch\.([a-z_]+)", src))
    optional_extensions = {"shared_prefix_len"}  # read with getattr(..., 0): an extension, absent upstream
    fb_fields = set(ref["ForwardBatch"].__dataclass_fields__)
    missing = fb_attrs - fb_fields - optional_extensions
    assert not missing, f"ForwardBatch lacks {missing}"
    layer = ref["RadixAttention"](num_heads=8, head_dim=128, scaling=1.0, num_kv_heads=2, layer_id=0)
    layer_attrs = set(re.findall(r"layer\.([a-z_]+)", src)) - {"k_scale_float", "v_scale_float"}
    assert all(hasattr(layer, a) for a in layer_attrs), [a for a in layer_attrs if not hasattr(layer, a)]
    pool_attrs = set(re.findall(r"pool\.([a-z_]+)", src)) - {"req_to_token", "py"}  # "memory_pool.py" in comments
    assert all(hasattr(ref["MHATokenToKVPool"], a) or a in ("dtype", "size") for a in pool_attrs), pool_attrs
