"""Tensor-parallel host logic on CPU: world_size-2 `gloo` processes (reference
python/sglang/srt/layers/linear.py:424-1309, distributed/communication_op.py:11-13).

The linears' `apply()` is CUDA-only (no CPU fallback in the product path), so what is checked here is everything
AROUND it: the TP group state, the column / merged / QKV / row shard loaders of every quant method, and the
all-reduce / all-gather wrappers — by evaluating each rank's shard with the ORACLE linear (test infrastructure)
and comparing the recombined result with the unsharded oracle."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _full_tensors(kind, n, k, seed):
    from sglang_npu_b200 import synth

    if kind == "int8":
        w, s = synth.int8_weight(n, k, seed)
        return {"weight": w, "weight_scale": s.view(n, 1)}
    if kind == "fp8":
        w, s = synth.fp8_block_weight(n, k, seed)
        return {"weight": w, "weight_scale_inv": s}
    qw, sc, qz = synth.awq_weight(k, n, 128, torch.float16, seed)
    return {"qweight": qw, "scales": sc, "qzeros": qz}


def _oracle_linear(kind, t, x):
    from oracle import quant_oracle as qo

    if kind == "int8":
        return qo.w8a8_int8_linear(x, t["weight"], t["weight_scale"]).float()
    if kind == "fp8":
        return qo.w8a8_block_fp8_linear(x, t["weight"], t["weight_scale_inv"]).float()
    return qo.awq_linear(x, t["qweight"], t["scales"], t["qzeros"], 128).float()


def _shard_tensors(kind, lin):
    if kind == "int8":
        return {"weight": lin.weight.data, "weight_scale": lin.weight_scale.data}
    if kind == "fp8":
        return {"weight": lin.weight.data, "weight_scale_inv": lin.weight_scale_inv.data}
    return {"qweight": lin.qweight.data, "scales": lin.scales.data, "qzeros": lin.qzeros.data}


def _worker(rank, world, port, kind, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from sglang_npu_b200.srt.distributed import parallel_state as ps
        from sglang_npu_b200.srt.layers.linear import MergedColumnParallelLinear, QKVParallelLinear, RowParallelLinear
        from sglang_npu_b200.srt.layers.quantization import get_quantization_config

        ps.init_tensor_parallel()
        assert ps.get_tensor_model_parallel_world_size() == world and ps.get_tensor_model_parallel_rank() == rank
        cfg_cls = get_quantization_config({"int8": "b200_w8a8_int8", "fp8": "b200_fp8", "awq": "b200_awq"}[kind])
        qc = cfg_cls.from_config({"weight_block_size": [128, 128], "activation_scheme": "dynamic", "quant_method": "fp8",
                                  "bits": 4, "group_size": 128, "zero_point": True, "w_bit": 4, "q_group_size": 128})
        dtype = torch.float16 if kind == "awq" else torch.bfloat16
        hidden, inter, hq, hkv, d = 512, 1024, 4, 2, 128
        x = (torch.randn(6, hidden, generator=torch.Generator().manual_seed(3)) * 0.3).to(dtype)

        # ---- column-parallel (merged gate_up): rank r holds [gate_r | up_r]; all-gather restores the full output
        gu = MergedColumnParallelLinear(hidden, [inter, inter], bias=False, quant_config=qc, params_dtype=dtype)
        fulls = [_full_tensors(kind, inter, hidden, 10 + i) for i in range(2)]
        for sid, t in enumerate(fulls):
            for name, val in t.items():
                gu.weight_loader(getattr(gu, name), val, sid)
        local = _oracle_linear(kind, _shard_tensors(kind, gu), x)  # [6, 2*inter/world]
        gathered = ps.tensor_model_parallel_all_gather(local, dim=-1)
        per = inter // world
        parts = gathered.view(6, world, 2, per)
        gate = parts[:, :, 0].reshape(6, inter)
        up = parts[:, :, 1].reshape(6, inter)
        torch.testing.assert_close(gate, _oracle_linear(kind, fulls[0], x), rtol=2e-2, atol=2e-2)
        torch.testing.assert_close(up, _oracle_linear(kind, fulls[1], x), rtol=2e-2, atol=2e-2)

        # ---- QKV: q heads split, kv heads split (tp <= Hkv)
        qkv = QKVParallelLinear(hidden, d, hq, hkv, bias=False, quant_config=qc, params_dtype=dtype)
        assert qkv.num_heads == hq // world and qkv.num_kv_heads == max(1, hkv // world)
        fq = {"q": _full_tensors(kind, hq * d, hidden, 20), "k": _full_tensors(kind, hkv * d, hidden, 21),
              "v": _full_tensors(kind, hkv * d, hidden, 22)}
        for sid, t in fq.items():
            for name, val in t.items():
                qkv.weight_loader(getattr(qkv, name), val, sid)
        local = _oracle_linear(kind, _shard_tensors(kind, qkv), x)
        qs, ks = qkv.num_heads * d, qkv.num_kv_heads * d
        ref_q = _oracle_linear(kind, fq["q"], x)[:, rank * qs:(rank + 1) * qs]
        ref_k = _oracle_linear(kind, fq["k"], x)[:, rank * ks:(rank + 1) * ks]
        ref_v = _oracle_linear(kind, fq["v"], x)[:, rank * ks:(rank + 1) * ks]
        torch.testing.assert_close(local, torch.cat([ref_q, ref_k, ref_v], -1), rtol=2e-2, atol=2e-2)

        # ---- row-parallel: K split, partial sums all-reduced (bf16 sum in the product; fp32 here)
        down = RowParallelLinear(inter, hidden, bias=False, quant_config=qc, params_dtype=dtype)
        fd = _full_tensors(kind, hidden, inter, 30)
        for name, val in fd.items():
            down.weight_loader(getattr(down, name), val)
        xin = (torch.randn(6, inter, generator=torch.Generator().manual_seed(4)) * 0.3).to(dtype)
        shard = inter // world
        x_local = xin[:, rank * shard:(rank + 1) * shard]
        partial = _oracle_linear(kind, _shard_tensors(kind, down), x_local)
        total = ps.tensor_model_parallel_all_reduce(partial.clone())
        if kind == "awq":  # exact: per-group dequantisation does not depend on the K split
            ref = _oracle_linear(kind, fd, xin)
            torch.testing.assert_close(total, ref, rtol=2e-2, atol=2e-2)
        else:  # int8 / block-fp8 activation scales are per (token, shard): compare with the sum of shard-wise oracles
            ref = sum(_oracle_linear(kind, {n_: (v.narrow(1, r * (v.shape[1] // world), v.shape[1] // world)
                                                 if v.shape[1] > 1 else v) for n_, v in fd.items()},
                                     xin[:, r * shard:(r + 1) * shard]) for r in range(world))
            torch.testing.assert_close(total, ref, rtol=1e-3, atol=1e-3)
        ps.destroy_tensor_parallel()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # surface the failure in the parent
        import traceback

        q.put((rank, traceback.format_exc()))


@pytest.mark.parametrize("kind", ["int8", "fp8", "awq"])
def test_tp2_shard_loaders_and_collectives(kind):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, kind, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", f"rank {rank}: {msg}"


def test_kv_head_replication_when_tp_exceeds_kv_heads():
    """tp > Hkv: one KV head per rank, replicated (llama.py:118-132, linear.py QKVParallelLinear)."""
    from sglang_npu_b200.srt.layers.linear import QKVParallelLinear
    from sglang_npu_b200.srt.model_executor.layer_stack import CONFIGS

    lin = QKVParallelLinear(256, 64, 8, 2, bias=False, params_dtype=torch.bfloat16, tp_rank=3, tp_size=4)
    assert lin.num_heads == 2 and lin.num_kv_heads == 1 and lin.num_kv_head_replicas == 2
    full_k = torch.arange(2 * 64 * 256, dtype=torch.float32).view(128, 256).bfloat16()
    lin.weight.data.zero_()
    lin.weight_loader(lin.weight, full_k, "k")  # the layer's own loader on the un-quantised fallback parameter
    assert torch.equal(lin.weight.data[128:192], full_k[64:128])  # rank 3 -> kv head 3 // 2 = 1
    assert CONFIGS["llama-3.1-70b"].get_num_kv_heads(8) == 1 and CONFIGS["qwen2.5-32b"].get_num_kv_heads(2) == 4


def test_unquantized_fallback_layers_shard_through_their_own_loaders():
    """Layers on the un-quantised fallback (quant ignore lists, quant_config=None) load at tp > 1 through the layer's
    own weight_loader: row-parallel narrows the input dim, merged-column shards each logical sub-matrix
    (reference linear.py:181-194 registers input_dim=1 / output_dim=0 on the plain weight)."""
    from sglang_npu_b200.srt.layers.linear import MergedColumnParallelLinear, RowParallelLinear

    full = torch.arange(32 * 64, dtype=torch.float32).view(32, 64).bfloat16()
    row = RowParallelLinear(64, 32, bias=False, params_dtype=torch.bfloat16, tp_rank=1, tp_size=2)
    row.weight_loader(row.weight, full)
    assert torch.equal(row.weight.data, full[:, 32:64])
    merged = MergedColumnParallelLinear(64, [32, 32], bias=False, params_dtype=torch.bfloat16, tp_rank=1, tp_size=2)
    merged.weight.data.zero_()
    merged.weight_loader(merged.weight, full, 0)
    merged.weight_loader(merged.weight, full + 1, 1)
    assert torch.equal(merged.weight.data[:16], full[16:32]) and torch.equal(merged.weight.data[16:], (full + 1)[16:32])
