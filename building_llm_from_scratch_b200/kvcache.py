"""KV cache for the generation path.

The reference's generate.py:4-75 re-runs the model on the whole window for every new token. Here the
prompt is run once (prefill, the training-path attention kernels) while the rotated keys and the values
of every block are kept; each further token runs one-row projections, RoPE at its own position and
`b200_attention_decode_bf16` against the cache. Inference only (no autograd)."""
from __future__ import annotations

import torch

from . import kernels as K
from . import ops


class KVCache:
    def __init__(self, n_layers: int, batch: int, t_max: int, kv_width: int, device, dtype=torch.bfloat16):
        self.k = [torch.empty((batch, t_max, kv_width), device=device, dtype=dtype) for _ in range(n_layers)]
        self.v = [torch.empty((batch, t_max, kv_width), device=device, dtype=dtype) for _ in range(n_layers)]
        self.t_max = t_max
        self.pos = 0          # tokens already cached (advanced by the model after all blocks ran)


def packed_qkv(x, modules):
    """[B*T, q | k | v] of the three projections of x (one packed GEMM when they can be fused)."""
    rows = x.numel() // x.shape[-1]
    qkv = ops.linears_packed(x, modules)
    if qkv is not None:
        return qkv.reshape(rows, -1)
    return torch.cat([t.reshape(rows, -1) for t in ops.linears_of_same_input(x, modules)], dim=1)


def attend_with_cache(qkv, cos, sin, cache: KVCache, layer: int, B: int, T: int, h: int, hkv: int, hd: int):
    """Attention of T new tokens per sequence (prefill: T > 1 into an empty cache; decode: T == 1) ->
    context [B*T, h*hd]. cos/sin: bf16 RoPE tables [ctx, hd] or None."""
    nq, nk = h * hd, hkv * hd
    pos = cache.pos
    if pos + T > cache.t_max:
        raise RuntimeError(f"KV cache holds {cache.t_max} positions, {pos + T} needed")
    if T > 1:
        if pos != 0:
            raise RuntimeError("multi-token steps are only supported as the prefill of an empty cache")
        o, _, qk = ops.attention_qkv(qkv, cos, sin, B, T, h, hkv, hd)
        k_new = (qk if cos is not None else qkv)[:, nq:nq + nk]
        cache.k[layer][:, :T].copy_(k_new.view(B, T, nk))
        cache.v[layer][:, :T].copy_(qkv[:, nq + nk:].view(B, T, nk))
        return o
    if cos is not None:
        qk = K.rope(qkv, cos[pos:pos + 1], sin[pos:pos + 1], B, 1, h + hkv, hd)   # table row 0 = position pos
    else:
        qk = qkv
    cache.k[layer][:, pos].copy_(qk[:, nq:nq + nk])
    cache.v[layer][:, pos].copy_(qkv[:, nq + nk:])
    return K.attention_decode(qk[:, :nq], cache.k[layer], cache.v[layer], pos + 1, h, hkv, hd)
