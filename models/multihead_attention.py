"""Drop-in for the reference's models/multihead_attention.py: `CustomMultiheadAttention` keeps q/k/v/out projections
as separate nn.Linear leaves (so the CIM pipeline can swap them for quantised layers) behind the
nn.MultiheadAttention call signature. Inside the fused blocks the engine reads its leaves directly; the standalone
`forward` below serves callers that use it as a module (batch_first inputs, CUDA only)."""
import math

import torch
import torch.nn as nn

import sdb200  # noqa: F401
from sdb200 import engine as E
from sdb200 import runtime as rt


class CustomMultiheadAttention(nn.Module):
    def __init__(self, embed_dim, num_heads, dropout=0.0, bias=True, batch_first=False):
        super().__init__()
        if embed_dim % num_heads != 0:
            raise ValueError("embed_dim must be divisible by num_heads")
        self.embed_dim = embed_dim
        self.num_heads = num_heads
        self.head_dim = embed_dim // num_heads
        self.batch_first = batch_first
        self.q_proj = nn.Linear(embed_dim, embed_dim, bias=bias)
        self.k_proj = nn.Linear(embed_dim, embed_dim, bias=bias)
        self.v_proj = nn.Linear(embed_dim, embed_dim, bias=bias)
        self.out_proj = nn.Linear(embed_dim, embed_dim, bias=bias)
        self.dropout = nn.Dropout(dropout) if dropout > 0.0 else nn.Identity()
        self.scaling = 1.0 / math.sqrt(self.head_dim)

    def forward(self, query, key=None, value=None, attn_mask=None, key_padding_mask=None, need_weights=True,
                average_attn_weights=True):
        """Returns (attn_output, None): the averaged attention weights the reference can return are never consumed
        on the hot path (every caller does `out, _ = ...`) and are not materialised by the flash kernel."""
        if attn_mask is not None or key_padding_mask is not None:
            raise rt.B200Error("CustomMultiheadAttention: masks have no CUDA path")
        if not isinstance(self.dropout, nn.Identity) and self.training:
            raise rt.B200Error("CustomMultiheadAttention: attention dropout has no CUDA path")
        key = query if key is None else key
        value = key if value is None else value
        if value is not key:
            raise rt.B200Error("CustomMultiheadAttention: key and value must be the same tensor")
        if not self.batch_first:
            query, key = query.transpose(0, 1), key.transpose(0, 1)
        b, lq, c = query.shape
        lk = key.shape[1]
        q2 = rt.cast_bf16(query.float().contiguous()).reshape(b * lq, c)
        pk = E._mha_weights(self)
        if key is query:
            qkv = rt.linear(q2, pk["qkv"], bias=pk["b_qkv"])
            q, k, v = qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:]
        else:
            k2 = rt.cast_bf16(key.float().contiguous()).reshape(b * lk, c)
            q = rt.linear(q2, pk["q"], bias=pk["b_q"])
            kv = rt.linear(k2, pk["kv"], bias=pk["b_kv"])
            k, v = kv[:, :c], kv[:, c:]
        o = rt.attention(q, k, v, b, self.num_heads, lq, lk, self.head_dim, scale=self.scaling)
        out = E.linear(o, self.out_proj, out_dtype=torch.float32).reshape(b, lq, c)
        if not self.batch_first:
            out = out.transpose(0, 1)
        return out, None
