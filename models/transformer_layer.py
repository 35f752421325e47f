"""Drop-in for the reference's models/transformer_layer.py: one DiT block —
  x += gate_a * Attn(LN(x) * (1 + scale_a) + shift_a)
  x += CrossAttn(LN(x), context_proj(context))                  (text conditioning only; un-gated)
  x += gate_m * MLP(LN(x) * (1 + scale_m) + shift_m)
with the six adaLN vectors chunked from Linear(ReLU(condition)). Every LayerNorm is one row kernel, every residual
update rides in the epilogue of the GEMM that produces it."""
import torch
import torch.nn as nn

import sdb200  # noqa: F401
from sdb200 import engine as E
from sdb200 import runtime as rt
from models.attention import Attention
from models.multihead_attention import CustomMultiheadAttention


class TransformerLayer(nn.Module):
    def __init__(self, config, *, cross_attn=False, context_dim=None):
        super().__init__()
        self.hidden_size = config['hidden_size']
        self.cross_attn = cross_attn
        self.context_dim = context_dim
        h = self.hidden_size
        self.att_norm = nn.LayerNorm(h, elementwise_affine=False, eps=1E-6)
        self.attn_block = Attention(config)
        self.ff_norm = nn.LayerNorm(h, elementwise_affine=False, eps=1E-6)
        self.mlp_block = nn.Sequential(nn.Linear(h, 4 * h), nn.ReLU(), nn.Linear(4 * h, h))
        if self.cross_attn:
            assert self.context_dim is not None, "Context dimension must be provided for cross attention"
            self.cross_attn_norm = nn.LayerNorm(h, elementwise_affine=False, eps=1E-6)
            self.cross_attn_block = CustomMultiheadAttention(h, config['num_heads'], batch_first=True)
            self.context_proj = nn.Linear(self.context_dim, h)
        self.adaptive_norm_layer = nn.Sequential(nn.ReLU(), nn.Linear(h, 6 * h, bias=True))

        nn.init.xavier_uniform_(self.mlp_block[0].weight)
        nn.init.constant_(self.mlp_block[0].bias, 0)
        nn.init.xavier_uniform_(self.mlp_block[-1].weight)
        nn.init.constant_(self.mlp_block[-1].bias, 0)
        nn.init.constant_(self.adaptive_norm_layer[-1].weight, 0)
        nn.init.constant_(self.adaptive_norm_layer[-1].bias, 0)
        if self.cross_attn:
            nn.init.xavier_uniform_(self.context_proj.weight)
            nn.init.constant_(self.context_proj.bias, 0)

    def run(self, x, mod6, n, L, ctx=None):
        """x: bf16 rows [n*L, hidden]; mod6: bf16 [n, 6*hidden] = adaptive_norm_layer(condition) (any row pitch)."""
        h = self.hidden_size
        shift_a, scale_a, gate_a, shift_m, scale_m, gate_m = (mod6[:, i * h:(i + 1) * h] for i in range(6))
        y = E.layer_norm_mod(x, self.att_norm, n, L, shift_a, scale_a)
        x = E.gated_linear(self.attn_block.run(y, n, L), self.attn_block.output_proj[0], x, gate_a, n, L)
        if self.cross_attn and ctx is not None:
            assert ctx.dim == self.context_dim, "Context shape does not match B,_,CONTEXT_DIM"
            x = E.dit_cross_attention(x, self.cross_attn_norm, self.cross_attn_block, self.context_proj, ctx, n, L)
        y = E.layer_norm_mod(x, self.ff_norm, n, L, shift_m, scale_m)
        y = E.linear(y, self.mlp_block[0], act=E._act_name(self.mlp_block[1]))
        return E.gated_linear(y, self.mlp_block[2], x, gate_m, n, L)

    def forward(self, x, condition, context=None):
        if not x.is_cuda:
            raise rt.B200Error("TransformerLayer.forward: CUDA tensors required (no CPU implementation in this package)")
        n, L, c = x.shape
        with torch.autocast("cuda", enabled=False):
            cond = E.activate_temb(condition.reshape(-1, c), nn.ReLU)
            if cond.shape[0] == 1 and n > 1:
                cond = cond.expand(n, c).contiguous()
            mod6 = E.linear(cond, self.adaptive_norm_layer[1])
            ctx = E.prepare_context(context) if context is not None else None
            out = self.run(x.reshape(n * L, c).to(torch.bfloat16), mod6, n, L, ctx)
        return out.reshape(n, L, c).to(x.dtype)
