"""Drop-in for the reference's models/attention.py: the DiT self-attention block (packed qkv Linear ->
softmax(QK^T / sqrt(d)) V -> output Linear) on the QKV GEMM + flash-attention kernels."""
import torch
import torch.nn as nn

import sdb200  # noqa: F401
from sdb200 import engine as E
from sdb200 import runtime as rt


class Attention(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.n_heads = config['num_heads']
        self.hidden_size = config['hidden_size']
        self.head_dim = config['head_dim']
        self.att_dim = self.n_heads * self.head_dim
        self.qkv_proj = nn.Linear(self.hidden_size, 3 * self.att_dim, bias=True)
        self.output_proj = nn.Sequential(nn.Linear(self.att_dim, self.hidden_size))
        nn.init.xavier_uniform_(self.qkv_proj.weight)
        nn.init.constant_(self.qkv_proj.bias, 0)
        nn.init.xavier_uniform_(self.output_proj[0].weight)
        nn.init.constant_(self.output_proj[0].bias, 0)

    def run(self, rows, n, L):
        """bf16 token rows [n*L, hidden] -> attention output rows [n*L, att_dim] (before output_proj, which the
        caller fuses with its gated residual)."""
        return E.dit_self_attention(rows, self, n, L)

    def forward(self, x):
        if not x.is_cuda:
            raise rt.B200Error("Attention.forward: CUDA tensors required (no CPU implementation in this package)")
        n, L, c = x.shape
        with torch.autocast("cuda", enabled=False):
            rows = x.reshape(n * L, c).to(torch.bfloat16)
            out = E.linear(self.run(rows, n, L), self.output_proj[0])
        return out.reshape(n, L, self.hidden_size).to(x.dtype)
