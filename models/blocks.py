"""Drop-in for the reference's models/blocks.py: same class names, constructor signatures, sub-module names and
parameter creation order (so `state_dict()` keys, seeded initialisation and `load_state_dict(strict=True)` are
interchangeable with the reference), but the forward pass runs on the fused sm_100a engine
(stablediffusion-pytorch_b200/engine.py) instead of a chain of ATen / cuDNN / cuBLAS calls.

Each block has two entry points:
  forward(x, ...)  — the reference signature on NCHW tensors (converts at the boundary, for standalone use)
  run(act, ...)    — the engine-internal path on NHWC bf16 `Act`s used by Unet / VQVAE / DIT
"""
import torch
import torch.nn as nn

import sdb200  # noqa: F401
from sdb200 import engine as E
from sdb200 import runtime as rt


def get_time_embedding(time_steps, temb_dim):
    r"""Sinusoidal timestep embedding, (B,) int -> (B, temb_dim) fp32 — reference models/blocks.py:5-24.
    One kernel instead of arange/pow/repeat/div/sin/cos/cat."""
    assert temb_dim % 2 == 0, "time embedding dimension must be divisible by 2"
    if not time_steps.is_cuda:
        raise rt.B200Error("get_time_embedding: CUDA tensor required (no CPU implementation in this package)")
    t = time_steps.reshape(-1).to(torch.int64)
    return rt.time_embedding(t, t.numel(), temb_dim, dtype=torch.float32)


def _norm_act_conv(norm_channels, cin, cout, act_cls):
    return nn.Sequential(nn.GroupNorm(norm_channels, cin), act_cls(), nn.Conv2d(cin, cout, kernel_size=3, stride=1, padding=1))


class _Stack(nn.Module):
    """Shared machinery of the four block types: a sequence of ResNet units, each optionally followed by
    self-attention and cross-attention. Sub-classes only decide how many units exist and what wraps them."""
    ACT = nn.SiLU
    MHA = nn.MultiheadAttention

    # --- construction helpers (called in the reference's parameter-creation order) -----------------------
    def _make_resnets(self, cins, cout, norm_channels):
        self.resnet_conv_first = nn.ModuleList([_norm_act_conv(norm_channels, c, cout, self.ACT) for c in cins])
        if self.t_emb_dim is not None:
            self.t_emb_layers = nn.ModuleList(
                [nn.Sequential(self.ACT(), nn.Linear(self.t_emb_dim, cout)) for _ in cins])
        self.resnet_conv_second = nn.ModuleList([_norm_act_conv(norm_channels, cout, cout, self.ACT) for _ in cins])

    def _make_self_attn(self, n, cout, norm_channels, num_heads):
        self.attention_norms = nn.ModuleList([nn.GroupNorm(norm_channels, cout) for _ in range(n)])
        self.attentions = nn.ModuleList([self.MHA(cout, num_heads, batch_first=True) for _ in range(n)])

    def _make_cross_attn(self, n, cout, norm_channels, num_heads, context_dim):
        assert context_dim is not None, "Context Dimension must be passed for cross attention"
        self.cross_attention_norms = nn.ModuleList([nn.GroupNorm(norm_channels, cout) for _ in range(n)])
        self.cross_attentions = nn.ModuleList([self.MHA(cout, num_heads, batch_first=True) for _ in range(n)])
        self.context_proj = nn.ModuleList([nn.Linear(context_dim, cout) for _ in range(n)])

    def _make_shortcuts(self, cins, cout):
        self.residual_input_conv = nn.ModuleList([nn.Conv2d(c, cout, kernel_size=1) for c in cins])

    # --- execution helpers -------------------------------------------------------------------------------
    def _resnet(self, a, i, temb_act):
        tl = self.t_emb_layers[i] if self.t_emb_dim is not None else None
        if tl is not None and temb_act is None:
            raise rt.B200Error("block was built with a time embedding but t_emb is None")
        return E.resnet(a, self.resnet_conv_first[i], tl, self.resnet_conv_second[i], self.residual_input_conv[i],
                        temb_act)

    def _self_attn(self, a, i):
        return E.self_attention(a, self.attention_norms[i], self.attentions[i])

    def _cross_attn(self, a, i, ctx):
        assert ctx is not None, "context cannot be None if cross attention layers are used"
        ctx_t, lk, dim = ctx
        assert dim == self.context_dim, "Context shape does not match B,_,CONTEXT_DIM"
        return E.cross_attention(a, self.cross_attention_norms[i], self.cross_attentions[i], self.context_proj[i],
                                 ctx_t, lk, cache=ctx)

    # --- NCHW boundary helpers -----------------------------------------------------------------------------
    def _prep_temb(self, t_emb):
        if t_emb is None or self.t_emb_dim is None:
            return None
        return E.activate_temb(t_emb, self.ACT)

    @staticmethod
    def _prep_ctx(context, batch):
        if context is None:
            return None
        assert len(context.shape) == 3 and context.shape[0] == batch, "Context shape does not match B,_,CONTEXT_DIM"
        return E.prepare_context(context)


class DownBlock(_Stack):
    r"""ResNet (+ self-attention) (+ cross-attention) x num_layers, then a 4x4 stride-2 convolution —
    reference models/blocks.py:27-146."""

    def __init__(self, in_channels, out_channels, t_emb_dim, down_sample, num_heads, num_layers, attn,
                 norm_channels, cross_attn=False, context_dim=None):
        super().__init__()
        self.num_layers = num_layers
        self.down_sample = down_sample
        self.attn = attn
        self.context_dim = context_dim
        self.cross_attn = cross_attn
        self.t_emb_dim = t_emb_dim
        cins = [in_channels if i == 0 else out_channels for i in range(num_layers)]
        self._make_resnets(cins, out_channels, norm_channels)
        if self.attn:
            self._make_self_attn(num_layers, out_channels, norm_channels, num_heads)
        if self.cross_attn:
            self._make_cross_attn(num_layers, out_channels, norm_channels, num_heads, context_dim)
        self._make_shortcuts(cins, out_channels)
        self.down_sample_conv = nn.Conv2d(out_channels, out_channels, 4, 2, 1) if self.down_sample else nn.Identity()

    def run(self, a, temb_act=None, ctx=None):
        for i in range(self.num_layers):
            a = self._resnet(a, i, temb_act)
            if self.attn:
                a = self._self_attn(a, i)
            if self.cross_attn:
                a = self._cross_attn(a, i, ctx)
        if self.down_sample:
            a = E.conv2d(a, self.down_sample_conv)
        return a

    def forward(self, x, t_emb=None, context=None):
        return self.run(E.Act.from_nchw(x), self._prep_temb(t_emb), self._prep_ctx(context, x.shape[0])).to_nchw()


class MidBlock(_Stack):
    r"""ResNet, then (self-attention (+ cross-attention), ResNet) x num_layers — reference models/blocks.py:149-267."""

    def __init__(self, in_channels, out_channels, t_emb_dim, num_heads, num_layers, norm_channels, cross_attn=None,
                 context_dim=None):
        super().__init__()
        self.num_layers = num_layers
        self.t_emb_dim = t_emb_dim
        self.context_dim = context_dim
        self.cross_attn = cross_attn
        cins = [in_channels if i == 0 else out_channels for i in range(num_layers + 1)]
        self._make_resnets(cins, out_channels, norm_channels)
        self._make_self_attn(num_layers, out_channels, norm_channels, num_heads)
        if self.cross_attn:
            self._make_cross_attn(num_layers, out_channels, norm_channels, num_heads, context_dim)
        self._make_shortcuts(cins, out_channels)

    def run(self, a, temb_act=None, ctx=None):
        a = self._resnet(a, 0, temb_act)
        for i in range(self.num_layers):
            a = self._self_attn(a, i)
            if self.cross_attn:
                a = self._cross_attn(a, i, ctx)
            a = self._resnet(a, i + 1, temb_act)
        return a

    def forward(self, x, t_emb=None, context=None):
        return self.run(E.Act.from_nchw(x), self._prep_temb(t_emb), self._prep_ctx(context, x.shape[0])).to_nchw()


class UpBlock(_Stack):
    r"""ConvTranspose 4x4 s2 upsample, optional skip concat, then ResNet (+ self-attention) x num_layers — the VQVAE
    decoder block, reference models/blocks.py:270-370."""

    def __init__(self, in_channels, out_channels, t_emb_dim, up_sample, num_heads, num_layers, attn, norm_channels):
        super().__init__()
        self.num_layers = num_layers
        self.up_sample = up_sample
        self.t_emb_dim = t_emb_dim
        self.attn = attn
        cins = [in_channels if i == 0 else out_channels for i in range(num_layers)]
        self._make_resnets(cins, out_channels, norm_channels)
        if self.attn:
            self._make_self_attn(num_layers, out_channels, norm_channels, num_heads)
        self._make_shortcuts(cins, out_channels)
        self.up_sample_conv = nn.ConvTranspose2d(in_channels, in_channels, 4, 2, 1) if self.up_sample else nn.Identity()

    def run(self, a, skip=None, temb_act=None):
        a = E.upsample_concat(a, self.up_sample_conv if self.up_sample else None, skip)
        for i in range(self.num_layers):
            a = self._resnet(a, i, temb_act)
            if self.attn:
                a = self._self_attn(a, i)
        return a

    def forward(self, x, out_down=None, t_emb=None):
        skip = None if out_down is None else E.Act.from_nchw(out_down)
        return self.run(E.Act.from_nchw(x), skip, self._prep_temb(t_emb)).to_nchw()


class UpBlockUnet(_Stack):
    r"""ConvTranspose upsample of the C/2 input, concat with the C/2 skip, then ResNet + self-attention
    (+ cross-attention) x num_layers — reference models/blocks.py:373-499. Self-attention is unconditional here."""

    def __init__(self, in_channels, out_channels, t_emb_dim, up_sample, num_heads, num_layers, norm_channels,
                 cross_attn=False, context_dim=None):
        super().__init__()
        self.num_layers = num_layers
        self.up_sample = up_sample
        self.t_emb_dim = t_emb_dim
        self.cross_attn = cross_attn
        self.context_dim = context_dim
        cins = [in_channels if i == 0 else out_channels for i in range(num_layers)]
        self._make_resnets(cins, out_channels, norm_channels)
        self._make_self_attn(num_layers, out_channels, norm_channels, num_heads)
        if self.cross_attn:
            self._make_cross_attn(num_layers, out_channels, norm_channels, num_heads, context_dim)
        self._make_shortcuts(cins, out_channels)
        self.up_sample_conv = nn.ConvTranspose2d(in_channels // 2, in_channels // 2, 4, 2, 1) \
            if self.up_sample else nn.Identity()

    def run(self, a, skip=None, temb_act=None, ctx=None):
        a = E.upsample_concat(a, self.up_sample_conv if self.up_sample else None, skip)
        for i in range(self.num_layers):
            a = self._resnet(a, i, temb_act)
            a = self._self_attn(a, i)
            if self.cross_attn:
                a = self._cross_attn(a, i, ctx)
        return a

    def forward(self, x, out_down=None, t_emb=None, context=None):
        skip = None if out_down is None else E.Act.from_nchw(out_down)
        return self.run(E.Act.from_nchw(x), skip, self._prep_temb(t_emb),
                        self._prep_ctx(context, x.shape[0])).to_nchw()
