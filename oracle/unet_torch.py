"""ORACLE — test infrastructure only (imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs; never by the product path).

Plain PyTorch fp32 CPU restatement of the reference's denoiser forward, written functionally over a state_dict so
it needs no module classes: UNet (models/unet_base.py, models/unet_cond_base.py, models/unet_cond_base_relu.py with
models/blocks.py / blocks_relu.py) and the VQVAE decoder (models/vqvae.py:141-153). The floating-point arithmetic of
conv / GroupNorm / softmax lives in PyTorch (un-pinned third-party dependency of the reference), so
this restatement calls the same torch.nn.functional primitives the reference's nn.Modules call.

Pinning: the reference ships no golden vectors. oracle/gen_golden.py runs the REAL reference modules in
the build container on seeded inputs and (a) asserts this restatement reproduces them, (b) commits the outputs as
tests/golden/*.pt; tests/test_oracle.py re-checks this file against those fixtures on every run.
"""
import math

import torch
import torch.nn.functional as F


def time_embedding(t, dim):
    """get_time_embedding — models/blocks.py:5-24."""
    half = dim // 2
    factor = 10000 ** (torch.arange(0, half, dtype=torch.float32, device=t.device) / half)
    e = t.reshape(-1)[:, None].repeat(1, half) / factor
    return torch.cat([torch.sin(e), torch.cos(e)], dim=-1)


def _act(x, relu):
    return F.relu(x) if relu else F.silu(x)


class _SD:
    """state_dict view with a key prefix."""

    def __init__(self, sd, prefix=""):
        self.sd, self.p = sd, prefix

    def sub(self, name):
        return _SD(self.sd, f"{self.p}{name}.")

    def __getitem__(self, k):
        return self.sd[self.p + k]

    def get(self, k):
        return self.sd.get(self.p + k)

    def has(self, k):
        return (self.p + k) in self.sd


def _gn(x, sd, groups):
    return F.group_norm(x, groups, sd["weight"], sd["bias"], 1e-5)


USE_SDPA = False   # bench.py's library-bar leg: let torch pick its fused attention (what nn.MultiheadAttention's fast
                   # path / F.scaled_dot_product_attention run on a GPU) instead of the explicit matmul-softmax-matmul


def _mha(sd, q_in, kv_in, heads):
    """nn.MultiheadAttention(batch_first) forward (packed in_proj) or CustomMultiheadAttention (q/k/v_proj) —
    models/blocks.py:128,140; models/multihead_attention.py:41-80. Explicit softmax(QK^T/sqrt(d))V."""
    if sd.has("in_proj_weight"):
        wq, wk, wv = sd["in_proj_weight"].chunk(3, dim=0)
        bq, bk, bv = sd["in_proj_bias"].chunk(3, dim=0)
    else:
        wq, wk, wv = sd["q_proj.weight"], sd["k_proj.weight"], sd["v_proj.weight"]
        bq, bk, bv = sd["q_proj.bias"], sd["k_proj.bias"], sd["v_proj.bias"]
    B, Lq, C = q_in.shape
    Lk = kv_in.shape[1]
    d = C // heads
    q = F.linear(q_in, wq, bq).view(B, Lq, heads, d).transpose(1, 2)
    k = F.linear(kv_in, wk, bk).view(B, Lk, heads, d).transpose(1, 2)
    v = F.linear(kv_in, wv, bv).view(B, Lk, heads, d).transpose(1, 2)
    if USE_SDPA:
        o = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(B, Lq, C)
    else:
        p = torch.softmax(q @ k.transpose(-2, -1) / math.sqrt(d), dim=-1)
        o = (p @ v).transpose(1, 2).reshape(B, Lq, C)
    return F.linear(o, sd["out_proj.weight"], sd["out_proj.bias"])


def _resnet(sd, i, x, t_emb, groups, relu):
    """GN-act-conv3x3, + t_emb projection, GN-act-conv3x3, + 1x1 shortcut — models/blocks.py:113-119."""
    f, s = sd.sub(f"resnet_conv_first.{i}"), sd.sub(f"resnet_conv_second.{i}")
    h = F.conv2d(_act(_gn(x, f.sub("0"), groups), relu), f["2.weight"], f["2.bias"], padding=1)
    if t_emb is not None:
        te = sd.sub(f"t_emb_layers.{i}")
        h = h + F.linear(_act(t_emb, relu), te["1.weight"], te["1.bias"])[:, :, None, None]
    h = F.conv2d(_act(_gn(h, s.sub("0"), groups), relu), s["2.weight"], s["2.bias"], padding=1)
    r = sd.sub(f"residual_input_conv.{i}")
    return h + F.conv2d(x, r["weight"], r["bias"])


def _self_attn(sd, i, x, groups, heads):
    B, C, H, W = x.shape
    a = _gn(x.reshape(B, C, H * W), sd.sub(f"attention_norms.{i}"), groups).transpose(1, 2)
    o = _mha(sd.sub(f"attentions.{i}"), a, a, heads)
    return x + o.transpose(1, 2).reshape(B, C, H, W)


def _cross_attn(sd, i, x, ctx, groups, heads):
    B, C, H, W = x.shape
    a = _gn(x.reshape(B, C, H * W), sd.sub(f"cross_attention_norms.{i}"), groups).transpose(1, 2)
    cp = sd.sub(f"context_proj.{i}")
    c = F.linear(ctx, cp["weight"], cp["bias"])
    o = _mha(sd.sub(f"cross_attentions.{i}"), a, c, heads)
    return x + o.transpose(1, 2).reshape(B, C, H, W)


def down_block(sd, x, t_emb, ctx, layers, groups, heads, attn, down_sample, relu):
    """DownBlock.forward — models/blocks.py:111-146."""
    for i in range(layers):
        x = _resnet(sd, i, x, t_emb, groups, relu)
        if attn:
            x = _self_attn(sd, i, x, groups, heads)
        if ctx is not None:
            x = _cross_attn(sd, i, x, ctx, groups, heads)
    if down_sample:
        x = F.conv2d(x, sd["down_sample_conv.weight"], sd["down_sample_conv.bias"], stride=2, padding=1)
    return x


def mid_block(sd, x, t_emb, ctx, layers, groups, heads, relu):
    """MidBlock.forward — models/blocks.py:225-267."""
    x = _resnet(sd, 0, x, t_emb, groups, relu)
    for i in range(layers):
        x = _self_attn(sd, i, x, groups, heads)
        if ctx is not None:
            x = _cross_attn(sd, i, x, ctx, groups, heads)
        x = _resnet(sd, i + 1, x, t_emb, groups, relu)
    return x


def up_block(sd, x, skip, t_emb, ctx, layers, groups, heads, up_sample, relu, attn=True):
    """UpBlockUnet.forward (models/blocks.py:461-499) and UpBlock.forward (:343-370, attn flag, no cross-attn)."""
    if up_sample:
        x = F.conv_transpose2d(x, sd["up_sample_conv.weight"], sd["up_sample_conv.bias"], stride=2, padding=1)
    if skip is not None:
        x = torch.cat([x, skip], dim=1)
    for i in range(layers):
        x = _resnet(sd, i, x, t_emb, groups, relu)
        if attn:
            x = _self_attn(sd, i, x, groups, heads)
        if ctx is not None:
            x = _cross_attn(sd, i, x, ctx, groups, heads)
    return x


def unet_forward(state_dict, cfg, x, t, cond_input=None, relu=False):
    """Unet.forward — models/unet_cond_base.py:124-183 (conditioned) / models/unet_base.py:68-100 (no condition_config).
    x (B,C,H,W) fp32, t (B,) or (1,) int64, cond_input {'text': (B,L,D), 'image': (B,Cm,Hm,Wm)}."""
    sd = _SD(state_dict)
    groups, heads = cfg["norm_channels"], cfg["num_heads"]
    cc = cfg.get("condition_config") or {}
    kinds = cc.get("condition_types", [])
    x = x.float()
    if "image" in kinds:
        m = F.interpolate(cond_input["image"].float(), size=x.shape[-2:])
        m = F.conv2d(m, sd["cond_conv_in.weight"])
        out = F.conv2d(torch.cat([x, m], dim=1), sd["conv_in_concat.weight"], sd["conv_in_concat.bias"], padding=1)
    else:
        out = F.conv2d(x, sd["conv_in.weight"], sd["conv_in.bias"], padding=1)
    t_emb = time_embedding(torch.as_tensor(t).long(), cfg["time_emb_dim"])
    t_emb = F.linear(_act(F.linear(t_emb, sd["t_proj.0.weight"], sd["t_proj.0.bias"]), relu), sd["t_proj.2.weight"],
                     sd["t_proj.2.bias"])
    if "class" in kinds:
        t_emb = t_emb + cond_input["class"].float() @ sd["class_emb.weight"]
    ctx = cond_input["text"].float() if "text" in kinds else None
    nd = len(cfg["down_channels"]) - 1
    skips = []
    for i in range(nd):
        skips.append(out)
        out = down_block(sd.sub(f"downs.{i}"), out, t_emb, ctx, cfg["num_down_layers"], groups, heads,
                         cfg["attn_down"][i], cfg["down_sample"][i], relu)
    for i in range(len(cfg["mid_channels"]) - 1):
        out = mid_block(sd.sub(f"mids.{i}"), out, t_emb, ctx, cfg["num_mid_layers"], groups, heads, relu)
    ups = list(reversed(cfg["down_sample"]))
    for i in range(nd):
        out = up_block(sd.sub(f"ups.{i}"), out, skips.pop(), t_emb, ctx, cfg["num_up_layers"], groups, heads, ups[i],
                       relu)
    out = _act(F.group_norm(out, groups, sd["norm_out.weight"], sd["norm_out.bias"], 1e-5), relu)
    return F.conv2d(out, sd["conv_out.weight"], sd["conv_out.bias"], padding=1)


def vqvae_decode(state_dict, cfg, z):
    """VQVAE.decode — models/vqvae.py:141-153: post_quant 1x1 -> conv_in -> MidBlocks -> UpBlocks (ConvTranspose
    first, no skip, no time embedding) -> GroupNorm / SiLU / conv_out."""
    sd = _SD(state_dict)
    groups, heads = cfg["norm_channels"], cfg["num_heads"]
    out = F.conv2d(z.float(), sd["post_quant_conv.weight"], sd["post_quant_conv.bias"])
    out = F.conv2d(out, sd["decoder_conv_in.weight"], sd["decoder_conv_in.bias"], padding=1)
    for i in range(len(cfg["mid_channels"]) - 1):
        out = mid_block(sd.sub(f"decoder_mids.{i}"), out, None, None, cfg["num_mid_layers"], groups, heads, False)
    nd = len(cfg["down_channels"]) - 1
    for j, i in enumerate(reversed(range(1, nd + 1))):
        out = up_block(sd.sub(f"decoder_layers.{j}"), out, None, None, None, cfg["num_up_layers"], groups, heads,
                       cfg["down_sample"][i - 1], False, attn=cfg["attn_down"][i - 1])
    out = F.silu(F.group_norm(out, groups, sd["decoder_norm_out.weight"], sd["decoder_norm_out.bias"], 1e-5))
    return F.conv2d(out, sd["decoder_conv_out.weight"], sd["decoder_conv_out.bias"], padding=1)


def vqvae_encode_latent(state_dict, cfg, x):
    """VQVAE.encode up to the quantiser — models/vqvae.py:130-138: conv_in -> DownBlocks (no time embedding) ->
    MidBlocks -> GroupNorm / SiLU / conv_out -> pre_quant 1x1."""
    sd = _SD(state_dict)
    groups, heads = cfg["norm_channels"], cfg["num_heads"]
    out = F.conv2d(x.float(), sd["encoder_conv_in.weight"], sd["encoder_conv_in.bias"], padding=1)
    for i in range(len(cfg["down_channels"]) - 1):
        out = down_block(sd.sub(f"encoder_layers.{i}"), out, None, None, cfg["num_down_layers"], groups, heads,
                         cfg["attn_down"][i], cfg["down_sample"][i], False)
    for i in range(len(cfg["mid_channels"]) - 1):
        out = mid_block(sd.sub(f"encoder_mids.{i}"), out, None, None, cfg["num_mid_layers"], groups, heads, False)
    out = F.silu(F.group_norm(out, groups, sd["encoder_norm_out.weight"], sd["encoder_norm_out.bias"], 1e-5))
    out = F.conv2d(out, sd["encoder_conv_out.weight"], sd["encoder_conv_out.bias"], padding=1)
    return F.conv2d(out, sd["pre_quant_conv.weight"], sd["pre_quant_conv.bias"])


def vq_quantize(x, codebook):
    """VQVAE.quantize — models/vqvae.py:93-128, with the distances as float64 sums of squared differences (the exact
    argmin; the reference's cdist picks the same code except between numerically tied ones).
    -> (quant_out NCHW fp32 as `x + (e - x)` evaluates, mse, indices (B, H, W), squared distances (B*H*W, K) float64)."""
    B, C, H, W = x.shape
    rows = x.float().permute(0, 2, 3, 1).reshape(-1, C)
    d2 = ((rows.double()[:, None, :] - codebook.double()[None, :, :]) ** 2).sum(-1)
    idx = torch.argmin(d2, dim=-1)
    e = codebook.float()[idx]
    mse = torch.mean((e - rows) ** 2)
    q = rows + (e - rows)
    return q.reshape(B, H, W, C).permute(0, 3, 1, 2).contiguous(), mse, idx.reshape(B, H, W), d2
