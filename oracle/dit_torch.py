"""ORACLE — test infrastructure only (imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs; never by the product path).

Plain PyTorch fp32 CPU restatement of the reference's DiT forward (models/transformer.py:153-213,
models/transformer_layer.py:80-106, models/attention.py:33-78, models/patch_embed.py:5-96,
models/multihead_attention.py:41-80), written functionally over a state_dict. The floating-point arithmetic of
linear / layer_norm / softmax lives in PyTorch (un-pinned third-party dependency of the reference), so
this restatement calls the same torch.nn.functional primitives.

Pinning: oracle/gen_golden.py runs the REAL reference DIT in the build container on seeded inputs (with the
zero-initialised adaLN / proj_out parameters re-randomised — SURVEY §8c trap (i): a default-init DiT outputs exactly
0), asserts this restatement reproduces it, and commits the outputs as tests/golden/dit_*.pt; tests/test_oracle.py
re-checks this file against those fixtures on every run.
"""
import torch
import torch.nn.functional as F


def time_embedding(t, dim):
    """get_time_embedding — models/transformer.py:18-40."""
    half = dim // 2
    factor = 10000 ** (torch.arange(0, half, dtype=torch.float32) / half)
    e = t.reshape(-1)[:, None].repeat(1, half) / factor
    return torch.cat([torch.sin(e), torch.cos(e)], dim=-1)


def position_embedding(dim, gh, gw):
    """get_patch_position_embedding — models/patch_embed.py:5-37."""
    q = dim // 4
    factor = 10000 ** (torch.arange(0, q, dtype=torch.float32) / q)
    ys = torch.arange(gh, dtype=torch.float32)[:, None].expand(gh, gw).reshape(-1)
    xs = torch.arange(gw, dtype=torch.float32)[None, :].expand(gh, gw).reshape(-1)
    ey, ex = ys[:, None] / factor, xs[:, None] / factor
    return torch.cat([torch.sin(ey), torch.cos(ey), torch.sin(ex), torch.cos(ex)], dim=-1)


# Optional leaf hook: LEAF(name, x, weight, bias, conv) -> y replaces every nn.Linear / the mask nn.Conv2d — how the
# reference's module surgery (cim_qn_train/progressive_qn_train.py:576-651) swaps in cim_layers drop-ins.
LEAF = None


def _lin(sd, name, x):
    if LEAF is not None:
        return LEAF(name, x, sd[name + ".weight"], sd.get(name + ".bias"), None)
    return F.linear(x, sd[name + ".weight"], sd.get(name + ".bias"))


def _heads(x, heads):
    b, l, c = x.shape
    return x.reshape(b, l, heads, c // heads).transpose(1, 2)


def _sdpa(q, k, v, heads):
    q, k, v = _heads(q, heads), _heads(k, heads), _heads(v, heads)
    att = torch.softmax(q @ k.transpose(-2, -1) * (q.shape[-1] ** -0.5), dim=-1)
    out = att @ v
    return out.transpose(1, 2).reshape(out.shape[0], out.shape[2], -1)


def _ln(x):
    return F.layer_norm(x, x.shape[-1:], None, None, 1e-6)


def layer_forward(sd, pre, x, cond, ctx, heads):
    """TransformerLayer.forward — models/transformer_layer.py:80-106."""
    mod = _lin(sd, pre + "adaptive_norm_layer.1", F.relu(cond)).chunk(6, dim=1)
    shift_a, scale_a, gate_a, shift_m, scale_m, gate_m = (m.unsqueeze(1) for m in mod)
    h = _ln(x) * (1 + scale_a) + shift_a
    q, k, v = _lin(sd, pre + "attn_block.qkv_proj", h).chunk(3, dim=-1)
    x = x + gate_a * _lin(sd, pre + "attn_block.output_proj.0", _sdpa(q, k, v, heads))
    if ctx is not None and (pre + "context_proj.weight") in sd:
        c = _lin(sd, pre + "context_proj", ctx)
        h = _ln(x)
        o = _sdpa(_lin(sd, pre + "cross_attn_block.q_proj", h), _lin(sd, pre + "cross_attn_block.k_proj", c),
                  _lin(sd, pre + "cross_attn_block.v_proj", c), heads)
        x = x + _lin(sd, pre + "cross_attn_block.out_proj", o)
    h = _ln(x) * (1 + scale_m) + shift_m
    h = _lin(sd, pre + "mlp_block.2", F.relu(_lin(sd, pre + "mlp_block.0", h)))
    return x + gate_m * h


def dit_forward(state_dict, cfg, x, t, cond_input=None):
    """DIT.forward — models/transformer.py:153-213."""
    sd = state_dict
    p, hidden, heads = cfg['patch_size'], cfg['hidden_size'], cfg['num_heads']
    b, c, H, W = x.shape
    src = x
    if "cond_conv_in.weight" in sd:
        im = F.interpolate(cond_input['image'].to(x.dtype), size=(H, W))
        if LEAF is not None:
            imc = LEAF("cond_conv_in", im, sd["cond_conv_in.weight"], None, dict(stride=1, padding=0))
        else:
            imc = F.conv2d(im, sd["cond_conv_in.weight"])
        src = torch.cat([x, imc], dim=1)
    gh, gw = H // p, W // p
    cin = src.shape[1]
    tok = src.reshape(b, cin, gh, p, gw, p).permute(0, 2, 4, 3, 5, 1).reshape(b, gh * gw, p * p * cin)
    tok = _lin(sd, "patch_embed_layer.patch_embed.0", tok) + position_embedding(hidden, gh, gw)
    t = torch.as_tensor(t)
    emb = time_embedding(t.reshape(-1).long(), cfg['timestep_emb_dim'])
    if "class_emb.weight" in sd:
        emb = emb + cond_input['class'].float() @ sd["class_emb.weight"]
    emb = _lin(sd, "t_proj.2", F.relu(_lin(sd, "t_proj.0", emb)))
    ctx = cond_input['text'] if (cond_input is not None and 'text' in cond_input) else None
    for i in range(cfg['num_layers']):
        tok = layer_forward(sd, f"transformer_layers.{i}.", tok, emb, ctx, heads)
    shift, scale = _lin(sd, "adaptive_norm_layer.1", F.relu(emb)).chunk(2, dim=1)
    tok = _ln(tok) * (1 + scale.unsqueeze(1)) + shift.unsqueeze(1)
    tok = _lin(sd, "proj_out", tok)
    return tok.reshape(b, gh, gw, p, p, c).permute(0, 5, 1, 3, 2, 4).reshape(b, c, H, W)
