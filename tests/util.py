import os

import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def relerr(a, b):
    """max |a-b| / max |b|: the 'max relative error' of the parity gate (per-element ratios diverge at zeros)."""
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-30)).item()


def l2err(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-30)).item()


def load_golden(name):
    return torch.load(os.path.join(GOLD, name + ".pt"), weights_only=False)


def expand_mask(classes, channels):
    oh = torch.nn.functional.one_hot(classes.long(), channels + 1)[..., 1:]
    return oh.permute(0, 3, 1, 2).float().contiguous()


def cond_of(inp, cfg, device=None):
    cc = cfg.get('condition_config')
    if not cc:
        return None
    c = {}
    if "text" in inp:
        c["text"] = inp["text"]
    if "mask_classes" in inp:
        c["image"] = expand_mask(inp["mask_classes"], cc['image_condition_config']['image_condition_input_channels'])
    if device is not None:
        c = {k: v.to(device) for k, v in c.items()}
    return c


def randomise_zero_init(model, seed):
    """Same draw as oracle/gen_golden.py: the DiT's zero-initialised adaLN / proj_out parameters from N(0, 0.02)."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if "adaptive_norm_layer" in name or name.startswith("proj_out"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)
