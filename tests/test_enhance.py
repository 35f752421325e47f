"""Enhance wrappers (Linear family): CPU checks — the group / rank search against the reference's recorded
results, the oracle against the reference fixture, and the drop-in modules' structure (state_dict keys and shapes)."""
import json
import os

import torch

from util import GOLD, load_golden


def test_group_rank_search_matches_reference_table():
    import cim_qn_train.layers_enhance as en
    table = json.load(open(os.path.join(GOLD, "enhance_search_table.json")))
    fn = {"lora_linear": en.analyze_lora_linear_ops, "enhance_linear": en.analyze_enhance_linear_ops,
          "lora_conv": lambda i, o, k, f: en.analyze_lora_conv_ops(i, o, k, k, f),
          "enhance_conv": lambda i, o, k, f: en.analyze_enhance_conv_ops(i, o, k, k, f)}
    assert len(table) > 400
    for kind, *args, want in table:
        assert fn[kind](*args) == want, (kind, args)


def _oracle_forward(c, ps, x):
    from oracle import enhance_torch as OE
    if c["kind"] == "branch":
        return OE.enhance_branch_linear(x, ps["original_linear.weight"], ps["original_linear.bias"],
                                        ps["enhance_branch_1.weight"], ps["enhance_branch_1.bias"],
                                        ps["enhance_branch_2.weight"], ps["enhance_branch_2.bias"], ps["blend_factor"],
                                        c["relu"], c["sigmoid"])[0]
    return OE.enhance_layer_linear(x, ps["original_linear.weight"], ps["original_linear.bias"], ps["fc1.weight"],
                                   ps["fc1.bias"])


def test_oracle_matches_reference_fixture():
    for c in load_golden("enhance_linear")["cases"]:
        ps = {k: v.clone().requires_grad_(True) for k, v in c["state"].items()}
        x = c["x"].clone().requires_grad_(True)
        y = _oracle_forward(c, ps, x)
        assert torch.equal(y, c["y"])
        (y * c["r"]).sum().backward()
        for got, want in [(x.grad, c["dx"])] + [(ps[k].grad, g) for k, g in c["grads"].items()]:
            assert (got - want).abs().max() <= 1e-5 * want.abs().max()


def test_drop_in_modules_have_the_reference_structure():
    import cim_qn_train.layers_enhance as en
    for c in load_golden("enhance_linear")["cases"]:
        lin = torch.nn.Linear(c["in"], c["out"])
        mod = en.EnhanceBranchLinear_LoR(lin, c["ops_factor"], relu=c["relu"], sigmoid=c["sigmoid"]) \
            if c["kind"] == "branch" else en.EnhanceLayerLinear(lin, c["ops_factor"])
        sd = mod.state_dict()
        assert list(sd.keys()) == list(c["state"].keys())
        assert all(sd[k].shape == v.shape for k, v in c["state"].items())
        mod.load_state_dict(c["state"], strict=True)
        if c["kind"] == "branch":
            assert mod.enhance_branch_1.layer_flag == "enhance_branch" and mod.enhance_branch_2.distance == 1
    # initialisation: truncated identities, zero bias, blend at sigmoid^-1(1e-3)
    gl = en.GroupedLinear(12, 8, 2)
    assert torch.equal(gl.weight[1], torch.eye(4, 6)) and float(gl.bias.abs().sum()) == 0.0
    fresh = en.EnhanceBranchLinear_LoR(torch.nn.Linear(96, 64))
    assert abs(float(fresh.blend_factor) + 6.9068) < 1e-6


def _oracle_conv_forward(c, ps, x, groups):
    from oracle import enhance_torch as OE
    if c["kind"] == "branch":
        return OE.enhance_branch_conv1x1(x, ps["original_conv.weight"], ps["original_conv.bias"],
                                         ps["enhance_branch_1.weight"], ps["enhance_branch_1.bias"], groups[0],
                                         ps["enhance_branch_2.weight"], ps["enhance_branch_2.bias"], groups[1],
                                         ps["blend_factor"], c["relu"], c["sigmoid"])
    return OE.enhance_layer_conv(x, ps["original_conv.weight"], ps["original_conv.bias"], 1, c["pad"],
                                 ps["conv_enhance.weight"], ps["conv_enhance.bias"], groups[0])


def _build_conv(c):
    import cim_qn_train.layers_enhance as en
    conv = torch.nn.Conv2d(c["cin"], c["cout"], c["k"], padding=c["pad"])
    mod = en.EnhanceBranchConv2d_LoR(conv, c["ops_factor"], relu=c["relu"], sigmoid=c["sigmoid"]) \
        if c["kind"] == "branch" else en.EnhanceLayerConv2d(conv, c["ops_factor"])
    groups = (mod.enhance_branch_1.groups, mod.enhance_branch_2.groups) if c["kind"] == "branch" else (mod.groups,)
    return mod, groups


def test_conv_wrappers_structure_and_oracle():
    import cim_qn_train.layers_enhance as en
    import pytest
    for c in load_golden("enhance_linear")["conv_cases"]:
        mod, groups = _build_conv(c)
        sd = mod.state_dict()
        assert list(sd.keys()) == list(c["state"].keys()) and all(sd[k].shape == v.shape for k, v in c["state"].items())
        mod.load_state_dict(c["state"], strict=True)
        ps = {k: v.clone().requires_grad_(True) for k, v in c["state"].items()}
        x = c["x"].clone().requires_grad_(True)
        y = _oracle_conv_forward(c, ps, x, groups)
        assert torch.equal(y, c["y"])
        (y * c["r"]).sum().backward()
        for got, want in [(x.grad, c["dx"])] + [(ps[k].grad, g) for k, g in c["grads"].items()]:
            assert (got - want).abs().max() <= 1e-5 * want.abs().max()
    fresh = en.EnhanceLayerConv2d(torch.nn.Conv2d(8, 16, 3, padding=1), 0.05)   # identity initialisation (:399-405)
    g = fresh.groups
    assert torch.equal(fresh.conv_enhance.weight.view(16, 16 // g), torch.eye(16 // g).repeat(g, 1))
    kxk = en.EnhanceBranchConv2d_LoR(torch.nn.Conv2d(8, 8, 3, padding=1))      # k x k originals: same search as the reference
    assert kxk.enhance_branch_1.padding == (1, 1) and kxk.pool is None
    assert en.EnhanceBranchConv2d_LoR(torch.nn.Conv2d(8, 8, 4, stride=2, padding=1)).pool is not None


def test_kxk_branch_conv_oracle_vs_reference_golden():
    """oracle/enhance_torch.py enhance_branch_conv against the real module's outputs and gradients for padded / strided
    originals (the frame / pool / top-left crop of layers_enhance.py:509-531), and the product class's structure."""
    import cim_qn_train.layers_enhance as en
    from oracle import enhance_torch as OE
    for c in load_golden("enhance_conv_kxk")["conv_cases"]:
        conv = torch.nn.Conv2d(c["cin"], c["cout"], c["k"], stride=c["stride"], padding=c["pad"])
        mod = en.EnhanceBranchConv2d_LoR(conv, c["ops_factor"], relu=c["relu"], sigmoid=c["sigmoid"])
        assert (mod.enhance_branch_1.groups, mod.enhance_branch_2.groups) == tuple(c["groups"])
        sd = mod.state_dict()
        assert list(sd.keys()) == list(c["state"].keys()) and all(sd[k].shape == v.shape for k, v in c["state"].items())
        ps = {k: v.clone().requires_grad_(True) for k, v in c["state"].items()}
        x = c["x"].clone().requires_grad_(True)
        y = OE.enhance_branch_conv(x, ps["original_conv.weight"], ps["original_conv.bias"], c["stride"], c["pad"],
                                   ps["enhance_branch_1.weight"], ps["enhance_branch_1.bias"], c["groups"][0],
                                   ps["enhance_branch_2.weight"], ps["enhance_branch_2.bias"], c["groups"][1],
                                   ps["blend_factor"], c["relu"], c["sigmoid"])
        assert torch.equal(y, c["y"])
        (y * c["r"]).sum().backward()
        for got, want in [(x.grad, c["dx"])] + [(ps[k].grad, g) for k, g in c["grads"].items()]:
            assert (got - want).abs().max() <= 1e-5 * want.abs().max()


def test_block_diagonal_dense_form_and_its_gradient():
    """GroupedLinear.dense_weight / _conv1x1_dense (one cached-index scatter) equal torch.block_diag padded to 8 columns,
    and the gradient that flows back to the (groups, out_g, in_g) parameter is the block-diagonal part of d(dense)."""
    import cim_qn_train.layers_enhance as en
    g = torch.Generator().manual_seed(2)
    for fin, fout, groups in ((12, 20, 4), (96, 11 * 3, 3), (8, 8, 1), (288, 288, 96)):
        gl = en.GroupedLinear(fin, fout, groups)
        with torch.no_grad():
            gl.weight.copy_(torch.randn(gl.weight.shape, generator=g))
        dense = gl.dense_weight()
        want = torch.block_diag(*gl.weight.detach().unbind(0))
        kp = (fin + 7) // 8 * 8
        assert dense.shape == (fout, kp) and torch.equal(dense[:, :fin], want) and float(dense[:, fin:].abs().sum()) == 0
        up = torch.randn(dense.shape, generator=g)
        (dense * up).sum().backward()
        blocks = torch.stack([up[k * gl.out_group:(k + 1) * gl.out_group, k * gl.in_group:(k + 1) * gl.in_group]
                              for k in range(groups)])
        assert torch.equal(gl.weight.grad, blocks)
    conv = torch.nn.Conv2d(16, 24, 1, groups=4)
    assert torch.equal(en._conv1x1_dense(conv), torch.block_diag(*conv.weight.detach().reshape(4, 6, 4).unbind(0)))
