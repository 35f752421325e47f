"""TEST INFRASTRUCTURE (oracle): CPU restatement of the Linear-side Enhance wrappers of the reference
(cim_qn_train/layers_enhance.py:324-365 GroupedLinear, :413-436 EnhanceLayerLinear, :556-621 EnhanceBranchLinear_LoR) as
pure functions over tensors. Pinned against the real reference modules by oracle/gen_golden.py --only-enhance
(tests/golden/enhance_linear.pt: outputs and gradients identical). Only tests/, __graft_entry__.smoke() and bench.py's
CPU legs may import this package."""
import torch


def grouped_linear(x, weight, bias):
    """weight (G, out_g, in_g), bias (G, out_g): x (..., G*in_g) -> (..., G*out_g)   (layers_enhance.py:348-365)"""
    g, _, ig = weight.shape
    xg = x.view(*x.shape[:-1], g, ig)
    y = torch.einsum('...gi,goi->...go', xg, weight) + bias
    return y.reshape(*x.shape[:-1], -1)


def enhance_layer_linear(x, lin_w, lin_b, fc_w, fc_b):
    """layers_enhance.py:433-436"""
    return grouped_linear(torch.nn.functional.linear(x, lin_w, lin_b), fc_w, fc_b)


def enhance_branch_linear(x, lin_w, lin_b, w1, b1, w2, b2, blend, relu=False, sigmoid=True):
    """layers_enhance.py:601-621; returns (combined, original_output, enhanced_output)"""
    orig = torch.nn.functional.linear(x, lin_w, lin_b)
    enh = grouped_linear(x, w1, b1)
    if relu:
        enh = torch.relu(enh)
    enh = grouped_linear(enh, w2, b2)
    s = torch.sigmoid(blend) if sigmoid else blend
    return orig * (1 - s) + enh * s, orig, enh


def enhance_layer_conv(x, conv_w, conv_b, stride, padding, enh_w, enh_b, groups):
    """layers_enhance.py:407-410: original conv, then the grouped 1x1 conv_enhance"""
    y = torch.nn.functional.conv2d(x, conv_w, conv_b, stride=stride, padding=padding)
    return torch.nn.functional.conv2d(y, enh_w, enh_b, groups=groups)


def enhance_branch_conv1x1(x, conv_w, conv_b, w1, b1, g1, w2, b2, g2, blend, relu=False, sigmoid=True):
    """layers_enhance.py:509-535 for a 1x1 / stride-1 / unpadded original conv (pool is None and the branch output has
    the original's size, so the pad / crop block :523-531 does nothing)."""
    orig = torch.nn.functional.conv2d(x, conv_w, conv_b)
    enh = torch.nn.functional.conv2d(x, w1, b1, groups=g1)
    if relu:
        enh = torch.relu(enh)
    enh = torch.nn.functional.conv2d(enh, w2, b2, groups=g2)
    s = torch.sigmoid(blend) if sigmoid else blend
    return orig * (1 - s) + enh * s
