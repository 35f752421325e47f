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


def enhance_branch_conv(x, conv_w, conv_b, stride, padding, w1, b1, g1, w2, b2, g2, blend, relu=False, sigmoid=True):
    """layers_enhance.py:439-535 for ANY original conv: both 1x1 branch convs carry the ORIGINAL's padding (:467,:473), so
    the branch output is (H + 4p) x (W + 4p); a strided original average-pools it by the stride (:484-487,:516-517); the
    result is cropped from the top-left corner (:527-528) or zero-padded around (:530-531) to the original's output size."""
    F = torch.nn.functional
    orig = F.conv2d(x, conv_w, conv_b, stride=stride, padding=padding)
    enh = F.conv2d(x, w1, b1, padding=padding, groups=g1)
    if relu:
        enh = torch.relu(enh)
    enh = F.conv2d(enh, w2, b2, padding=padding, groups=g2)
    if stride != 1:
        enh = F.avg_pool2d(enh, kernel_size=stride, stride=stride, padding=0)
    dy, dx = orig.shape[2] - enh.shape[2], orig.shape[3] - enh.shape[3]
    if dy != 0 or dx != 0:
        if dy < 0 or dx < 0:
            enh = enh[:, :, :orig.shape[2], :orig.shape[3]]
        elif dy > 0 or dx > 0:
            enh = F.pad(enh, (dx // 2, dx - dx // 2, dy // 2, dy - dy // 2))
    s = torch.sigmoid(blend) if sigmoid else blend
    return orig * (1 - s) + enh * s
