"""The drop-in boundary exercised by the REAL reference caller (SURVEY §8b invariant 3).

With this repo ahead of /root/reference on sys.path, the reference's own module surgery —
`ProgressiveTrain.convert_to_layers` (cim_qn_train/progressive_qn_train.py:576-651), `add_enhance_branch_LoR` /
`add_enhance_layers` (:884-950), `convert_to_layers(... 'layers_qn_lsq_adda_cim')` and `map_weight_for_model`
(cim_weight_mapper/weight_process.py:177), looked up through `cim_layers/register_dict.py:17-110` by the `cls.__module__`
component — runs UNCHANGED on the product `DIT` and must end with every leaf being THIS repo's class. CPU only (module
construction, no compute); skipped where the reference checkout does not exist (the GPU box).
`onnx` / `matplotlib` are absent from the image: pre-seeded with MagicMock as BASELINE.md §4 describes."""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"

SCRIPT = textwrap.dedent('''
    import os, sys, io, contextlib
    from unittest.mock import MagicMock
    ROOT, REF = sys.argv[1], sys.argv[2]
    sys.path[:0] = [ROOT, REF]                      # this repo first, the reference checkout behind it
    for m in ("onnx", "onnx.shape_inference", "onnx.numpy_helper", "matplotlib", "matplotlib.pyplot"):
        sys.modules[m] = MagicMock()
    import torch, torch.nn as nn
    quiet = io.StringIO()
    with contextlib.redirect_stdout(quiet):
        from cim_layers.layers_all import *           # the reference's star import (train scripts do exactly this)
        import cim_layers.register_dict as reg_dict
        from cim_qn_train.progressive_qn_train import ProgressiveTrain
        from cim_weight_mapper.weight_process import map_weight_for_model
    import cim_layers.layers_qn_lsq as ours_lsq, cim_layers.layers_qn_lsq_adda_cim as ours_adda
    import cim_qn_train.layers_enhance as ours_en
    import cim_layers.IBA_optimizer as ours_opt
    here = lambda mod: os.path.abspath(mod.__file__).startswith(ROOT + os.sep)
    assert here(ours_lsq) and here(ours_adda) and here(ours_en) and here(ours_opt), "repo modules shadowed by the reference"
    assert not os.path.abspath(reg_dict.__file__).startswith(ROOT + os.sep), "register_dict must be the reference's own file"
    # the registry tuples (register_dict.py:17-110) hold THIS repo's classes
    assert ours_lsq.Linear_qn_lsq in reg_dict.qn_linear_layers and ours_lsq.Conv2d_qn_lsq in reg_dict.qn_conv_layers
    assert ours_adda.Linear_lsq_adda_cim in reg_dict.cim_linear_layers and ours_adda.Conv2d_lsq_adda_cim in reg_dict.cim_conv_layers
    assert ours_opt.PercentOptimizerFP.__module__.startswith(("sdb200", "stablediffusion-pytorch_b200", "cim_layers")), ours_opt.PercentOptimizerFP.__module__
    import sdb200.optim
    assert ours_opt.PercentOptimizerFP is sdb200.optim.PercentOptimizerFP, "the reference's exec'd file must not override the fused optimizer"

    from models.transformer import DIT
    cc = {"condition_types": ["text", "image"],
          "text_condition_config": {"text_embed_model": "clip", "train_text_embed_model": False, "text_embed_dim": 512, "cond_drop_prob": 0.9},
          "image_condition_config": {"image_condition_input_channels": 18, "image_condition_output_channels": 3,
                                     "image_condition_h": 512, "image_condition_w": 512, "cond_drop_prob": 0.9}}
    cfg = {"hidden_size": 288, "patch_size": 2, "timestep_emb_dim": 192, "num_layers": 9, "num_heads": 9, "head_dim": 32,
           "condition_config": cc}
    torch.manual_seed(0)
    model = DIT(4, cfg)
    n_leaves = sum(1 for m in model.modules() if type(m) in (nn.Linear, nn.Conv2d))
    with contextlib.redirect_stdout(quiet):
        trainer = ProgressiveTrain(model)
        trainer.convert_to_layers(convert_layer_type_list=reg_dict.nn_layers, tar_layer_type="layers_qn_lsq",
                                  noise_scale=0.08, input_bit=8, output_bit=8, weight_bit=4)
    leaves = [m for m in trainer.model.modules() if type(m) in (ours_lsq.Linear_qn_lsq, ours_lsq.Conv2d_qn_lsq)]
    assert len(leaves) == n_leaves and n_leaves == 96, (len(leaves), n_leaves)
    assert not any(type(m) in (nn.Linear, nn.Conv2d) for m in trainer.model.modules())
    assert all("layers_qn_lsq" in type(m).__module__.split(".") and here(sys.modules[type(m).__module__]) for m in leaves)
    with contextlib.redirect_stdout(quiet):
        trainer.add_enhance_branch_LoR(ops_factor=0.05)
        trainer.add_enhance_layers(ops_factor=0.05)
    kinds = {type(m) for m in trainer.model.modules()}
    assert ours_en.EnhanceBranchLinear_LoR in kinds and ours_en.EnhanceLayerLinear in kinds
    assert ours_en.EnhanceBranchConv2d_LoR in kinds and ours_en.EnhanceLayerConv2d in kinds
    n_branch = sum(1 for m in trainer.model.modules() if type(m) in (ours_en.EnhanceBranchLinear_LoR, ours_en.EnhanceBranchConv2d_LoR))
    n_layer = sum(1 for m in trainer.model.modules() if type(m) in (ours_en.EnhanceLayerLinear, ours_en.EnhanceLayerConv2d))
    assert n_branch == 96 and n_layer == 96, (n_branch, n_layer)
    sd_keys = len(trainer.model.state_dict())
    with contextlib.redirect_stdout(quiet):
        trainer.convert_to_layers(convert_layer_type_list=reg_dict.custom_layers, tar_layer_type="layers_qn_lsq_adda_cim",
                                  noise_scale=0.08, input_bit=8, output_bit=8, weight_bit=4, dac_bit=5, adc_bit=8,
                                  adc_gain_1_scale=9.071428571, adc_gain_range=[1 / 64, 1 / 64], adc_adjust_mode="current")
        map_weight_for_model(trainer.model, array_device_name="TC05", array_size=[576, 2048],
                             weight_block_size=[576, 2048], draw_weight_block=False)
    adda = [m for m in trainer.model.modules() if type(m) in (ours_adda.Linear_lsq_adda_cim, ours_adda.Conv2d_lsq_adda_cim)]
    assert len(adda) == 96, len(adda)
    assert all(getattr(m, "weight_mapping_info", None) for m in adda), "map_weight_for_model did not reach every ADDA leaf"
    blocks = sum(len(m.weight_mapping_info) for m in adda)
    print("OK", n_leaves, len(adda), blocks, sd_keys)
''')


@pytest.mark.skipif(not os.path.isdir(REF), reason="the reference checkout only exists in the build container")
def test_reference_progressive_train_converts_the_product_dit_to_this_repos_layers():
    r = subprocess.run([sys.executable, "-c", SCRIPT, ROOT, REF], capture_output=True, text=True, timeout=600,
                       cwd="/tmp")
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    tag, leaves, adda, blocks, _ = r.stdout.strip().splitlines()[-1].split()
    assert tag == "OK" and int(leaves) == 96 and int(adda) == 96
    assert int(blocks) == 105    # SURVEY §8c [probed]: 96 layers -> 105 blocks; only mlp_block.2 (K = 1152) splits
