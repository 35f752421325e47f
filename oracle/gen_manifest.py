"""Generate tests/golden/state_dict_manifest.json from the REAL reference (build container only):
state_dict keys + shapes (in order) of every drop-in model class, constructed on the meta device.

    python oracle/gen_manifest.py
"""
import json
import os
import sys

import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from gen_golden import DIT_9L, DIT_12L  # noqa: E402


def main():
    sys.path.insert(0, REF)
    from models.unet_cond_base import Unet as RefCond
    from models.unet_base import Unet as RefUncond
    from models.unet_cond_base_relu import Unet as RefCondRelu
    from models.vqvae import VQVAE as RefVQVAE
    from models.transformer import DIT as RefDIT
    sys.path.pop(0)
    full_cond = yaml.safe_load(open(os.path.join(REF, "config/celebhq_text_image_cond.yaml")))
    full_uncond = yaml.safe_load(open(os.path.join(REF, "config/celebhq.yaml")))
    entries = {
        "unet_cond_base": (RefCond, (4, full_cond['ldm_params'])),
        "unet_base": (RefUncond, (4, full_uncond['ldm_params'])),
        "unet_cond_base_relu": (RefCondRelu, (4, full_cond['ldm_params'])),
        "vqvae": (RefVQVAE, (3, full_uncond['autoencoder_params'])),
        "dit_9l_text_image": (RefDIT, (4, DIT_9L)),
        "dit_12l_image": (RefDIT, (4, DIT_12L)),
    }
    out = {}
    for name, (cls, args) in entries.items():
        with torch.device("meta"):
            m = cls(*args)
        out[name] = {"config": args[1], "state_dict": [[k, list(v.shape)] for k, v in m.state_dict().items()]}
        print(name, len(out[name]["state_dict"]), "tensors")
    json.dump(out, open(os.path.join(ROOT, "tests", "golden", "state_dict_manifest.json"), "w"))


if __name__ == "__main__":
    main()
