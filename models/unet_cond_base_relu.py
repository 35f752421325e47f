"""Drop-in for the reference's models/unet_cond_base_relu.py (ReLU / CustomMultiheadAttention conditioned UNet)."""
import torch.nn as nn

from models._unet import *  # noqa: F401,F403
from models._unet import UnetCore
from models.blocks_relu import DownBlock, MidBlock, UpBlockUnet, get_time_embedding  # noqa: F401


class Unet(UnetCore):
    CONDITIONED = True
    ACT = nn.ReLU

    def _blocks(self):
        return DownBlock, MidBlock, UpBlockUnet
