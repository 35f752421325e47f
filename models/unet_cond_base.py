"""Drop-in for the reference's models/unet_cond_base.py: `Unet(im_channels, model_config).forward(x, t, cond_input)`
(text / mask-image / class conditioned LDM UNet) on the fused sm_100a engine."""
from models._unet import *  # noqa: F401,F403  (the reference star-exports its config helpers from this module)
from models._unet import UnetCore
from models.blocks import DownBlock, MidBlock, UpBlockUnet, get_time_embedding  # noqa: F401


class Unet(UnetCore):
    CONDITIONED = True

    def _blocks(self):
        return DownBlock, MidBlock, UpBlockUnet
