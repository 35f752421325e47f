"""Drop-in for the reference's models/unet_base.py: the unconditional LDM UNet `Unet(im_channels, model_config)
.forward(x, t)` used by tools/train_ddpm_vqvae.py and tools/sample_ddpm_vqvae.py, on the fused sm_100a engine."""
from models._unet import UnetCore
from models.blocks import DownBlock, MidBlock, UpBlockUnet, get_time_embedding  # noqa: F401


class Unet(UnetCore):
    CONDITIONED = False

    def _blocks(self):
        return DownBlock, MidBlock, UpBlockUnet

    def forward(self, x, t):
        return super().forward(x, t, None)
