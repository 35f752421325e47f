"""Drop-in for the reference's models/vqvae.py: `VQVAE(im_channels, model_config)` with the same sub-module names and
parameter creation order (state_dict / seeded init interchangeable). Only `decode(z)` — the one VQVAE call on the
DDPM sampling path (tools/sample_ddpm_text_image_cond.py:94-98; reference models/vqvae.py:141-153) — has a CUDA
path here; it reuses the implicit-GEMM convolution, GroupNorm and attention kernels of the UNet.

Stage-1 autoencoder training (`encode`, `quantize`, `forward`) is outside this package's scope (SURVEY §2, row
"Autoencoders (train side)" and §8 f-4): those methods raise instead of silently running something else.
"""
import torch
import torch.nn as nn

import sdb200  # noqa: F401
from sdb200 import engine as E
from sdb200 import runtime as rt
from models.blocks import DownBlock, MidBlock, UpBlock


class VQVAE(nn.Module):
    def __init__(self, im_channels, model_config):
        super().__init__()
        cfg = model_config
        self.down_channels = cfg['down_channels']
        self.mid_channels = cfg['mid_channels']
        self.down_sample = cfg['down_sample']
        self.num_down_layers = cfg['num_down_layers']
        self.num_mid_layers = cfg['num_mid_layers']
        self.num_up_layers = cfg['num_up_layers']
        self.attns = cfg['attn_down']
        self.z_channels = cfg['z_channels']
        self.codebook_size = cfg['codebook_size']
        self.norm_channels = cfg['norm_channels']
        self.num_heads = cfg['num_heads']

        assert self.mid_channels[0] == self.down_channels[-1]
        assert self.mid_channels[-1] == self.down_channels[-1]
        assert len(self.down_sample) == len(self.down_channels) - 1
        assert len(self.attns) == len(self.down_channels) - 1
        self.up_sample = list(reversed(self.down_sample))

        # ---- encoder (constructed for checkpoint compatibility; no CUDA path in this package)
        self.encoder_conv_in = nn.Conv2d(im_channels, self.down_channels[0], kernel_size=3, padding=(1, 1))
        self.encoder_layers = nn.ModuleList([
            DownBlock(self.down_channels[i], self.down_channels[i + 1], t_emb_dim=None,
                      down_sample=self.down_sample[i], num_heads=self.num_heads, num_layers=self.num_down_layers,
                      attn=self.attns[i], norm_channels=self.norm_channels)
            for i in range(len(self.down_channels) - 1)])
        self.encoder_mids = nn.ModuleList([
            MidBlock(self.mid_channels[i], self.mid_channels[i + 1], t_emb_dim=None, num_heads=self.num_heads,
                     num_layers=self.num_mid_layers, norm_channels=self.norm_channels)
            for i in range(len(self.mid_channels) - 1)])
        self.encoder_norm_out = nn.GroupNorm(self.norm_channels, self.down_channels[-1])
        self.encoder_conv_out = nn.Conv2d(self.down_channels[-1], self.z_channels, kernel_size=3, padding=1)
        self.pre_quant_conv = nn.Conv2d(self.z_channels, self.z_channels, kernel_size=1)
        self.embedding = nn.Embedding(self.codebook_size, self.z_channels)

        # ---- decoder
        self.post_quant_conv = nn.Conv2d(self.z_channels, self.z_channels, kernel_size=1)
        self.decoder_conv_in = nn.Conv2d(self.z_channels, self.mid_channels[-1], kernel_size=3, padding=(1, 1))
        self.decoder_mids = nn.ModuleList([
            MidBlock(self.mid_channels[i], self.mid_channels[i - 1], t_emb_dim=None, num_heads=self.num_heads,
                     num_layers=self.num_mid_layers, norm_channels=self.norm_channels)
            for i in reversed(range(1, len(self.mid_channels)))])
        self.decoder_layers = nn.ModuleList([
            UpBlock(self.down_channels[i], self.down_channels[i - 1], t_emb_dim=None,
                    up_sample=self.down_sample[i - 1], num_heads=self.num_heads, num_layers=self.num_up_layers,
                    attn=self.attns[i - 1], norm_channels=self.norm_channels)
            for i in reversed(range(1, len(self.down_channels)))])
        self.decoder_norm_out = nn.GroupNorm(self.norm_channels, self.down_channels[0])
        self.decoder_conv_out = nn.Conv2d(self.down_channels[0], im_channels, kernel_size=3, padding=1)

    def quantize(self, x):
        raise rt.B200Error("VQVAE.quantize: stage-1 autoencoder path is out of scope of this package (SURVEY §8 f-4)")

    def encode(self, x):
        raise rt.B200Error("VQVAE.encode: stage-1 autoencoder path is out of scope of this package (SURVEY §8 f-4)")

    def decode(self, z):
        r"""(B, z_channels, h, w) latent -> (B, im_channels, 8h, 8w) image, fp32 NCHW — reference :141-153."""
        if not z.is_cuda:
            raise rt.B200Error("VQVAE.decode: CUDA tensors required (no CPU implementation in this package)")
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.post_quant_conv.parameters()) \
                and z.requires_grad:
            raise rt.B200Error("VQVAE.decode: no backward path (decoder training is out of scope)")
        a = E.Act.from_nchw(z)
        a = E.conv2d(a, self.post_quant_conv)
        a = E.conv2d(a, self.decoder_conv_in)
        for mid in self.decoder_mids:
            a = mid.run(a)
        for up in self.decoder_layers:
            a = up.run(a)
        a = E.group_norm(a, self.decoder_norm_out, act="silu")
        a = E.conv2d(a, self.decoder_conv_out, out_dtype=torch.float32)
        return a.to_nchw()

    def forward(self, x):
        raise rt.B200Error("VQVAE.forward: stage-1 autoencoder path is out of scope of this package (SURVEY §8 f-4)")
