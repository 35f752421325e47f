"""Drop-in for the reference's models/vqvae.py: `VQVAE(im_channels, model_config)` with the same sub-module names and
parameter creation order (state_dict / seeded init interchangeable).

`decode(z)` is the VQVAE call on the DDPM sampling path (tools/sample_ddpm_text_image_cond.py:94-98; reference
models/vqvae.py:141-153); `encode(x)` / `quantize(x)` / `forward(x)` are the inference calls that produce the latents
the LDM trainers consume (gen_vqvae_latents_multi_GPU.py:42-188; reference :93-139,155-158). All of them
reuse the implicit-GEMM convolution, GroupNorm and attention kernels of the UNet; the codebook search is
`b200sd_vq_nearest`. Stage-1 autoencoder TRAINING (gradients through encode / quantize / decode) is outside this
package's scope (SURVEY §2, row "Autoencoders (train side)"): calling these with autograd recording raises.
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

        # ---- encoder
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

    def _inference_only(self, what, x):
        if not x.is_cuda:
            raise rt.B200Error(f"VQVAE.{what}: CUDA tensors required (no CPU implementation in this package)")
        if torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in self.parameters())):
            raise rt.B200Error(f"VQVAE.{what}: no backward path (stage-1 autoencoder training is out of scope); "
                               "call it under torch.no_grad()")

    def _quantize_rows(self, rows, n, h, w):
        c = self.z_channels
        zq, idx, mse = rt.vq_nearest(rows, n, h * w, c, self.embedding.weight.detach().float().contiguous())
        # codebook_loss and commitment_loss differ only in which side is detached (:113-114): same value
        losses = {'codebook_loss': mse, 'commitment_loss': mse.clone()}
        return zq.view(n, c, h, w), losses, idx.view(n, h, w)

    def quantize(self, x):
        r"""(B, z_channels, H, W) -> (nearest codebook vectors (B, C, H, W), {'codebook_loss', 'commitment_loss'},
        indices (B, H, W) int64) — reference :93-128."""
        self._inference_only("quantize", x)
        n, c, h, w = x.shape
        rows = x.detach().float().permute(0, 2, 3, 1).reshape(n * h * w, c).contiguous()
        return self._quantize_rows(rows, n, h, w)

    def _encode_rows(self, x):
        a = E.Act.from_nchw(x)
        a = E.conv2d(a, self.encoder_conv_in)
        for down in self.encoder_layers:
            a = down.run(a)
        for mid in self.encoder_mids:
            a = mid.run(a)
        a = E.group_norm(a, self.encoder_norm_out, act="silu")
        a = E.conv2d(a, self.encoder_conv_out)
        return E.conv2d(a, self.pre_quant_conv, out_dtype=torch.float32)

    def encode_latent(self, x):
        r"""The encoder output before quantisation, fp32 (B, z_channels, H/8, W/8) (not a reference method: what
        `pre_quant_conv` returns inside reference :130-138; used by the parity tests)."""
        self._inference_only("encode_latent", x)
        return self._encode_rows(x).to_nchw()

    def encode(self, x):
        r"""(B, im_channels, H, W) image -> (quantised latent (B, z_channels, H/8, W/8), losses) — reference :130-139."""
        self._inference_only("encode", x)
        a = self._encode_rows(x)
        out, losses, _ = self._quantize_rows(a.t, a.n, a.h, a.w)
        return out, losses

    def decode(self, z):
        r"""(B, z_channels, h, w) latent -> (B, im_channels, 8h, 8w) image, fp32 NCHW — reference :141-153."""
        if not z.is_cuda:
            raise rt.B200Error("VQVAE.decode: CUDA tensors required (no CPU implementation in this package)")
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.post_quant_conv.parameters()) \
                and z.requires_grad:
            raise rt.B200Error("VQVAE.decode: no backward path (decoder training is out of scope)")
        if E.precise():
            from sdb200 import precise as P
            a = P.PAct.from_nchw(z)
        else:
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
        z, quant_losses = self.encode(x)
        out = self.decode(z)
        return out, z, quant_losses
