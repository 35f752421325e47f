"""Drop-in for the reference's models/patch_embed.py: `PatchEmbedding` (fold p x p patches, Linear, add the 2-D
sinusoidal position embedding) and `get_patch_position_embedding`. The fold + Linear + position add run as ONE
implicit GEMM over strided views of the NHWC input (stablediffusion-pytorch_b200/ops.py: patch_embed)."""
import torch
import torch.nn as nn

import sdb200  # noqa: F401
from sdb200 import engine as E
from sdb200 import runtime as rt


def _sincos(pos, quarter, device):
    freq = 10000 ** (torch.arange(0, quarter, dtype=torch.float32, device=device) / quarter)
    ang = pos[:, None].repeat(1, quarter) / freq
    return torch.cat([torch.sin(ang), torch.cos(ang)], dim=-1)


def get_patch_position_embedding(pos_emb_dim, grid_size, device):
    """(grid_h*grid_w, pos_emb_dim) fp32: [sin|cos of the row index, sin|cos of the column index], each half
    pos_emb_dim/2 wide with frequencies 10000^(i / (pos_emb_dim/4)) — reference models/patch_embed.py:5-37."""
    assert pos_emb_dim % 4 == 0, 'Position embedding dimension must be divisible by 4'
    gh, gw = grid_size
    rows = torch.arange(gh, dtype=torch.float32, device=device).repeat_interleave(gw)
    cols = torch.arange(gw, dtype=torch.float32, device=device).repeat(gh)
    q = pos_emb_dim // 4
    return torch.cat([_sincos(rows, q, device), _sincos(cols, q, device)], dim=-1)


class PatchEmbedding(nn.Module):
    def __init__(self, image_height, image_width, im_channels, patch_height, patch_width, hidden_size):
        super().__init__()
        self.image_height = image_height
        self.image_width = image_width
        self.im_channels = im_channels
        self.hidden_size = hidden_size
        self.patch_height = patch_height
        self.patch_width = patch_width
        self.patch_embed = nn.Sequential(nn.Linear(im_channels * patch_height * patch_width, hidden_size))
        nn.init.xavier_uniform_(self.patch_embed[0].weight)
        nn.init.constant_(self.patch_embed[0].bias, 0)

    def position_rows(self, gh, gw, device):
        """bf16 [gh*gw, hidden] position embedding (what `pos_embed.to(out.dtype)` is under the trainers' bf16
        autocast, patch_embed.py:95), built once per grid."""
        cache = self.__dict__.setdefault("_pos", {})
        key = (gh, gw, str(device))
        if key not in cache:
            cache[key] = get_patch_position_embedding(self.hidden_size, (gh, gw), device).to(torch.bfloat16).contiguous()
        return cache[key]

    def run(self, xbuf, n, H, W, cin):
        """xbuf: NHWC bf16 [n*H*W, ld] holding `cin` channels -> token rows bf16 [n*gh*gw, hidden]."""
        if self.patch_height != self.patch_width:
            raise rt.B200Error("PatchEmbedding: only square patches have a CUDA path")
        assert H % self.patch_height == 0, "Input height must be divisible by patch height"
        assert W % self.patch_width == 0, "Input width must be divisible by patch width"
        p = self.patch_height
        pos = self.position_rows(H // p, W // p, xbuf.device)
        return E.patch_embed(xbuf, n, H, W, cin, p, self.patch_embed[0], pos)

    def forward(self, x):
        if not x.is_cuda:
            raise rt.B200Error("PatchEmbedding.forward: CUDA tensors required (no CPU implementation in this package)")
        n, c, H, W = x.shape
        with torch.autocast("cuda", enabled=False):
            a = E.Act.from_nchw(x)
            out = self.run(a.t, n, H, W, c)
        return out.reshape(n, -1, self.hidden_size).float()
