"""Drop-in for the reference's models/blocks_relu.py: the block family with ReLU activations and
CustomMultiheadAttention (separate q/k/v/out nn.Linear leaves) — the variant the CIM quantisation pipeline trains,
because every projection is then a plain nn.Linear its module surgery can replace."""
import torch.nn as nn

from models import blocks as _b
from models.blocks import get_time_embedding  # noqa: F401
from models.multihead_attention import CustomMultiheadAttention


class _Relu:
    ACT = nn.ReLU
    MHA = CustomMultiheadAttention


class DownBlock(_Relu, _b.DownBlock):
    pass


class MidBlock(_Relu, _b.MidBlock):
    pass


class UpBlock(_Relu, _b.UpBlock):
    pass


class UpBlockUnet(_Relu, _b.UpBlockUnet):
    pass
