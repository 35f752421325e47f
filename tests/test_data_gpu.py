"""DevicePrefetcher: same batches, same order, same structure as iterating the loader and calling .to(device)."""
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu

DATA = importlib.import_module("stablediffusion-pytorch_b200.data")


def test_prefetcher_yields_every_batch_in_order_with_structure():
    g = torch.Generator().manual_seed(3)
    host = [(torch.randn(4, 4, 32, 32, generator=g).pin_memory(),
             {"text": torch.randn(4, 77, 16, generator=g).pin_memory(), "image": torch.rand(4, 3, 64, 64, generator=g).pin_memory(),
              "name": f"batch{i}"}) for i in range(5)]
    got = []
    for im, cond in DATA.DevicePrefetcher(host, "cuda:0"):
        assert im.is_cuda and cond["text"].is_cuda and cond["image"].is_cuda and isinstance(cond["name"], str)
        got.append((im * 2.0, cond))     # work on the compute stream while the next upload is in flight
    assert len(got) == 5 and len(DATA.DevicePrefetcher(host, "cuda:0")) == 5
    for (im2, cond), (him, hcond) in zip(got, host):
        assert torch.equal(im2.cpu(), him * 2.0)
        assert torch.equal(cond["text"].cpu(), hcond["text"]) and torch.equal(cond["image"].cpu(), hcond["image"])
        assert cond["name"] == hcond["name"]
    assert list(DATA.DevicePrefetcher([], "cuda:0")) == []
