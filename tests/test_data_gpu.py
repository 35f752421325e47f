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


def test_shrunk_mask_upload_is_bit_identical_to_the_full_resolution_path():
    """shrink={"image": (32, 32)}: the prefetcher uploads only the pixels F.interpolate(mode="nearest") selects
    (unet_cond_base.py:134-135); the model output must not change by a bit, over several batches (the two pinned
    staging buffers rotate)."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from util import cond_of, load_golden
    from models.unet_cond_base import Unet
    dev = torch.device("cuda:0")
    fx = load_golden("unet_small_cond")
    cfg, inp = fx["config"], fx["inputs"]
    torch.manual_seed(fx["seed"])
    model = Unet(4, cfg).eval().to(dev)
    base = cond_of(inp, cfg)
    host = []
    for i in range(4):
        mask = torch.roll(base["image"], shifts=7 * i, dims=3).contiguous()
        host.append((inp["x"] + 0.1 * i, {"text": base["text"] * (1 + i), "image": mask.pin_memory()}))
    h, w = inp["x"].shape[-2:]
    small = list(DATA.DevicePrefetcher(host, dev, shrink={"image": (h, w)}))
    full = list(DATA.DevicePrefetcher(host, dev))
    ratio = host[0][1]["image"].shape[2] // h
    assert ratio > 1
    t = inp["t"].to(dev)
    for (xs, cs), (xf, cf), (_, hc) in zip(small, full, host):
        assert cs["image"].shape[-2:] == (h, w) and cf["image"].shape == hc["image"].shape
        assert torch.equal(cs["image"].cpu(), hc["image"][:, :, ::ratio, ::ratio])
        with torch.no_grad():
            a, b = model(xs, t, cs), model(xf, t, cf)
        assert torch.equal(a, b)
    # the reference resize picks the same pixels
    ref = torch.nn.functional.interpolate(host[1][1]["image"], size=(h, w))
    assert torch.equal(ref, small[1][1]["image"].cpu())
