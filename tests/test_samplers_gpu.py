"""DDPMSampler / DDIMSampler drop-ins (scheduler/linear_noise_scheduler.py in this repo) against outputs of the REAL
reference samplers (tests/golden/samplers.pt, oracle/gen_golden.py:gen_samplers): same stand-in model, same recorded
randn_like draws -> bit-identical fp32 results, including the intermediate-sample trace."""
import os

import pytest
import torch

from scheduler.linear_noise_scheduler import DDIMSampler, DDPMSampler, extract
from test_oracle import _sampler_model
from util import GOLD

pytestmark = pytest.mark.gpu


@pytest.fixture()
def fx():
    return torch.load(os.path.join(GOLD, "samplers.pt"), weights_only=False)


class _Replay:
    """Stands in for torch.randn_like: hands out the draws the reference run consumed, in order."""

    def __init__(self, zs):
        self.zs, self.k = zs, 0

    def __call__(self, x, *a, **k):
        z = self.zs[self.k].to(x.device)
        self.k += 1
        return z


@pytest.mark.parametrize("name", ["ddim_linear_eta0", "ddim_quadratic_eta05", "ddim_linear_eta1_trace"])
def test_ddim_bit_exact_vs_reference(fx, name, monkeypatch):
    c = fx[name]
    rep = _Replay(c["z"])
    monkeypatch.setattr(torch, "randn_like", rep)
    cond = {"image": fx["cond_image"].cuda()}
    smp = DDIMSampler(_sampler_model, c["beta"], c["T"])
    out = smp.forward(fx["x_T"].cuda(), cond, None, **c["kw"])
    assert rep.k == len(c["z"])          # consumed the generator exactly like the reference (one draw per step)
    assert out.shape == c["out"].shape
    assert torch.equal(out.cpu(), c["out"])
    assert smp.cond_input is cond and smp.uncond_input is None


def test_ddim_eta0_without_the_draw_gives_the_same_values(fx, monkeypatch):
    c = fx["ddim_linear_eta0"]
    monkeypatch.setattr(DDIMSampler, "draw_noise_when_eta_is_zero", False)
    smp = DDIMSampler(_sampler_model, c["beta"], c["T"])
    out = smp.forward(fx["x_T"].cuda(), {"image": fx["cond_image"].cuda()}, None, **c["kw"])
    assert torch.equal(out.cpu(), c["out"])


def test_ddpm_bit_exact_vs_reference(fx, monkeypatch):
    c = fx["ddpm_T40"]
    rep = _Replay(c["z"])
    monkeypatch.setattr(torch, "randn_like", rep)
    smp = DDPMSampler(_sampler_model, c["beta"], c["T"]).cuda()
    assert list(smp.state_dict().keys()) == c["state_keys"]
    out = smp.forward(fx["x_T"].cuda())
    assert torch.equal(out.cpu(), c["out"])
    trace = None
    rep.k = 0
    trace = smp.forward(fx["x_T"].cuda(), only_return_x_0=False, interval=10)
    assert trace.shape[1] == 1 + 4 and torch.equal(trace[:, -1].cpu(), c["out"].clamp(-1, 1))


def test_ddpm_raises_on_nan_and_bad_method():
    smp = DDPMSampler(lambda x, t: x * float("nan"), (1e-4, 0.02), 5).cuda()
    with pytest.raises(ValueError, match="nan in tensor"):
        smp.forward(torch.randn(1, 4, 8, 8, device="cuda"))
    with pytest.raises(NotImplementedError):
        DDIMSampler(_sampler_model, (1e-4, 0.02), 100).forward(torch.randn(1, 4, 8, 8, device="cuda"), None, None,
                                                                steps=5, method="cosine")
    v = torch.arange(10.0)
    assert extract(v, torch.tensor([3, 7]), (2, 4, 8, 8)).shape == (2, 1, 1, 1)


def test_ddim_drives_the_graphed_unet_with_a_guidance_wrapper():
    """The batch generator's pattern (batch_condition_image_generator.py:241-268): model = CFG wrapper around our Unet,
    cond / uncond dicts alternate every step; both must keep replaying their captured forward."""
    import importlib
    import yaml  # noqa: F401
    from models.unet_cond_base import Unet
    from util import load_golden
    fxu = load_golden("unet_small_cond")
    torch.manual_seed(fxu["seed"])
    net = Unet(4, fxu["config"]).cuda().eval()
    n = 2
    g = torch.Generator().manual_seed(5)
    cond = {"text": torch.randn(n, 77, 128, generator=g).cuda(), "image": (torch.rand(n, 18, 64, 64, generator=g) > 0.5).float().cuda()}
    uncond = {"text": torch.zeros_like(cond["text"]), "image": torch.zeros_like(cond["image"])}

    class Guided(torch.nn.Module):
        def __init__(self, base):
            super().__init__()
            self.base_model = base

        def forward(self, x_t, t, c):
            cp = self.base_model(x_t, t, c)
            up = self.base_model(x_t, t, uncond)
            return up + 2.0 * (cp - up)

    rt = importlib.import_module("stablediffusion-pytorch_b200.runtime")
    smp = DDIMSampler(Guided(net), (0.00085, 0.012), 1000)
    x = torch.randn(n, 4, 32, 32, generator=g).cuda()
    with torch.no_grad():
        out = smp.forward(x, cond, uncond, steps=6, method="linear", eta=0.0)
    assert out.shape == x.shape and torch.isfinite(out).all()
    assert len(net.__dict__["_graphed"]) == 2 and len(net.__dict__["_cond_cache"]) == 2
    assert rt.launch_count() > 0
