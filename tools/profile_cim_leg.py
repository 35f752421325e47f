"""Kernel census of one quant-noise training step of the DiT-9L with cim_layers leaves (bench.py's cim_* legs):
torch.profiler over 3 steps (captured graphs replayed -> kernels attributed by name), top kernels by device time.

    python tools/profile_cim_leg.py [lsq|adda|enhance] [--eager]"""
import os
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("-") else "lsq"
    dev = torch.device("cuda:0")
    import torch.nn as nn
    from cim_layers.layers_qn_lsq import Conv2d_qn_lsq, Linear_qn_lsq
    from models.transformer import DIT
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    from sdb200.optim import FusedAdam
    from sdb200 import graphs as G
    if "--eager" in sys.argv:
        G.ENABLE_TRAIN_GRAPHS = False
    ucfg, _ = bench.load_cfg()
    torch.manual_seed(bench.SEED)
    m = DIT(4, bench.dit_config(9, ["text", "image"]))
    g = torch.Generator().manual_seed(3)
    with torch.no_grad():
        for name, p in m.named_parameters():
            if "adaptive_norm_layer" in name or name.startswith("proj_out"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)
    kw = dict(weight_bit=4, input_bit=8, output_bit=8, noise_scale=0.08)
    lin_cls, conv_cls = Linear_qn_lsq, Conv2d_qn_lsq
    if kind == "adda":
        from cim_layers.layers_qn_lsq_adda_cim import Conv2d_lsq_adda_cim, Linear_lsq_adda_cim, gen_weight_split_dict
        kw.update(dac_bit=5, adc_bit=8, adc_gain_1_scale=9.071428571, adc_gain_range=[1 / 64, 1 / 64], adc_adjust_mode="current")
        lin_cls, conv_cls = Linear_lsq_adda_cim, Conv2d_lsq_adda_cim
    for parent in list(m.modules()):
        for name, c in list(parent.named_children()):
            new = None
            if type(c) is nn.Linear:
                new = lin_cls(c.in_features, c.out_features, bias=c.bias is not None, **kw)
            elif type(c) is nn.Conv2d:
                new = conv_cls(c.in_channels, c.out_channels, c.kernel_size, c.stride, c.padding, c.groups, bias=c.bias is not None, **kw)
            if new is not None:
                new.weight, new.bias = c.weight, c.bias
                if kind == "adda":
                    new.weight_mapping_info = gen_weight_split_dict(new, [576, 2048])
                setattr(parent, name, new)
    if kind == "enhance":       # the wrapping bench.py's cim_lsq_enhance_train leg applies (Model_DiT_9L_train.py:541-553)
        import cim_qn_train.layers_enhance as en
        for n, c in [(n, c) for n, c in m.named_modules() if type(c) in (Linear_qn_lsq, Conv2d_qn_lsq)]:
            parent = m.get_submodule(n.rsplit(".", 1)[0]) if "." in n else m
            if type(c) is Linear_qn_lsq:
                br = en.EnhanceBranchLinear_LoR(original_linear=c, relu=False, sigmoid=True, ops_factor=0.05)
                br.original_linear = en.EnhanceLayerLinear(c, ops_factor=0.05)
            else:
                br = en.EnhanceBranchConv2d_LoR(original_conv=c, relu=False, sigmoid=True, ops_factor=0.05)
                br.original_conv = en.EnhanceLayerConv2d(c, ops_factor=0.05)
            setattr(parent, n.rsplit(".", 1)[-1], br)
    m = m.train().to(dev)
    if kind == "enhance":
        from sdb200.optim import PercentOptimizerFP
        opt = PercentOptimizerFP(m.parameters(), lr=1e-5)
    else:
        opt = FusedAdam(m.parameters(), lr=1e-5, max_grad_norm=1.0)
    sched = LinearNoiseScheduler(1000, 0.00085, 0.012)
    B = 16
    x, text, mask = (t.to(dev) for t in bench.synth_inputs(B, ucfg, pin=False))
    noise = torch.randn_like(x)
    t = torch.randint(0, 1000, (B,), device=dev)

    def step():
        noisy = sched.add_noise(x, noise, t)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = torch.nn.functional.mse_loss(m(noisy, t, {"text": text, "image": mask}), noise)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()

    for _ in range(4):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        step()
    e1.record()
    torch.cuda.synchronize()
    print(f"{kind}: {e0.elapsed_time(e1) / 5:.2f} ms / step")
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            step()
        torch.cuda.synchronize()
    agg = defaultdict(lambda: [0, 0.0])
    for ev in prof.events():
        if ev.device_type is not None and "cuda" in str(ev.device_type).lower():
            agg[ev.name][0] += 1
            agg[ev.name][1] += ev.device_time if hasattr(ev, "device_time") else ev.cuda_time
    tot = sum(v[1] for v in agg.values())
    print(f"kernels per step {sum(v[0] for v in agg.values()) / 3:.0f}, device time per step {tot / 3 / 1e3:.2f} ms")
    for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
        print(f"{us / 3 / 1e3:8.3f} ms {n / 3:7.0f} x  {us / tot * 100:5.1f}%  {name[:110]}")


if __name__ == "__main__":
    main()
