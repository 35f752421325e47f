"""Per-kernel CUDA-event timing of one training step (forward + backward) of the conditioned UNet at batch 16:
    python tools/train_profile.py [--batch 16]
Prints the per-family totals of the library's kernels and the step's wall time (host-bound vs device-bound)."""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
import sdb200  # noqa: E402,F401
from sdb200 import runtime as rt  # noqa: E402
from models.unet_cond_base import Unet  # noqa: E402
from scheduler.linear_noise_scheduler import LinearNoiseScheduler  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--ncu-steps", type=int, default=0,
                    help="run N captured steps (forward, backward, clip, Adam), each preceded by an advance_t_kernel "
                         "marker, and exit: the launch list for `ncu --metrics gpu__time_duration.sum`; "
                         "tools/summarize_launches.py --step k then cuts out one step")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    ucfg, _ = bench.load_cfg()
    torch.manual_seed(1111)
    model = Unet(4, ucfg).train().to(dev)
    x, text, mask = (v.to(dev) for v in bench.synth_inputs(a.batch, ucfg))
    noise = torch.randn_like(x)
    t = torch.randint(0, 1000, (a.batch,), device=dev)
    sched = LinearNoiseScheduler(1000, 0.00085, 0.012)

    def step():
        noisy = sched.add_noise(x, noise, t)
        pred = model(noisy, t, {"text": text, "image": mask})
        loss = torch.nn.functional.mse_loss(pred, noise)
        model.zero_grad(set_to_none=True)
        loss.backward()

    if a.ncu_steps:
        opt = torch.optim.Adam(model.parameters(), lr=1e-5)
        mark = torch.zeros(1, device=dev, dtype=torch.int64)
        for i in range(a.ncu_steps + 1):
            rt.advance_t(mark, 1)
            if i == a.ncu_steps:
                break
            step()
            torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
            opt.step()
        torch.cuda.synchronize()
        print("done", a.ncu_steps, "steps")
        return
    from sdb200 import graphs as G
    G.ENABLE_TRAIN_GRAPHS = False   # per-launch events need eager launches
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / 5
    with rt.KernelTimer() as kt:
        step()
    s = kt.summary()
    tot = sum(v["ms"] for v in s.values())
    print(f"batch {a.batch}: wall {wall * 1e3:.2f} ms/step (fwd+bwd, eager); library kernels {tot:.2f} ms in {sum(v['launches'] for v in s.values())} timed launches")
    for k, v in sorted(s.items(), key=lambda kv: -kv[1]["ms"]):
        unit = "TFLOP/s" if k in ("igemm", "attention", "wgrad", "attention_bwd") else "GB/s"
        perf = v["work"] / (v["ms"] * 1e-3) / (1e12 if unit == "TFLOP/s" else 1e9)
        print(f"  {k:16s} n={v['launches']:4d} {v['ms']:8.3f} ms  {perf:8.1f} {unit}")


if __name__ == "__main__":
    main()
