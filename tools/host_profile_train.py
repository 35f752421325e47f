"""Host-side cost of one training step of the drop-in loop (reference train_ddpm_cond_celebhq_multi_gpu.py:327-378) when
the loss is read back every step: cProfile over 30 steps of add_noise -> forward -> mse -> backward -> FusedAdam.step ->
loss.item(). The GPU work is two graph replays plus a few kernels; whatever Python spends between the read-back and the
next replay is exposed time.

    python tools/host_profile_train.py"""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
import sdb200  # noqa: E402,F401
from models.unet_cond_base import Unet  # noqa: E402
from scheduler.linear_noise_scheduler import LinearNoiseScheduler  # noqa: E402
from sdb200.optim import FusedAdam  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    ucfg, _ = bench.load_cfg()
    torch.manual_seed(1)
    m = Unet(4, ucfg).train().to(dev)
    opt = FusedAdam(m.parameters(), lr=1e-5, max_grad_norm=1.0)
    sched = LinearNoiseScheduler(1000, 0.00085, 0.012)
    B = 16
    x = torch.randn(B, 4, 32, 32, device=dev)
    text = torch.randn(B, 77, 512, device=dev)
    mask = (torch.rand(B, 18, 32, 32, device=dev) > 0.5).float()
    noise = torch.randn_like(x)
    t = torch.randint(0, 1000, (B,), device=dev)

    def step():
        noisy = sched.add_noise(x, noise, t)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = torch.nn.functional.mse_loss(m(noisy, t, {"text": text, "image": mask}), noise)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        return loss.item()

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(30):
        step()
    wall = (time.perf_counter() - t0) / 30 * 1e3
    # host time alone: the same calls with the read-back replaced by nothing, timed until the host is done enqueueing
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(30):
        step()
    pr.disable()
    print(f"wall per step with loss.item(): {wall:.2f} ms")
    st = pstats.Stats(pr)
    st.sort_stats("cumulative").print_stats(28)


if __name__ == "__main__":
    main()
