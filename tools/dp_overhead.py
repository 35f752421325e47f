"""Host-side cost of wrapping the model in sdb200.parallel.DataParallel, measured with a one-rank process group (no
collective runs): forward + backward + FusedAdam step time with and without the wrapper.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29544 tools/dp_overhead.py
"""
import copy
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
import sdb200  # noqa: E402,F401
from sdb200.optim import FusedAdam  # noqa: E402
from sdb200.parallel import DataParallel  # noqa: E402
from models.unet_cond_base import Unet  # noqa: E402


def main():
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    ucfg, _ = bench.load_cfg()
    torch.manual_seed(1111)
    base = Unet(4, ucfg).train().to(dev)
    x, text, mask = (v.to(dev) for v in bench.synth_inputs(16, ucfg, pin=False))
    t = torch.randint(0, 1000, (16,), device=dev)
    noise = torch.randn_like(x)
    for name, wrap in (("plain", False), ("DataParallel", True)):
        model = copy.deepcopy(base)
        net = DataParallel(model, device_ids=[dev.index]) if wrap else model
        opt = FusedAdam(model.parameters(), lr=1e-5, max_grad_norm=1.0)

        def step():
            opt.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = torch.nn.functional.mse_loss(net(x, t, {"text": text, "image": mask}), noise)
            loss.backward()
            opt.step()

        for _ in range(4):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            step()
        e1.record()
        torch.cuda.synchronize()
        print(f"{name:14s} {e0.elapsed_time(e1) / 20:.3f} ms/step", flush=True)
        del net, model, opt
    bench.shutdown(dist)


if __name__ == "__main__":
    main()
