"""bench.py — DDPM sampling throughput of the text+mask conditioned LDM UNet (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--cfg S] [--impl ours|reference]

Workload: the reference's `config/celebhq_text_image_cond.yaml` UNet (118.5 M parameters,
random init, seed 1111), batch 64 latents (4x32x32) per GPU, synthetic CLIP-shaped text context (B,77,512) and one-hot
mask (B,18,512,512); ancestral DDPM sampling over T=1000 timesteps followed by one VQVAE decode to 3x256x256.

A *step* is one DDPM timestep of the whole per-GPU batch: UNet forward (+ a second, unconditional forward when
--cfg > 1) -> sample_prev_timestep. K steps are timed with CUDA events; with the default K = 1000 the timed region is
literally the whole job (t = 999..0 and the VQVAE decode). For other K the images/s figure is
B / (1000 * t_step + t_decode) with both terms measured in the same run (stated in config.extrapolated).

    value : inputs resident in HBM, the device-resident loop (one CUDA-graph replay per timestep)
    e2e   : the reference-facing drop-in calls `model(xt, t, cond_input)` + `scheduler.sample_prev_timestep(...)`
            + `vae.decode(xt)` from pinned HOST buffers (H2D of latents/text/mask, per-step D2H of the x0 preview the
            reference script saves every step, D2H of the decoded images) inside the timed region

Multi-GPU (torchrun): the latent batch is sharded, B per rank, no collective on the data path (weak scaling);
time = max over ranks.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

T_STEPS = 1000
SEED = 1111
# algorithmic forward FLOPs per sample (FlopCounterMode on the reference modules)
UNET_GFLOP = 40.59
VQVAE_DECODE_GFLOP = 99.2


def load_cfg():
    """Model configs come from the committed golden fixtures (they embed the reference YAML sections), because
    /root/reference does not exist on the GPU box."""
    g = os.path.join(ROOT, "tests", "golden")
    unet = torch.load(os.path.join(g, "unet_full_cond.pt"), weights_only=False)["config"]
    vae = torch.load(os.path.join(g, "vqvae_decode.pt"), weights_only=False)["config"]
    return unet, vae


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tf_burst": d["bf16_tflops"], "tf_sustained": d["bf16_tflops_sustained"],
                "src": "measured"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "src": "fallback"}


def synth_inputs(batch, ucfg, rank=0, pin=True):
    """Seeded synthetic inputs on the HOST (pinned): x_T, text context, one-hot mask."""
    g = torch.Generator().manual_seed(SEED + 17 * rank)
    cc = ucfg["condition_config"]
    xt = torch.randn(batch, 4, 32, 32, generator=g)
    text = torch.randn(batch, 77, cc["text_condition_config"]["text_embed_dim"], generator=g)
    ic = cc["image_condition_config"]
    ch, mh, mw = ic["image_condition_input_channels"], ic["image_condition_h"], ic["image_condition_w"]
    cls = torch.randint(0, ch + 1, (batch, mh // 16, mw // 16), generator=g)  # blocky class map, 16x16 cells
    cls = cls.repeat_interleave(16, 1).repeat_interleave(16, 2)
    mask = torch.zeros(batch, ch, mh, mw)
    for c in range(ch):
        mask[:, c] = (cls == c + 1).float()
    if pin and torch.cuda.is_available():
        xt, text, mask = xt.pin_memory(), text.pin_memory(), mask.pin_memory()
    return xt, text, mask


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms while the timed region runs (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            self.th.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        # "under load": the upper half of the samples (idle samples before/after the region sit at low clocks)
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------- CPU baseline
def cpu_reference_run(ucfg, vcfg, steps, warmup, budget_s, cfg_scale=1.0, sample_batch=2):
    """The reference's CPU path for this workload: oracle/unet_torch.py + oracle/scheduler_np.py are the fp32 CPU
    restatement of models/unet_cond_base.py, scheduler/linear_noise_scheduler.py and models/vqvae.py:decode, pinned
    to the real reference by oracle/gen_golden.py (the reference itself is Python and cannot travel to the GPU box).
    Each step is one DDPM timestep on a BOUNDED sample of the workload (`sample_batch` images of the 64); the loop
    stops early when `budget_s` is spent. Returns per-image seconds for a timestep and for the decode."""
    import numpy as np
    from oracle import scheduler_np as onp
    from oracle import unet_torch as O
    from models.unet_cond_base import Unet
    from models.vqvae import VQVAE
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(SEED)
    usd = {k: v.clone() for k, v in Unet(4, ucfg).state_dict().items()}
    vsd = {k: v.clone() for k, v in VQVAE(3, vcfg).state_dict().items()}
    xt, text, mask = synth_inputs(sample_batch, ucfg, pin=False)
    cond = {"text": text, "image": mask}
    uncond = {"text": torch.zeros_like(text), "image": torch.zeros_like(mask)}
    tb = onp.make_tables(T_STEPS, 0.00085, 0.012)
    rng = np.random.default_rng(SEED)
    x = xt.clone()
    done, t_spent, t0_all = 0, 0.0, time.perf_counter()
    with torch.no_grad():
        for k in range(warmup + steps):
            i = max(T_STEPS - 1 - k, 0)
            t0 = time.perf_counter()
            tt = torch.full((sample_batch,), i, dtype=torch.long)
            eps = O.unet_forward(usd, ucfg, x, tt, cond)
            if cfg_scale > 1:
                eu = O.unet_forward(usd, ucfg, x, tt, uncond)
                eps = eu + cfg_scale * (eps - eu)
            z = rng.standard_normal(x.shape, dtype=np.float32)
            xp, _ = onp.sample_prev_timestep(tb, x.numpy(), eps.numpy(), i, z)
            x = torch.from_numpy(np.ascontiguousarray(xp))
            dt = time.perf_counter() - t0
            if k >= warmup:
                done += 1
                t_spent += dt
            if time.perf_counter() - t0_all > budget_s and done >= 1:
                break
        t0 = time.perf_counter()
        O.vqvae_decode(vsd, vcfg, x[:1].clamp(-1, 1))
        t_dec = time.perf_counter() - t0
    return {"t_step_img": t_spent / done / sample_batch, "t_dec_img": t_dec, "steps_done": done, "cores": cores,
            "sample_batch": sample_batch}


def images_per_s(batch, t_step_s, t_dec_s, t_in_s=0.0):
    return batch / (T_STEPS * t_step_s + t_dec_s + t_in_s)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ucfg, vcfg = load_cfg()
    sb = min(args.cpu_sample, args.batch)
    r2 = cpu_reference_run(ucfg, vcfg, 2, 1, budget_s=args.cpu_budget / 4, cfg_scale=args.cfg, sample_batch=2)
    r = cpu_reference_run(ucfg, vcfg, args.steps, min(args.warmup, 1), budget_s=args.cpu_budget, cfg_scale=args.cfg,
                          sample_batch=sb)
    ips = images_per_s(1, r["t_step_img"], r["t_dec_img"])
    sample = (f"{r['steps_done']} timed DDPM timesteps (of {args.steps} requested; stops at {args.cpu_budget:.0f} s) on "
              f"{r['sample_batch']} of the {args.batch} images + VQVAE decode of 1 image, fp32, oracle port of the "
              f"reference on {r['cores']} host threads; images/s = 1 / (1000 * s_per_image_step + s_per_image_decode); "
              f"per-image step time {1e3 * r2['t_step_img']:.1f} ms at batch 2 -> {1e3 * r['t_step_img']:.1f} ms at batch "
              f"{r['sample_batch']}")
    line = {
        "impl": "reference", "metric": "ddpm_sampling_images_per_s", "value": ips, "unit": "images/s",
        "n_gpus": args.gpus, "steps": r["steps_done"], "warmup": min(args.warmup, 1),
        "ms_per_step": 1e3 * r["t_step_img"] * args.batch, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, ucfg),
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": r["cores"], "kind": "port", "sample": sample},
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(args, ucfg):
    return {
        "workload": "configs[1]: text+mask conditioned LDM UNet (config/celebhq_text_image_cond.yaml), 1000-step DDPM "
                    "ancestral sampling + VQVAE decode, batch 64 per GPU",
        "batch_per_gpu": args.batch, "timesteps": T_STEPS, "cf_guidance_scale": args.cfg,
        "latent": [4, 32, 32], "text_ctx": [77, 512], "mask": [18, 512, 512], "image": [3, 256, 256],
        "unet_params": 118513466, "parallelism": f"batch-sharded x{args.gpus}, no collective",
        "l2": "inputs larger than L2: 237 MB of bf16 weights + >1 GB of activations are touched per step (L2 = 126 MB)",
    }


# ---------------------------------------------------------------------------------------------- our arm
def run_ours(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this package has no CPU path; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(vals):
        t = torch.tensor(vals, device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()

    import __graft_entry__ as ge
    ge.build()
    import sdb200  # noqa: F401
    from sdb200 import runtime as rt
    from sdb200.graphs import SamplingPipeline
    from models.unet_cond_base import Unet
    from models.vqvae import VQVAE
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler

    ucfg, vcfg = load_cfg()
    B, K, W = args.batch, args.steps, max(args.warmup, 3)
    torch.manual_seed(SEED)
    model = Unet(4, ucfg).eval().to(dev)
    vae = VQVAE(3, vcfg).eval().to(dev)
    sched = LinearNoiseScheduler(T_STEPS, 0.00085, 0.012)
    sched.seed = SEED + rank
    xt_h, text_h, mask_h = synth_inputs(B, ucfg, rank)
    xt_d, text_d, mask_d = xt_h.to(dev), text_h.to(dev), mask_h.to(dev)
    cond = {"text": text_d, "image": mask_d}
    uncond = {"text": torch.zeros_like(text_d), "image": torch.zeros_like(mask_d)} if args.cfg > 1 else None

    # ------------------------------------------------------------------ value: device-resident loop
    pipe = SamplingPipeline(model, sched, vae, use_graph=not args.no_graph)
    with torch.no_grad():
        P = pipe.plan(xt_d, cond, uncond, args.cfg)
        P.xt.copy_(xt_d)
        P.t.fill_(T_STEPS - 1)
        for _ in range(W):
            pipe.step(P)
        vae.decode(P.xt.clamp(-1, 1))
        l0 = rt.launch_count()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        P.xt.copy_(xt_d)
        t_first = min(K, T_STEPS) - 1
        P.t.fill_(t_first)
        barrier()
        with ClockSampler(local) as clk:
            ev[0].record()
            for k in range(K):
                if k and k % T_STEPS == 0:
                    P.t.fill_(T_STEPS - 1)
                pipe.step(P)
            ev[1].record()
            ims = vae.decode(P.xt)
            ev[2].record()
            barrier()
        launches = rt.launch_count() - l0
        ms_steps, ms_dec = max_over_ranks([ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])])
        finite = bool(torch.isfinite(ims).all().item())
    t_step, t_dec = ms_steps / K / 1e3, ms_dec / 1e3
    value = world * images_per_s(B, t_step, t_dec)

    # ------------------------------------------------------------------ e2e: drop-in API from host buffers
    x0_host = torch.empty(B, 4, 32, 32).pin_memory()
    ims_host = torch.empty(B, 3, 256, 256).pin_memory()
    Ke = min(K, args.e2e_steps) if args.e2e_steps > 0 else K

    def e2e_job(nsteps, events=None):
        rec = (lambda i: events[i].record()) if events else (lambda i: None)
        with torch.no_grad():
            rec(0)
            xt = xt_h.to(dev, non_blocking=True)
            c = {"text": text_h.to(dev, non_blocking=True), "image": mask_h.to(dev, non_blocking=True)}
            u = {"text": torch.zeros_like(c["text"]), "image": torch.zeros_like(c["image"])} if args.cfg > 1 else None
            rec(1)
            for i in reversed(range(nsteps)):
                t = torch.full((B,), i, device=dev, dtype=torch.long)
                eps = model(xt, t, c)
                if u is not None:
                    eps_u = model(xt, t, u)
                    xt, x0 = sched.sample_prev_timestep(xt, eps, torch.as_tensor(i), noise_pred_uncond=eps_u,
                                                        cf_guidance_scale=args.cfg)
                else:
                    xt, x0 = sched.sample_prev_timestep(xt, eps, torch.as_tensor(i))
                x0_host.copy_(x0, non_blocking=True)  # the per-step preview the reference script pulls to the host
            rec(2)
            out = vae.decode(xt)
            ims_host.copy_(out, non_blocking=True)
            rec(3)

    e2e_job(W)
    ee = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    barrier()
    wall0 = time.perf_counter()
    e2e_job(Ke, ee)
    barrier()
    wall = time.perf_counter() - wall0
    ms_in, ms_loop, ms_out = max_over_ranks([ee[0].elapsed_time(ee[1]), ee[1].elapsed_time(ee[2]),
                                             ee[2].elapsed_time(ee[3])])
    e2e_value = world * images_per_s(B, ms_loop / Ke / 1e3, ms_out / 1e3, ms_in / 1e3)
    h2d = (xt_h.numel() + text_h.numel() + mask_h.numel()) * 4 / T_STEPS
    d2h = x0_host.numel() * 4 + ims_host.numel() * 4 / T_STEPS

    # ------------------------------------------------------------------ roofline of the dominant kernel
    with torch.no_grad():
        P.xt.copy_(xt_d)
        P.t.fill_(500)
        P.body()
        torch.cuda.synchronize()
        # hold the stream busy while the host enqueues the whole step, so that the event pairs bracket kernels that run
        # back to back as they do inside the captured graph (an idle stream adds the launch latency to every pair)
        torch.cuda._sleep(int(3.0e7))
        with rt.KernelTimer() as kt:
            P.body()
        ksum = kt.summary()
        rows = kt.rows()
    pk = peaks()
    traffic = None  # DRAM bytes per igemm launch from the committed ncu pass of one step (profiles/README.md)
    tpath = os.path.join(ROOT, "profiles", "igemm_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
        except (OSError, ValueError):
            traffic = None
    ig = ksum.get("igemm", {"launches": 0, "ms": 1e-9, "work": 0.0})
    ach = ig["work"] / (ig["ms"] * 1e-3) / 1e12
    eager_ms = sum(v["ms"] for v in ksum.values())
    roofline = {
        "kernel": "igemm_kernel (tcgen05 implicit GEMM: every conv / linear of the step)", "bound": "tensor",
        "achieved": ach, "peak": pk["tf_sustained"], "unit": "TFLOP/s", "frac": ach / pk["tf_sustained"],
        "peak_source": f"{pk['src']} MEASURED_PEAKS.json bf16_tflops_sustained (kernel timed inside a long step)",
        "traffic": traffic, "launches_per_step": ig["launches"], "avg_launch_us": 1e3 * ig["ms"] / max(ig["launches"], 1),
        "algorithmic_gflop_per_step": ig["work"] / 1e9,
    }
    kernels = {k: {"launches": v["launches"], "ms": round(v["ms"], 4), "share": round(v["ms"] / eager_ms, 4),
                   "achieved": (round(v["work"] / (v["ms"] * 1e-3) / 1e12, 2) if k in ("igemm", "attention")
                                else round(v["work"] / (v["ms"] * 1e-3) / 1e9, 1)),
                   "unit": "TFLOP/s" if k in ("igemm", "attention") else "GB/s"} for k, v in ksum.items()}
    if rank == 0 and args.dump:
        os.makedirs(os.path.dirname(os.path.abspath(args.dump)) or ".", exist_ok=True)
        with open(args.dump, "w") as f:
            json.dump({"per_launch": [{"kernel": n, "work": w, "us": round(ms * 1e3, 2), "shape": d}
                                      for n, w, ms, d in rows], "summary": kernels}, f)

    parity = None
    if rank == 0:
        try:
            parity = parity_check(dev)
        except Exception as exc:   # noqa: BLE001 - reported, never fatal for the measurement
            parity = {"error": f"{type(exc).__name__}: {exc}"[:200]}
    step_flops = B * UNET_GFLOP * 1e9 * (2 if args.cfg > 1 else 1)
    line = {
        "metric": "ddpm_sampling_images_per_s", "value": value, "unit": "images/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": 1e3 * t_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": dict(workload_config(args, ucfg), extrapolated=(K != T_STEPS), decode_ms=ms_dec,
                       cuda_graph=not args.no_graph, outputs_finite=finite, parity_mode="fast", parity=parity),
        "step_tflops": step_flops / t_step / 1e12, "step_frac_of_bf16_sustained": step_flops / t_step / 1e12 / pk["tf_sustained"],
        "decode_tflops": B * VQVAE_DECODE_GFLOP * 1e9 / t_dec / 1e12,
        "roofline": roofline, "kernels": kernels,
        "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": Ke, "ms_per_step": ms_loop / Ke, "inputs_h2d_ms": ms_in, "decode_d2h_ms": ms_out,
                "wall_s": wall, "api": "models.unet_cond_base.Unet.forward + LinearNoiseScheduler.sample_prev_timestep "
                                        "+ VQVAE.decode (drop-in module calls, pinned host buffers)"},
        "gpu_launches": launches, "clocks": clk.summary(),
    }
    if rank == 0 and world == 1 and not args.no_cpu:
        r2 = cpu_reference_run(ucfg, vcfg, 2, 1, budget_s=args.cpu_budget / 4, cfg_scale=args.cfg, sample_batch=2)
        r = cpu_reference_run(ucfg, vcfg, 3, 1, budget_s=args.cpu_budget, cfg_scale=args.cfg,
                              sample_batch=min(args.cpu_sample, B))
        ips = images_per_s(1, r["t_step_img"], r["t_dec_img"])
        line["cpu_baseline"] = {
            "value": ips, "unit": "images/s", "cores": r["cores"], "kind": "port",
            "sample": f"{r['steps_done']} DDPM timesteps on {r['sample_batch']} of the {B} images + VQVAE decode of 1 "
                      f"image, fp32 oracle port on {r['cores']} host threads, extrapolated to 1000 steps per image; "
                      f"per-image step time {1e3 * r2['t_step_img']:.1f} ms at batch 2 -> {1e3 * r['t_step_img']:.1f} ms "
                      f"at batch {r['sample_batch']}"}
    del pipe, P, model, vae
    torch.cuda.empty_cache()

    def leg(name, fn):
        """A secondary leg: its full record is printed as its own JSON line ({"leg": name, ...}) BEFORE the headline line,
        the headline keeps a one-line summary; a failing leg is recorded, the headline still prints."""
        try:
            rec = fn()
        except Exception as exc:   # noqa: BLE001
            rec = {"error": f"{type(exc).__name__}: {exc}"[:400]}
        if rank == 0:
            print(json.dumps({"leg": name, "n_gpus": world, **rec}), flush=True)
        return rec

    def brief(rec, keys=("value", "unit", "ms_per_step", "batch_per_gpu")):
        if "error" in rec:
            return {"error": rec["error"][:120]}
        return {k: (round(rec[k], 3) if isinstance(rec[k], float) else rec[k]) for k in keys if k in rec}

    if args.train_steps > 0:
        tr = leg("train", lambda: train_leg(args, ucfg, dev, rank, world, barrier, max_over_ranks))
        line["train"] = dict(brief(tr), algorithmic_tflops=round(tr.get("algorithmic_tflops", 0.0), 1),
                             e2e=brief(tr.get("e2e", {}), ("value", "ms_per_step", "h2d_bytes_per_step")))
        if args.train_sweep and "error" not in tr:
            sweep = {}
            for tb in (32, 64, 128):
                a2 = argparse.Namespace(**vars(args))
                a2.train_batch, a2.train_steps = tb, max(5, args.train_steps // 2)
                r = leg(f"train_b{tb}", lambda: train_leg(a2, ucfg, dev, rank, world, barrier, max_over_ranks, e2e=False))
                sweep[str(tb)] = dict(brief(r, ("value", "ms_per_step")), algorithmic_tflops=round(r.get("algorithmic_tflops", 0.0), 1))
            line["train"]["batch_sweep"] = sweep
    if args.dit_steps > 0:
        dl = leg("dit", lambda: dit_leg(args, ucfg, dev, rank, world, barrier, max_over_ranks, t_dec))
        line["dit"] = {k: brief(v, ("value", "unit", "ms_per_step", "batch_per_gpu", "step_tflops")) for k, v in dl.items()
                       if isinstance(v, dict)} if "error" not in dl else brief(dl)
    if rank == 0 and world == 1:
        if not args.no_hbm:
            hb = leg("hbm_kernels", lambda: hbm_leg(dev, pk))
            line["hbm_kernels"] = {k: {"gbps": round(v["gbps"], 1), "frac": round(v["frac"], 3), "us": round(v["us"], 1)}
                                   for k, v in hb.items() if isinstance(v, dict)} if "error" not in hb else brief(hb)
        if not args.no_config0:
            c0 = leg("config0", lambda: config0_leg(args, dev, 0 if args.no_cpu else args.cpu_budget / 2))
            line["config0"] = {"gpu_job_s": round(c0["gpu"]["job_s"], 4), "cpu_job_s": round(c0.get("cpu", {}).get("job_s", 0.0), 2),
                               "gpu_over_cpu": round(c0.get("gpu_over_cpu", 0.0), 1)} if "error" not in c0 else brief(c0)
        if not args.no_library_bar:
            lb = leg("library_bar", lambda: library_bar_leg(args, ucfg, vcfg, dev))
            if "error" not in lb:
                sb16, s32 = lb["sampling_bf16_autocast"], lb["sampling_fp32"]
                line["library_bar"] = {
                    "what": "reference's stock PyTorch path (cuDNN/cuBLAS/SDPA, eager) on this GPU",
                    "sampling_bf16_autocast": {"images_per_s": round(sb16["value"], 3), "ms_per_step": round(sb16["ms_per_step"], 2)},
                    "sampling_fp32": {"images_per_s": round(s32["value"], 3), "ms_per_step": round(s32["ms_per_step"], 2)},
                    "ours_over_bar_sampling": round(value / sb16["value"], 2)}
                if "train_bf16_autocast" in lb and "value" in line.get("train", {}):
                    tb16 = lb["train_bf16_autocast"]
                    line["library_bar"]["train_bf16_autocast"] = {"samples_per_s": round(tb16["value"], 1),
                                                                  "ms_per_step": round(tb16["ms_per_step"], 2)}
                    line["library_bar"]["ours_over_bar_train"] = round(line["train"]["value"] / tb16["value"], 2)
            else:
                line["library_bar"] = brief(lb)
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        shutdown(dist)


def shutdown(dist):
    """Leave the process group without risking a hang at exit: captured graphs that still hold NCCL nodes (the in-graph
    gradient buckets of the training legs) can keep destroy_process_group() waiting; give it ten seconds after a
    collect + barrier, then exit the process."""
    import gc
    import threading
    gc.collect()
    torch.cuda.synchronize()
    try:
        dist.barrier()
    except Exception:
        pass
    sys.stdout.flush()
    sys.stderr.flush()
    th = threading.Thread(target=dist.destroy_process_group, daemon=True)
    th.start()
    th.join(10.0)
    if th.is_alive():
        os._exit(0)


# Model_DiT_12L_config.py:125-133 (image-conditioned, 12 layers) and Model_DiT_9L_config.py:139-148 (text + image, 9 layers)
def dit_config(layers, kinds):
    cc = {"condition_types": kinds}
    if "text" in kinds:
        cc["text_condition_config"] = {"text_embed_model": "clip", "train_text_embed_model": False, "text_embed_dim": 512,
                                       "cond_drop_prob": 0.9}
    if "image" in kinds:
        cc["image_condition_config"] = {"image_condition_input_channels": 18, "image_condition_output_channels": 3,
                                        "image_condition_h": 512, "image_condition_w": 512, "cond_drop_prob": 0.9}
    return {"hidden_size": 288, "patch_size": 2, "timestep_emb_dim": 192, "num_layers": layers, "num_heads": 9,
            "head_dim": 32, "condition_config": cc}


DIT12_GFLOP, DIT9_GFLOP = 7.04, 6.69     # forward per sample, SURVEY §8d [probed]


def dit_leg(args, ucfg, dev, rank, world, barrier, max_over_ranks, t_dec):
    """BASELINE.json configs[3] (DiT-12L, image-conditioned: sampling + data-parallel training, batch sharded) and
    configs[4] (DiT-9L text+image with every Linear / the mask conv replaced by the cim_layers drop-ins — first the LSQ
    layers, then the ADDA compute-in-memory layers — 8-bit activations, 4-bit weights, weight noise 0.08: one
    quant-noise training step; the Enhance wrappers and PercentOptimizerFP of that pipeline are not part of it). Same step definitions as the
    UNet legs; the VQVAE decode time of the main leg is reused for images/s."""
    import torch.nn as nn
    from sdb200 import runtime as rt
    from sdb200.graphs import SamplingPipeline
    from models.transformer import DIT
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    from cim_layers.layers_qn_lsq import Conv2d_qn_lsq, Linear_qn_lsq

    def fresh(layers, kinds):
        torch.manual_seed(SEED)
        m = DIT(4, dit_config(layers, kinds))
        g = torch.Generator().manual_seed(SEED + 7)
        with torch.no_grad():   # the adaLN / proj_out parameters are zero-initialised: a default-init DiT outputs 0
            for name, p in m.named_parameters():
                if "adaptive_norm_layer" in name or name.startswith("proj_out"):
                    p.copy_(torch.randn(p.shape, generator=g) * 0.02)
        return m

    B, K, W = args.batch, args.dit_steps, 5
    model = fresh(12, ["image"]).eval().to(dev)
    sched = LinearNoiseScheduler(T_STEPS, 0.00085, 0.012)
    sched.seed = SEED + rank
    xt_h, text_h, mask_h = synth_inputs(B, ucfg, rank)
    xt_d, mask_d = xt_h.to(dev), mask_h.to(dev)
    pipe = SamplingPipeline(model, sched, None)
    with torch.no_grad():
        P = pipe.plan(xt_d, {"image": mask_d}, None, 1.0)
        P.xt.copy_(xt_d)
        P.t.fill_(T_STEPS - 1)
        for _ in range(W):
            pipe.step(P)
        with rt.KernelTimer() as kt:
            P.body()
        ksum = kt.summary()
        if rank == 0 and args.dump:
            with open(args.dump.replace(".json", "") + "_dit.json", "w") as f:
                json.dump({"per_launch": [{"kernel": n, "work": w, "us": round(ms_ * 1e3, 2), "shape": d}
                                          for n, w, ms_, d in kt.rows()]}, f)
        l0 = rt.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        P.t.fill_(min(K, T_STEPS) - 1)
        barrier()
        e0.record()
        for _ in range(K):
            pipe.step(P)
        e1.record()
        barrier()
        (ms,) = max_over_ranks([e0.elapsed_time(e1)])
        launches = rt.launch_count() - l0
    t_step = ms / K / 1e3
    tot = sum(v["ms"] for v in ksum.values())
    kernels = {k: {"launches": v["launches"], "ms": round(v["ms"], 4), "share": round(v["ms"] / tot, 4)}
               for k, v in sorted(ksum.items(), key=lambda kv: -kv[1]["ms"])}
    ln = ksum.get("layernorm")
    out = {
        "sampling": {"model": "DiT-12L image-conditioned (Model_DiT_12L_config)", "metric": "ddpm_sampling_images_per_s",
                     "value": world * images_per_s(B, t_step, t_dec), "unit": "images/s", "ms_per_step": ms / K,
                     "steps": K, "batch_per_gpu": B, "step_tflops": DIT12_GFLOP * 1e9 * B / t_step / 1e12,
                     "gpu_launches": launches, "kernels": kernels,
                     "layernorm_gbps": None if ln is None else ln["work"] / (ln["ms"] * 1e-3) / 1e9},
    }
    del pipe, P, model
    if args.train_steps <= 0:
        return out
    out["train"] = train_leg(args, ucfg, dev, rank, world, barrier, max_over_ranks, make_model=lambda: fresh(12, ["image"]),
                             gflop_fwd=DIT12_GFLOP, tag="DiT-12L image-conditioned", batch=args.dit_train_batch)

    def cim_model(kind):
        from cim_layers.layers_qn_lsq_adda_cim import Conv2d_lsq_adda_cim, Linear_lsq_adda_cim, gen_weight_split_dict
        m = fresh(9, ["text", "image"])
        kw = dict(weight_bit=4, input_bit=8, output_bit=8, noise_scale=0.08)      # Model_DiT_9L_train.py:541-548
        lin_cls, conv_cls = Linear_qn_lsq, Conv2d_qn_lsq
        if kind == "adda":                                                          # :578-590
            kw.update(dac_bit=5, adc_bit=8, adc_gain_1_scale=9.071428571, adc_gain_range=[1 / 64, 1 / 64],
                      adc_adjust_mode="current")
            lin_cls, conv_cls = Linear_lsq_adda_cim, Conv2d_lsq_adda_cim
        for parent in list(m.modules()):
            for name, c in list(parent.named_children()):
                new = None
                if type(c) is nn.Linear:
                    new = lin_cls(c.in_features, c.out_features, bias=c.bias is not None, **kw)
                elif type(c) is nn.Conv2d:
                    new = conv_cls(c.in_channels, c.out_channels, c.kernel_size, c.stride, c.padding, c.groups,
                                   bias=c.bias is not None, **kw)
                if new is not None:
                    new.weight, new.bias = c.weight, c.bias
                    if kind == "adda":                                              # array [576, 2048], :594-600
                        new.weight_mapping_info = gen_weight_split_dict(new, [576, 2048])
                    setattr(parent, name, new)
        return m

    out["cim_lsq_train"] = train_leg(args, ucfg, dev, rank, world, barrier, max_over_ranks,
                                     make_model=lambda: cim_model("lsq"), gflop_fwd=DIT9_GFLOP,
                                     tag="DiT-9L text+image, all 95 Linear + mask conv as LSQ layers")
    def enhanced_lsq_model():
        """LSQ stage of the pipeline as Model_DiT_9L_train.py:541-553 leaves it: every quantised Linear wrapped in the
        low-rank blend branch, then followed by a grouped enhance layer; the 1x1 mask conv likewise with the Conv2d
        wrappers (progressive_qn_train.py:884-950)."""
        import cim_qn_train.layers_enhance as en
        m = cim_model("lsq")
        for n, c in [(n, c) for n, c in m.named_modules() if type(c) in (Linear_qn_lsq, Conv2d_qn_lsq)]:
            parent = m.get_submodule(n.rsplit(".", 1)[0]) if "." in n else m
            if type(c) is Linear_qn_lsq:
                br = en.EnhanceBranchLinear_LoR(original_linear=c, relu=False, sigmoid=True, ops_factor=0.05)
                br.original_linear = en.EnhanceLayerLinear(c, ops_factor=0.05)
            else:
                br = en.EnhanceBranchConv2d_LoR(original_conv=c, relu=False, sigmoid=True, ops_factor=0.05)
                br.original_conv = en.EnhanceLayerConv2d(c, ops_factor=0.05)
            setattr(parent, n.rsplit(".", 1)[-1], br)
        return m

    out["cim_lsq_enhance_train"] = train_leg(
        args, ucfg, dev, rank, world, barrier, max_over_ranks, make_model=enhanced_lsq_model, gflop_fwd=DIT9_GFLOP,
        optimizer="percent",
        tag="DiT-9L text+image, LSQ leaves + EnhanceBranch*_LoR + EnhanceLayer* (ops_factor 0.05) around all 95 Linear "
            "and the mask conv, PercentOptimizerFP")
    out["cim_adda_train"] = train_leg(args, ucfg, dev, rank, world, barrier, max_over_ranks,
                                      make_model=lambda: cim_model("adda"), gflop_fwd=DIT9_GFLOP,
                                      tag="DiT-9L text+image, all 95 Linear + mask conv as ADDA compute-in-memory layers "
                                          "(2 DAC slices, 8-bit ADC, array 576 x 2048)")
    return out


def train_leg(args, ucfg, dev, rank, world, barrier, max_over_ranks, make_model=None, gflop_fwd=None, tag="unet",
              optimizer=None, e2e=True, batch=None):
    """BASELINE.json configs[2]: data-parallel training of the same conditioned UNet as
    train_ddpm_cond_celebhq_multi_gpu.py:341-378 does it — add_noise, bf16-autocast forward, MSE, backward (DDP's
    bucketed NCCL all-reduce overlapped with it when world > 1), clip_grad_norm_(1.0), Adam, EMA — per-GPU batch 16
    (config/celebhq_text_image_cond.py:76). Forward and backward of every heavy op run on libb200sd.so.
    value: inputs resident on the device. e2e: latents / text / mask copied from pinned host memory every step (on the
    copy stream of sdb200.data.DevicePrefetcher, one batch ahead) and
    loss.item() read back every step, as the reference loop does (:327-354)."""
    import copy
    import torch.distributed as dist
    from sdb200 import runtime as rt
    from models.unet_cond_base import Unet
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    B, K, W = (batch or args.train_batch), args.train_steps, 3
    torch.manual_seed(SEED)
    model = (make_model() if make_model is not None else Unet(4, ucfg)).train().to(dev)
    gflop_fwd = UNET_GFLOP if gflop_fwd is None else gflop_fwd
    cond_keys = [k for k, on in (("text", model.text_cond), ("image", model.image_cond)) if on]
    ema = copy.deepcopy(model).eval()
    net = model
    if world > 1:
        if args.train_ddp == "torch":
            net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[dev.index], broadcast_buffers=False)
        else:
            from sdb200.parallel import DataParallel
            net = DataParallel(model, device_ids=[dev.index], broadcast_buffers=False)
    fused = args.train_opt == "fused" and optimizer is None
    percent = optimizer == "percent"     # the CIM trainers' optimizer (Model_DiT_9L_train.py:266), no clip / EMA there
    if percent:
        from cim_layers.IBA_optimizer import PercentOptimizerFP
        opt = PercentOptimizerFP(model.parameters(), lr=1e-5)
    elif fused:
        from sdb200.optim import FusedAdam
        opt = FusedAdam(model.parameters(), lr=1e-5, max_grad_norm=1.0, ema_params=ema.parameters(), ema_decay=0.999)
    else:
        opt = torch.optim.Adam(model.parameters(), lr=1e-5)
    sched = LinearNoiseScheduler(T_STEPS, 0.00085, 0.012)
    x_h, text_h, mask_h = synth_inputs(B, ucfg, rank)
    g = torch.Generator().manual_seed(SEED + 5 + rank)
    noise_h = torch.randn(x_h.shape, generator=g).pin_memory()
    t_h = torch.randint(0, T_STEPS, (B,), generator=g).pin_memory()
    res = {"x": x_h.to(dev), "text": text_h.to(dev), "mask": mask_h.to(dev), "noise": noise_h.to(dev), "t": t_h.to(dev)}

    def step(batch, from_host):
        x0, text, mask, noise, t = batch
        noisy = sched.add_noise(x0, noise, t)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            pred = net(noisy, t, {k: v for k, v in (("text", text), ("image", mask)) if k in cond_keys})
            loss = torch.nn.functional.mse_loss(pred, noise)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        if percent:
            opt.step()      # Adam moments + per-tensor max-scaled update: two launches (csrc/optim.cu)
        elif fused:
            opt.step()      # clip_grad_norm_(1.0) + finite check + Adam + EMA: two launches (sdb200/optim.py)
        else:
            torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
            opt.step()
            with torch.no_grad():
                torch._foreach_mul_(list(ema.parameters()), 0.999)
                torch._foreach_add_(list(ema.parameters()), list(model.parameters()), alpha=0.001)
        return loss.item() if from_host else loss.detach()

    def batches(k, from_host):
        """k batches: device-resident tensors, or the pinned host buffers uploaded every step — batch i+1 on the
        prefetcher's copy stream while step i computes (sdb200/data.py)."""
        if not from_host:
            return ((res["x"], res["text"], res["mask"], res["noise"], res["t"]) for _ in range(k))
        from sdb200.data import DevicePrefetcher
        # the mask travels under its dict key so that the prefetcher can pre-select the 32x32 pixels the model reads
        gen = ((x_h, text_h, {"image": mask_h}, noise_h, t_h) for _ in range(k))
        pf = DevicePrefetcher(gen, dev, shrink={"image": (32, 32)} if not args.full_mask_upload else None)
        return ((x, tx, m["image"], nz, tt) for x, tx, m, nz, tt in pf)

    out = {}
    for name, from_host in ((("value", False), ("e2e", True)) if e2e else (("value", False),)):
        for b in batches(W, from_host):
            step(b, from_host)
        l0 = rt.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for b in batches(K, from_host):
            last = step(b, from_host)
        e1.record()
        barrier()
        (ms,) = max_over_ranks([e0.elapsed_time(e1)])
        out[name] = (world * B * K / (ms / 1e3), ms / K, rt.launch_count() - l0, float(last))
    mask_up = mask_h.numel() if args.full_mask_upload else mask_h.shape[0] * mask_h.shape[1] * 32 * 32
    h2d = (x_h.numel() + text_h.numel() + mask_up + noise_h.numel()) * 4 + t_h.numel() * 8
    ema_params = list(ema.parameters())
    del net, model, opt
    return {
        "metric": "ddp_train_samples_per_s", "value": out["value"][0], "unit": "samples/s", "ms_per_step": out["value"][1],
        "steps": K, "warmup": W, "batch_per_gpu": B, "global_batch": B * world, "dtype": "bf16",
        "model": tag, "params": sum(p.numel() for p in ema_params),
        "algorithmic_tflops": 3 * gflop_fwd * 1e9 * B / (out["value"][1] / 1e3) / 1e12,
        "optimizer": ("cim_layers.IBA_optimizer.PercentOptimizerFP on b200sd_percent_opt_step: two launches" if percent else
                      "sdb200.optim.FusedAdam: clip_grad_norm_(1.0) + finite check + Adam + EMA in two launches" if fused
                      else "Adam + clip_grad_norm_(1.0) + EMA (torch foreach)"), "gpu_launches": out["value"][2],
        "grad_allreduce": None if world == 1 else (
            "torch DistributedDataParallel: 25 MB NCCL buckets, reduced after the backward graph has replayed"
            if args.train_ddp == "torch" else
            "sdb200.parallel.DataParallel: in-place NCCL AVG all-reduces of 32 MB slices of the flat fp32 gradient "
            "buffer (laid out in gradient-arrival order), captured inside the backward graph on a side stream as each "
            "bucket's last gradient is produced"),
        "loss_finite": bool(out["value"][3] == out["value"][3]),
        "e2e": None if not e2e else {"value": out["e2e"][0], "unit": "samples/s", "ms_per_step": out["e2e"][1],
                                     "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4},
    }


def parity_check(dev):
    """Measured error of both parity modes on the committed full-config golden fixture (the REFERENCE's fp32 output of
    the conditioned UNet, tests/golden/unet_full_cond.pt): what `config.parity_mode` costs in accuracy."""
    from sdb200 import engine as E
    from models.unet_cond_base import Unet
    fx = torch.load(os.path.join(ROOT, "tests", "golden", "unet_full_cond.pt"), weights_only=False)
    torch.manual_seed(fx["seed"])
    m = Unet(4, fx["config"]).eval().to(dev)
    inp = fx["inputs"]
    ch = fx["config"]["condition_config"]["image_condition_config"]["image_condition_input_channels"]
    mask = torch.nn.functional.one_hot(inp["mask_classes"].long(), ch + 1)[..., 1:].permute(0, 3, 1, 2).float().contiguous()
    cond = {"text": inp["text"].to(dev), "image": mask.to(dev)}
    out = {"fixture": "tests/golden/unet_full_cond.pt (reference fp32 output)", "bound": 1e-2,
           "reference_bf16_autocast": round(float(fx["ref_bf16_autocast_relerr"]), 5)}
    ref = fx["output"].float()
    for mode in ("fast", "precise"):
        E.set_parity_mode(mode)
        try:
            with torch.no_grad():
                y = m(inp["x"].to(dev), inp["t"].to(dev), cond).cpu()
        finally:
            E.set_parity_mode("fast")
        out[mode] = {"max_norm": round(float((y - ref).abs().max() / ref.abs().max()), 5),
                     "rel_l2": round(float((y - ref).norm() / ref.norm()), 5)}
    del m
    return out


# ---------------------------------------------------------------------------------------------- library bar
def library_bar_leg(args, ucfg, vcfg, dev, steps=10, warmup=3):
    """The bar this repo has to beat on the same GPU (BASELINE.md §4 "second bar"): the reference's own
    stock path — PyTorch modules over cuDNN / cuBLAS / SDPA / ATen — restated functionally by oracle/unet_torch.py
    (pinned to the real reference), run on cuda:0 with none of this repo's kernels. Two regimes:
      bf16  torch.autocast(bfloat16) around the model call, as train_ddpm_cond_celebhq_multi_gpu.py:281-283,341-362 does
      fp32  no autocast, as tools/sample_ddpm_text_image_cond.py:79-91 runs (cuDNN convolutions may use TF32, matmuls fp32)
    Sampling: batch 64, eager per-timestep loop of the reference script (model -> sample_prev_timestep with host-side
    z and table uploads) + VQVAE decode. Training: batch 16, add_noise -> forward -> MSE -> backward -> clip -> Adam -> EMA.
    Attention is F.scaled_dot_product_attention (the library's fused kernel), which favours the bar."""
    from oracle import scheduler_torch as ST
    from oracle import unet_torch as O
    from models.unet_cond_base import Unet
    from models.vqvae import VQVAE
    O.USE_SDPA = True
    torch.manual_seed(SEED)
    usd = {k: v.to(dev) for k, v in Unet(4, ucfg).state_dict().items()}
    vsd = {k: v.to(dev) for k, v in VQVAE(3, vcfg).state_dict().items()}
    B = args.batch
    xt_h, text_h, mask_h = synth_inputs(B, ucfg, 0, pin=False)
    cond = {"text": text_h.to(dev), "image": mask_h.to(dev)}
    tb = ST.tables(T_STEPS, 0.00085, 0.012)
    out = {"what": "reference's stock PyTorch path (cuDNN/cuBLAS/SDPA) on the same GPU, eager, none of this repo's kernels",
           "torch": torch.__version__, "cudnn_allow_tf32": bool(torch.backends.cudnn.allow_tf32)}

    def sampling(autocast):
        x = xt_h.to(dev)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        with torch.no_grad():
            for k in range(warmup + steps):
                if k == warmup:
                    torch.cuda.synchronize()
                    e0.record()
                i = T_STEPS - 1 - k
                with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
                    eps = O.unet_forward(usd, ucfg, x, torch.full((B,), i, device=dev, dtype=torch.long), cond)
                x, _ = ST.sample_prev_timestep(tb, x, eps.float(), torch.as_tensor(i))
            e1.record()
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
                ims = O.vqvae_decode(vsd, vcfg, x.clamp(-1, 1))
            e2.record()
            torch.cuda.synchronize()
        t_step, t_dec = e0.elapsed_time(e1) / steps / 1e3, e1.elapsed_time(e2) / 1e3
        return {"value": images_per_s(B, t_step, t_dec), "unit": "images/s", "ms_per_step": 1e3 * t_step,
                "decode_ms": 1e3 * t_dec, "steps": steps, "batch": B, "finite": bool(torch.isfinite(ims).all().item())}

    out["sampling_bf16_autocast"] = sampling(True)
    out["sampling_fp32"] = sampling(False)

    if args.train_steps > 0:
        Bt = args.train_batch
        leaves = {k: v.clone().requires_grad_(v.dtype.is_floating_point) for k, v in usd.items()}
        params = [v for v in leaves.values() if v.requires_grad]
        ema = [p.detach().clone() for p in params]
        opt = torch.optim.Adam(params, lr=1e-5)
        x_h, text_t, mask_t = synth_inputs(Bt, ucfg, 0, pin=False)
        g = torch.Generator().manual_seed(SEED + 5)
        x0, noise = x_h.to(dev), torch.randn(x_h.shape, generator=g).to(dev)
        t = torch.randint(0, T_STEPS, (Bt,), generator=g).to(dev)
        c = {"text": text_t.to(dev), "image": mask_t.to(dev)}
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for k in range(3 + args.train_steps):
            if k == 3:
                torch.cuda.synchronize()
                e0.record()
            noisy = ST.add_noise(tb, x0, noise, t)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                pred = O.unet_forward(leaves, ucfg, noisy, t, c)
                loss = torch.nn.functional.mse_loss(pred, noise)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            torch.nn.utils.clip_grad_norm_(params, 1.0)
            opt.step()
            with torch.no_grad():
                torch._foreach_mul_(ema, 0.999)
                torch._foreach_add_(ema, params, alpha=0.001)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.train_steps
        out["train_bf16_autocast"] = {"value": Bt / (ms / 1e3), "unit": "samples/s", "ms_per_step": ms, "batch": Bt,
                                      "steps": args.train_steps, "loss_finite": bool(torch.isfinite(loss).item())}
        del leaves, params, ema, opt
    O.USE_SDPA = False
    del usd, vsd
    torch.cuda.empty_cache()
    return out


# ---------------------------------------------------------------------------------------------- configs[0]
def config0_leg(args, dev, cpu_budget_s):
    """BASELINE.json configs[0] exactly as BASELINE.md §4 words it: the unconditional LDM UNet of config/celebhq.yaml
    (tools/train_ddpm_vqvae.py:95-107, tools/sample_ddpm_vqvae.py:29-34), batch 4: ONE train step (add_noise -> model ->
    MSE -> backward -> Adam) + the last 50 timesteps of the sampling loop + one VQVAE decode. GPU: this repo's drop-in
    modules (bf16 tensor-core arithmetic; the reference runs this config in fp32). CPU: the oracle port on
    the host cores, timed on a bounded sample (the train step and 3 of the 50 timesteps, decode of 1 of the 4 images),
    extrapolated as stated."""
    import numpy as np
    from models.unet_base import Unet
    from models.vqvae import VQVAE
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    g0 = os.path.join(ROOT, "tests", "golden")
    ucfg = torch.load(os.path.join(g0, "unet_full_uncond.pt"), weights_only=False)["config"]
    vcfg = torch.load(os.path.join(g0, "vqvae_decode.pt"), weights_only=False)["config"]
    B, NS = 4, 50
    import copy
    torch.manual_seed(SEED)
    model = Unet(4, ucfg).train().to(dev)        # tools/train_ddpm_vqvae.py's process
    sampler = copy.deepcopy(model).eval()        # tools/sample_ddpm_vqvae.py's process (loads the checkpoint, eval mode)
    vae = VQVAE(3, vcfg).eval().to(dev)
    sched = LinearNoiseScheduler(T_STEPS, 0.0015, 0.0195)       # config/celebhq.yaml:9-10
    opt = torch.optim.Adam(model.parameters(), lr=5e-6)
    g = torch.Generator().manual_seed(SEED)
    x0, noise = torch.randn(B, 4, 32, 32, generator=g).to(dev), torch.randn(B, 4, 32, 32, generator=g).to(dev)
    t = torch.randint(0, T_STEPS, (B,), generator=g).to(dev)

    def job(timed):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record()
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(model(sched.add_noise(x0, noise, t), t), noise)
        loss.backward()
        opt.step()
        ev[1].record()
        with torch.no_grad():
            xt = noise.clone()
            for i in reversed(range(NS)):
                xt, _ = sched.sample_prev_timestep(xt, sampler(xt, torch.as_tensor(i).unsqueeze(0).to(dev)),
                                                   torch.as_tensor(i))
            ev[2].record()
            ims = vae.decode(xt)
        ev[3].record()
        torch.cuda.synchronize()
        return [ev[k].elapsed_time(ev[k + 1]) for k in range(3)], bool(torch.isfinite(ims).all().item() and
                                                                      torch.isfinite(loss).item())

    job(False)
    job(False)
    (ms_train, ms_samp, ms_dec), finite = job(True)
    gpu_s = (ms_train + ms_samp + ms_dec) / 1e3
    out = {"workload": "configs[0]: unconditional LDM UNet (config/celebhq.yaml), batch 4: 1 train step + 50 DDPM "
                       "timesteps + VQVAE decode", "dtype": "bf16 tensor-core arithmetic, fp32 accumulate / scheduler / optimizer",
           "gpu": {"job_s": gpu_s, "train_step_ms": ms_train, "sampling_50_steps_ms": ms_samp, "decode_ms": ms_dec,
                   "finite": finite}}
    del model, sampler, vae, opt
    torch.cuda.empty_cache()
    if cpu_budget_s > 0:
        from oracle import scheduler_np as onp
        from oracle import unet_torch as O
        torch.set_num_threads(os.cpu_count() or 1)
        torch.manual_seed(SEED)
        usd = {k: v.clone().requires_grad_(v.dtype.is_floating_point) for k, v in Unet(4, ucfg).state_dict().items()}
        vsd = {k: v.clone() for k, v in VQVAE(3, vcfg).state_dict().items()}
        tb = onp.make_tables(T_STEPS, 0.0015, 0.0195)
        x0c, nc, tc = x0.cpu(), noise.cpu(), t.cpu()
        popt = torch.optim.Adam([v for v in usd.values() if v.requires_grad], lr=5e-6)
        t0 = time.perf_counter()
        noisy = torch.from_numpy(onp.add_noise(tb, x0c.numpy(), nc.numpy(), tc.numpy()))
        loss = torch.nn.functional.mse_loss(O.unet_forward(usd, ucfg, noisy, tc), nc)
        loss.backward()
        popt.step()
        cpu_train = time.perf_counter() - t0
        rng = np.random.default_rng(SEED)
        xt = nc.clone()
        t0, done = time.perf_counter(), 0
        with torch.no_grad():
            for i in reversed(range(NS)):
                eps = O.unet_forward(usd, ucfg, xt, torch.full((B,), i, dtype=torch.long))
                xp, _ = onp.sample_prev_timestep(tb, xt.numpy(), eps.numpy(), i, rng.standard_normal(xt.shape, dtype=np.float32))
                xt = torch.from_numpy(np.ascontiguousarray(xp))
                done += 1
                if done >= 3 and time.perf_counter() - t0 > cpu_budget_s / 2:
                    break
            cpu_step = (time.perf_counter() - t0) / done
            t0 = time.perf_counter()
            O.vqvae_decode(vsd, vcfg, xt[:1].clamp(-1, 1))
            cpu_dec = (time.perf_counter() - t0) * B
        cpu_s = cpu_train + NS * cpu_step + cpu_dec
        out["cpu"] = {"job_s": cpu_s, "train_step_s": cpu_train, "sampling_step_s": cpu_step, "decode_s": cpu_dec,
                      "cores": os.cpu_count(), "kind": "port", "dtype": "f32",
                      "sample": f"the train step, {done} of the 50 timesteps (x50/{done}) and the decode of 1 of the 4 images (x4)"}
        out["gpu_over_cpu"] = cpu_s / gpu_s
    return out


# ---------------------------------------------------------------------------------------------- HBM-bound kernels
def hbm_leg(dev, pk):
    """Achieved bytes/s of the HBM-bound kernels against the measured copy bandwidth: algorithmic bytes /
    CUDA-event time, at the workload's own size (launch-latency bound) and at a size that cannot sit in L2."""
    from sdb200 import runtime as rt
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    sched = LinearNoiseScheduler(T_STEPS, 0.00085, 0.012)
    out = {}

    def timed(fn, reps):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps / 1e3

    for tag, nb, reps in (("batch64", 64, 50), ("batch65536", 65536, 10)):
        x, n = torch.randn(nb, 4, 32, 32, device=dev), torch.randn(nb, 4, 32, 32, device=dev)
        t = torch.randint(0, T_STEPS, (nb,), device=dev)
        s = timed(lambda: sched.add_noise(x, n, t), reps)
        out[f"add_noise_{tag}"] = {"us": 1e6 * s, "gbps": 49152.0 * nb / s / 1e9, "frac": 49152.0 * nb / s / 1e9 / pk["hbm_gbs"],
                                   "bytes_per_sample": 49152}
        s = timed(lambda: sched.sample_prev_timestep(x, n, 500), reps)
        out[f"prev_step_{tag}"] = {"us": 1e6 * s, "gbps": 65536.0 * nb / s / 1e9, "frac": 65536.0 * nb / s / 1e9 / pk["hbm_gbs"],
                                   "bytes_per_sample": 65536}
        del x, n, t
    from sdb200.optim import FusedAdam
    p = [torch.randn(118513466, device=dev).requires_grad_()]
    e = [p[0].detach().clone()]
    p[0].grad = torch.randn_like(p[0])
    opt = FusedAdam(p, lr=1e-5, max_grad_norm=1.0, ema_params=e, ema_decay=0.999)
    s = timed(opt.step, 10)
    nbytes = 36.0 * p[0].numel() + 4.0 * p[0].numel()        # Adam+EMA 36 B/param + the norm pass's gradient read
    out["clip_adam_ema_118M"] = {"us": 1e6 * s, "gbps": nbytes / s / 1e9, "frac": nbytes / s / 1e9 / pk["hbm_gbs"],
                                 "bytes_per_param": 40}
    out["peak_gbps"] = pk["hbm_gbs"]
    del p, e, opt
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=T_STEPS)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--cfg", type=float, default=1.0, help="classifier-free guidance scale (>1: two forwards per step)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-budget", type=float, default=90.0, help="seconds the CPU legs may spend")
    ap.add_argument("--e2e-steps", type=int, default=0, help="cap on the e2e leg's timesteps (0 = same as --steps)")
    ap.add_argument("--train-steps", type=int, default=20, help="timed steps of the training leg (0 = skip it)")
    ap.add_argument("--train-batch", type=int, default=16, help="per-GPU training batch (reference: 16)")
    ap.add_argument("--dit-steps", type=int, default=200,
                    help="DiT legs (configs[3], LSQ half of configs[4]): timed sampling steps; 0 disables them")
    ap.add_argument("--train-opt", choices=["fused", "torch"], default="fused",
                    help="optimizer side of the training legs: sdb200.optim.FusedAdam or torch Adam + clip + EMA")
    ap.add_argument("--train-ddp", choices=["ours", "torch"], default="ours",
                    help="N>1 gradient mean: sdb200.parallel.DataParallel (flat buffer) or torch DistributedDataParallel")
    ap.add_argument("--dit-train-batch", type=int, default=28,
                    help="per-GPU batch of the DiT-12L training leg (Model_DiT_12L_config.py:83)")
    ap.add_argument("--full-mask-upload", action="store_true",
                    help="training e2e legs upload the whole (B,18,512,512) mask as the reference loop does")
    ap.add_argument("--cpu-sample", type=int, default=8, help="images per step of the CPU legs (of the 64)")
    ap.add_argument("--train-sweep", action="store_true", help="also time the UNet training step at batch 32 / 64 / 128")
    ap.add_argument("--no-hbm", action="store_true", help="skip the HBM-bound kernel leg")
    ap.add_argument("--no-config0", action="store_true", help="skip the configs[0] leg")
    ap.add_argument("--no-library-bar", action="store_true", help="skip the reference-on-this-GPU leg")
    ap.add_argument("--dump", default="", help="write the per-launch kernel timing of one eager step to this JSON file")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
