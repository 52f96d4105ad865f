#!/usr/bin/env python
"""bench.py — the StableKeypoints hot path on B200: token-optimisation steps/s (BASELINE.json
configs[1]) or keypoint-extraction images/s (configs[2]), next to the reference algorithm on the
box's host cores.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload optimize|extract] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0 (contract: the task statement's bench section; DESIGN.md §Measurement
explains every key).  `value` = whole-job throughput with the images resident in HBM; `e2e` = the
same job through the public API (`optimize.optimize_embedding` / `eval.run_image_with_context_augmented`)
with the images in pinned host memory, copied in every step, and the step's loss / keypoints read
back; `roofline` = the fused attention-map forward (tcgen05) timed live with CUDA events around the
C-ABI call; `cpu_baseline` = the oracle port of the reference path (`oracle/cpu_pipeline.py`) timed
on the host cores on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

# configs[1] of BASELINE.json == the reference CLI defaults (main.py:74-194) on SD-2.1-base shapes
WL = dict(model="sd21-base", image=512, T=500, R=128, top_k=10, n_cand=25, sigma=2.0, w_sharp=100.0, w_equiv=1000.0,
          degrees=15.0, scale=(0.8, 1.0), translate=(0.25, 0.25), batch_size=4, layers=[0, 1, 2, 3], lr=5e-3,
          n_aug=10, pool=32)
TINY = dict(model="tiny-sd21", image=128, T=48, R=32, top_k=5, n_cand=12, pool=8)   # --tiny: CPU-sized smoke of the harness


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="optimize", choices=["optimize", "extract"])
    ap.add_argument("--images-per-pass", type=int, default=4, help="images per UNet pass on each GPU")
    ap.add_argument("--precision", default="auto", choices=["auto", "tc_bf16x3", "tc_bf16", "simt_f32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=150.0, help="wall-clock cap for the CPU legs")
    ap.add_argument("--tiny", action="store_true", help="tiny model/config (harness smoke, not a bench number)")
    ap.add_argument("--trunk-tf32", action="store_true",
                    help="INFORMATIONAL, not the reference's precision: let the stock UNet/VAE Linear layers use TF32 tensor cores "
                         "(torch.backends.cuda.matmul.allow_tf32).  The default keeps the reference's fp32 matmuls; a line produced "
                         "with this flag says so in `dtype` and `config.trunk_precision`.")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------
# clocks (B200_PROFILING.md "clocks DURING the timed region")
# ----------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.path = gpu_index, None, None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            p = [x.strip() for x in line.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for n, v in zip(names, p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        busy = [s for s in sm if s > 0.5 * max(sm)] if sm else []
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


# ----------------------------------------------------------------------------------------------
# synthetic data
# ----------------------------------------------------------------------------------------------
class ImagePool(torch.utils.data.Dataset):
    """A fixed pool of synthetic images held in ONE tensor: on the GPU (the `value` arm: inputs
    resident in HBM) or in pinned host memory (the `e2e` arm: copied in every step)."""

    def __init__(self, images: torch.Tensor):
        self.images = images

    def __len__(self):
        return self.images.shape[0]

    def __getitem__(self, i):
        return {"img": self.images[i]}


def make_pool(cfg, rank: int) -> torch.Tensor:
    from stablekeypoints_b200.datasets import SyntheticKeypoints
    ds = SyntheticKeypoints(length=cfg["pool"], image_size=cfg["image"], seed=100 + rank)
    return torch.stack([ds[i]["img"] for i in range(len(ds))])


def ncu_dram_bytes(profile: str, kernel: str, images: int):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` from a committed `ncu --set full` summary
    under profiles/ (captured at 8 images per launch; the traffic is proportional to the images).  None if absent."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", profile)
    if not os.path.exists(path):
        return None
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total, inside = 0.0, False
    for ln in open(path):
        if ln.startswith("# kernel:"):
            inside = kernel in ln
        for key in ("dram__bytes_read.sum = ", "dram__bytes_write.sum = "):
            if inside and ln.startswith(key):
                val, unit = ln[len(key):].split()[:2]
                total += float(val) * mult[unit]
    return total * images / 8.0 if total else None


def attn_algorithmic_flops(cfg_model: str, T: int, R: int, n_layers: int = 4) -> float:
    """2 * R^2 * T * sum_l C_l per image over the captured layers."""
    from stablekeypoints_b200.sd_standin import unet_config
    c = unet_config(cfg_model)
    ch = list(c.block_channels)
    # up path visits the attention levels from the deepest: 3 layers per level
    order = []
    for lvl in reversed(range(len(ch))):
        if c.attn_levels[lvl]:
            order += [ch[lvl]] * (c.layers_per_block + 1) * c.transformer_depth[lvl]
    return 2.0 * R * R * T * sum(order[:n_layers])


class StepTimer:
    """barrier + synchronize, CUDA events on the current stream around exactly K steps."""

    def __init__(self, warmup: int, steps: int, on_start=None, on_end=None):
        self.w, self.k = warmup, steps
        self.t0 = torch.cuda.Event(enable_timing=True)
        self.t1 = torch.cuda.Event(enable_timing=True)
        self.on_start, self.on_end = on_start, on_end
        self.ms = None
        if warmup == 0:
            self._begin()

    def _begin(self):
        from stablekeypoints_b200 import distributed as D
        torch.cuda.synchronize()
        D.barrier()
        torch.cuda.synchronize()
        if self.on_start:
            self.on_start()
        self.t0.record()

    def __call__(self, step: int):
        from stablekeypoints_b200 import distributed as D
        if step == self.w:
            self._begin()
        elif step == self.w + self.k:
            self.t1.record()
            torch.cuda.synchronize()
            if self.on_end:
                self.on_end()
            D.barrier()
            self.ms = self.t0.elapsed_time(self.t1)


def max_over_ranks(x: float) -> float:
    from stablekeypoints_b200 import distributed as D
    if not D.is_dist():
        return x
    t = torch.tensor([x], device="cuda", dtype=torch.float64)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())


# ----------------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------------
def setup_ours(args, cfg, local_rank):
    from stablekeypoints_b200 import _lib, optimize_token
    _lib.load()                                  # fail loudly if libskp_b200.so is missing: no fallback exists
    dev = f"cuda:{local_rank}"
    ldm, controllers, _ = optimize_token.load_ldm(dev, type=cfg["model"], feature_upsample_res=cfg["R"],
                                                  images_per_pass=args.images_per_pass, seed=0)
    from stablekeypoints_b200.sd_standin import unet_config
    ctx_dim = unet_config(cfg["model"]).context_dim
    context = torch.randn(1, cfg["T"], ctx_dim, generator=torch.Generator().manual_seed(1)).to(dev)
    return dev, ldm, controllers, context


def run_optimize(args, cfg, dev, ldm, controllers, context, dataset, world, log_fn, timer):
    from stablekeypoints_b200 import optimize
    return optimize.optimize_embedding(
        ldm, top_k_strategy="gaussian", wandb_log=False, context=context, device=dev, num_steps=args.warmup + args.steps,
        layers=cfg["layers"], lr=cfg["lr"], num_tokens=cfg["T"], top_k=cfg["top_k"], augment_degrees=cfg["degrees"],
        augment_scale=cfg["scale"], augment_translate=cfg["translate"], sigma=cfg["sigma"],
        sharpening_loss_weight=cfg["w_sharp"], equivariance_attn_loss_weight=cfg["w_equiv"],
        batch_size=cfg["batch_size"] * world, num_gpus=args.images_per_pass,
        furthest_point_num_samples=cfg["n_cand"], controllers=controllers, dataset=dataset, log_fn=log_fn, step_hook=timer)


def run_extract(args, cfg, dev, ldm, controllers, context, pool, indices, timer, read_back):
    """One step = one image: `n_aug` augmentations in one UNet pass, TTA ensemble, argmax readout."""
    from stablekeypoints_b200 import eval as skp_eval
    out = []
    for step in range(1, args.warmup + args.steps + 1):
        img = pool[(step - 1) % pool.shape[0]]
        if not img.is_cuda:
            img = img.to(dev, non_blocking=True)
        maps = skp_eval.run_image_with_context_augmented(
            ldm, img, context, indices, device=dev, layers=cfg["layers"], augmentation_iterations=cfg["n_aug"],
            augment_degrees=cfg["degrees"], augment_scale=cfg["scale"], augment_translate=cfg["translate"],
            controllers=controllers, num_gpus=cfg["n_aug"], upscale_size=cfg["image"])
        kp = skp_eval.find_max_pixel(maps) / float(cfg["image"])
        out.append(kp.cpu() if read_back else kp)
        timer(step)
    return out


def ours(args):
    from stablekeypoints_b200 import _lib
    from stablekeypoints_b200 import distributed as D
    rank, local_rank, world = D.init_from_env("nccl")
    torch.cuda.set_device(local_rank)
    torch.backends.cuda.matmul.allow_tf32 = bool(args.trunk_tf32)     # default False = the reference's fp32 Linear layers
    cfg = dict(WL, **(TINY if args.tiny else {}))
    dev, ldm, controllers, context = setup_ours(args, cfg, local_rank)
    pool_host = make_pool(cfg, rank).pin_memory()
    pool_dev = pool_host.to(dev)
    peaks = measured_peaks()
    sampler = ClockSampler(local_rank)
    launches = {}

    def on_start():
        launches["k0"], launches["c0"] = _lib.load().skp_kernel_launches(), _lib.launches
        _lib.timing_start()
        if rank == 0:
            sampler.start()

    def on_end():
        launches["timing"] = _lib.timing_summary()
        launches["k1"], launches["c1"] = _lib.load().skp_kernel_launches(), _lib.launches
        if rank == 0:
            launches["clocks"] = sampler.stop()

    # ---- value: images resident in HBM, no per-step host read --------------------------------
    timer = StepTimer(args.warmup, args.steps, on_start, on_end)
    if args.workload == "optimize":
        run_optimize(args, cfg, dev, ldm, controllers, context, ImagePool(pool_dev), world, False, timer)
        units_per_step = 1.0 * world           # batch-4 optimiser-step equivalents per global step (weak scaling)
        metric, unit = "token_opt_steps_per_sec", "steps/s"
    else:
        indices = torch.arange(cfg["top_k"]) * (cfg["T"] // cfg["top_k"])
        run_extract(args, cfg, dev, ldm, controllers, context, pool_dev, indices, timer, False)
        units_per_step = 1.0 * world
        metric, unit = "keypoint_extraction_images_per_sec", "images/s"
    ms = max_over_ranks(timer.ms)
    # ---- e2e: pinned host images copied in every step, loss / keypoints read back -------------
    timer2 = StepTimer(args.warmup, args.steps)
    if args.workload == "optimize":
        losses = []
        run_optimize(args, cfg, dev, ldm, controllers, context, ImagePool(pool_host), world, losses.append, timer2)
        h2d = cfg["batch_size"] * 3 * cfg["image"] ** 2 * 4
        d2h = 3 * 4
        assert all(l["loss"] == l["loss"] for l in losses), "NaN loss in the e2e arm"
    else:
        kps = run_extract(args, cfg, dev, ldm, controllers, context, pool_host, indices, timer2, True)
        h2d = 3 * cfg["image"] ** 2 * 4
        d2h = cfg["top_k"] * 2 * 4
    ms2 = max_over_ranks(timer2.ms)
    if rank != 0:
        return
    value = units_per_step * args.steps / (ms / 1e3)
    e2e = units_per_step * args.steps / (ms2 / 1e3)
    tm = launches["timing"]
    # ---- roofline of the dominant kernel of OUR path: the fused attention-map backward in the token
    # optimisation (it moves 2x the forward's algorithmic FLOPs), the forward in extraction -----------
    imgs_per_call = (2 * args.images_per_pass) if args.workload == "optimize" else cfg["n_aug"]
    flops_fwd = attn_algorithmic_flops(cfg["model"], cfg["T"], cfg["R"], len(cfg["layers"])) * imgs_per_call
    peak = peaks["bf16_tflops_sustained"]

    def roof(symbol, label, flops_call, kernel, profile="r1e_attn_tc_full.txt"):
        t = tm.get(symbol, {"calls": 0, "ms": 0.0})
        ach = flops_call * t["calls"] / (t["ms"] / 1e3) / 1e12 if t["ms"] > 0 else None
        return {"kernel": label, "bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                "frac": (ach / peak) if ach else None, "peak_source": f"{peaks['source']} bf16 dense, sustained",
                "traffic": ncu_dram_bytes(profile, kernel, imgs_per_call), "traffic_source": f"profiles/{profile} (ncu --set full, "
                "dram read+write of the main kernel at 8 images/launch, scaled to this launch's images)",
                "ms_per_launch": t["ms"] / max(1, t["calls"]), "algorithmic_flops_per_launch": flops_call,
                "ms_per_step": t["ms"] / args.steps}

    r_fwd = roof("skp_attn_map_fwd", "skp_attn_map_fwd (native-resolution logit GEMM + attn_map_fwd_tc_kernel, bf16x3 on tcgen05)",
                 flops_fwd, "attn_map_fwd_tc_kernel")
    r_bwd = roof("skp_attn_map_bwd", "skp_attn_map_bwd (attn_map_bwd_tc_kernel: softmax recompute + dz = dS^T W (16^2 layers) or dQ'/dK + bicubic adjoint (32^2 layer) on tcgen05, native-resolution dq/dk GEMMs)",
                 2.0 * flops_fwd, "attn_map_bwd_tc_kernel")
    roofline, roofline_other = (r_bwd, r_fwd) if r_bwd["ms_per_step"] > r_fwd["ms_per_step"] else (r_fwd, r_bwd)
    total_kernel_ms = sum(v["ms"] for v in tm.values())
    kernels = {k: {"calls": v["calls"], "ms_per_step": v["ms"] / args.steps} for k, v in sorted(tm.items())}
    line = {
        "metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": ("f32 (Q.K^T as bf16x3 on tcgen05 with fp32 accumulation; UNet trunk stock PyTorch fp32)" if not args.trunk_tf32 else
                  "f32 kernels (Q.K^T as bf16x3 on tcgen05) + TF32 matmuls in the stock UNet trunk [reduced precision, informational run]"),
        "data": "synthetic", "impl": "ours",
        "config": {"workload": ("configs[1] celeba_wild-shaped token optimisation" if args.workload == "optimize"
                                else "configs[2] extraction sweep (10 augmentations/image, argmax readout)"),
                   "unet": cfg["model"] + " random-init stand-in", "image": cfg["image"], "num_tokens": cfg["T"],
                   "feature_upsample_res": cfg["R"], "top_k": cfg["top_k"], "furthest_point_num_samples": cfg["n_cand"],
                   "batch_size_per_gpu": cfg["batch_size"], "images_per_unet_pass": imgs_per_call,
                   "step": ("one optimiser step = 4 image-iterations per GPU (reference batch_size=4); value counts "
                            "batch-4 step equivalents over all GPUs" if args.workload == "optimize" else "one image"),
                   "l2": "inputs larger than L2: each step streams >1 GB of activations", "tiny": bool(args.tiny),
                   "trunk_precision": "tf32 matmuls (informational)" if args.trunk_tf32 else "fp32 matmuls, TF32 cuDNN convolutions (PyTorch / reference default)"},
        "e2e": {"value": e2e, "unit": unit, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms2 / args.steps},
        "gpu_launches": launches["k1"] - launches["k0"], "abi_calls": launches["c1"] - launches["c0"],
        "roofline": roofline, "roofline_second_kernel": roofline_other, "kernels": kernels, "own_kernel_ms_per_step": total_kernel_ms / args.steps,
        "clocks": launches.get("clocks"),
    }
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_leg(args, cfg, args.cpu_budget_s, max_steps=1, warmup=0)
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# CPU legs: the oracle port of the reference path on the host cores (oracle/cpu_pipeline.py)
# ----------------------------------------------------------------------------------------------
def cpu_leg(args, cfg, budget_s: float, max_steps: int, warmup: int) -> dict:
    """Times image-iterations (optimize) or single-augmentation extractions (extract) of the
    reference algorithm on CPU; converts to the metric's unit.  Bounded: stops at `budget_s`."""
    from oracle import cpu_pipeline as P
    from oracle import keypoint_oracle as O
    from stablekeypoints_b200.datasets import SyntheticKeypoints
    from stablekeypoints_b200.sd_standin import PipelineStandIn, unet_config
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ucfg = unet_config(cfg["model"])
    vae_ch = (32, 32, 64, 64) if args.tiny else (128, 256, 512, 512)
    ldm = PipelineStandIn(ucfg, vae_channels=vae_ch, seed=0)
    for m in (ldm.unet, ldm.vae):
        for p in m.parameters():
            p.requires_grad = False
    store = P.CpuStore()
    P.register_cpu_hook(ldm.unet, store, cfg["R"])
    g = torch.Generator().manual_seed(1)
    context = torch.randn(1, cfg["T"], ucfg.context_dim, generator=g).requires_grad_(True)
    ds = SyntheticKeypoints(length=cfg["pool"], image_size=cfg["image"], seed=100)
    lat = (1, 4, ucfg.latent_side, ucfg.latent_side)
    times, t_begin = [], time.time()
    for i in range(warmup + max_steps):
        if i > 0 and time.time() - t_begin + (times[-1] if times else 0) > budget_s and len(times) > warmup:
            break
        image = ds[i]["img"][None]
        theta = O.affine_theta(10.0, 0.9, 0.1, -0.1)
        t0 = time.perf_counter()
        if args.workload == "optimize":
            context.grad = None
            P.cpu_token_opt_iteration(ldm, store, image, theta, context, torch.randn(lat, generator=g),
                                      torch.randn(lat, generator=g), n_cand=cfg["n_cand"], top_k=cfg["top_k"],
                                      sigma=cfg["sigma"], w_sharp=cfg["w_sharp"], w_equiv=cfg["w_equiv"])
        else:
            idx = torch.arange(cfg["top_k"]) * (cfg["T"] // cfg["top_k"])
            P.cpu_extract_image(ldm, store, image, context.detach(), idx, [theta], [torch.randn(lat, generator=g)],
                                upscale=cfg["image"])
        times.append(time.perf_counter() - t0)
    timed = times[warmup:] if len(times) > warmup else times
    per_sample = sum(timed) / len(timed)
    if args.workload == "optimize":
        value, unit = 1.0 / (per_sample * cfg["batch_size"]), "steps/s"
        sample = (f"{len(timed)} image-iteration(s) (2 UNet fwd+bwd + hook + collect_maps + selection + losses) of the "
                  f"{cfg['batch_size']} per optimiser step, {per_sample:.2f} s each; steps/s = 1/({cfg['batch_size']} x that)")
    else:
        value, unit = 1.0 / (per_sample * cfg["n_aug"]), "images/s"
        sample = (f"{len(timed)} augmentation pass(es) of the {cfg['n_aug']} per image, {per_sample:.2f} s each; "
                  f"images/s = 1/({cfg['n_aug']} x that)")
    return {"value": value, "unit": unit, "cores": cores, "threads": torch.get_num_threads(), "kind": "port",
            "sample": sample, "samples_timed": len(timed), "s_per_sample": per_sample}


def reference(args):
    """--impl reference: the reference algorithm's CPU path on this box's host cores (rank 0 only)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cfg = dict(WL, **(TINY if args.tiny else {}))
    cb = cpu_leg(args, cfg, args.cpu_budget_s, max_steps=max(1, args.steps), warmup=min(args.warmup, 1))
    optimize = args.workload == "optimize"
    line = {
        "metric": "token_opt_steps_per_sec" if optimize else "keypoint_extraction_images_per_sec",
        "value": cb["value"], "unit": cb["unit"], "n_gpus": args.gpus, "steps": cb["samples_timed"], "warmup": min(args.warmup, 1),
        "ms_per_step": 1e3 / cb["value"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "impl": "reference",
        "config": {"workload": ("configs[1] celeba_wild-shaped token optimisation" if optimize
                                else "configs[2] extraction sweep (10 augmentations/image, argmax readout)"),
                   "unet": cfg["model"] + " random-init stand-in", "image": cfg["image"], "num_tokens": cfg["T"],
                   "feature_upsample_res": cfg["R"], "top_k": cfg["top_k"], "furthest_point_num_samples": cfg["n_cand"],
                   "note": "CPU only: the reference (pure PyTorch) cannot travel to the GPU box; this is its algorithm "
                           "restated in oracle/cpu_pipeline.py (pinned to the live reference by tests/golden/e2e_tiny.npz), "
                           "all host threads; one GPU-count-independent measurement", "tiny": bool(args.tiny)},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": cb["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    if args.impl == "reference":
        reference(args)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device. The stablekeypoints_b200 hot path has no CPU fallback; "
                         "use --impl reference for the CPU port of the reference algorithm.")
    ours(args)
    from stablekeypoints_b200 import distributed as D
    if D.is_dist():
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
