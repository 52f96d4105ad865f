#!/usr/bin/env python
"""bench.py — the StableKeypoints hot path on B200 next to the reference algorithm on the box's host cores.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload all|optimize|extract|cub9|sdxl] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0 (contract: the task statement's bench section; DESIGN.md §6 explains every key).

BASELINE.json's metric is "images/sec keypoint extraction & token-opt steps/sec at 1/2/4/8 B200 vs host CPU": the default
invocation (`--workload all`) measures BOTH halves.  The top-level `metric` / `value` / `e2e` / `roofline` / `cpu_baseline`
are the token optimisation (configs[1], weak scaling: batch 4 per GPU); the block `extract` carries the same keys for the
keypoint-extraction sweep (configs[2], images sharded over the ranks, no collective); the block `strong` is the token
optimisation at a FIXED global batch of 8 (the batch of configs[3]) split over the N ranks.  `--workload cub9` times
configs[3] (9 captured layers incl. 64^2, global batch 8 strong-scaled), `--workload sdxl` configs[4] (SDXL stand-in,
1024^2 images, first four 32^2 layers, 16 keypoints).

`value` = whole-job throughput with the images resident in HBM; `e2e` = the same job through the public API
(`optimize.optimize_embedding` / `eval.extract_keypoints`) with the images in pinned host memory, copied in every
step, and the step's losses / keypoints read back; `roofline` = the fused attention-map kernel with the most device time,
timed live with CUDA events around the C-ABI call; `cpu_baseline` = the oracle port of the reference path
(`oracle/cpu_pipeline.py`) timed on the host cores on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

# the reference CLI defaults (main.py:74-194) shared by every workload
BASE = dict(image=512, T=500, R=128, top_k=10, n_cand=25, sigma=2.0, w_sharp=100.0, w_equiv=1000.0, degrees=15.0,
            scale=(0.8, 1.0), translate=(0.25, 0.25), lr=5e-3, n_aug=10, pool=32, max_query_tokens=32 ** 2)
WORKLOADS = {
    # configs[1]: celeba_wild-shaped token optimisation on SD-2.1-base shapes, batch 4 per GPU (weak scaling)
    "optimize": dict(BASE, name="configs[1] celeba_wild-shaped token optimisation", model="sd21-base", layers=[0, 1, 2, 3],
                     batch_size=4, images_per_pass=4),
    # configs[2]: extraction sweep, 10 augmentations per image, argmax readout, images sharded over the ranks
    "extract": dict(BASE, name="configs[2] extraction sweep (10 augmentations/image, argmax readout)", model="sd21-base",
                    layers=[0, 1, 2, 3], images_per_pass=4),
    # configs[3]: CUB-shaped, 9 captured layers incl. 64^2, global batch 8 sharded over the ranks (strong scaling)
    "cub9": dict(BASE, name="configs[3] CUB-shaped token optimisation, 9 captured layers incl. 64^2, global batch 8",
                 model="sd21-base", layers=list(range(9)), global_batch=8, images_per_pass=4, max_query_tokens=64 ** 2),
    # configs[4]: SDXL stand-in, 1024^2 images (128^2 latent), first four 32^2 up-path cross-attention layers, 16 keypoints
    "sdxl": dict(BASE, name="configs[4] SDXL monkey-patch path, 1024^2 images, 16 keypoints", model="sdxl", image=1024,
                 layers=[0, 1, 2, 3], top_k=16, batch_size=2, images_per_pass=2, pool=8, n_aug=10),
}
TINY = dict(image=128, T=48, R=32, top_k=5, n_cand=12, pool=8)   # --tiny: CPU-sized smoke of the harness
TOLERANCE = {
    "maps_forward": "rtol 1e-3 vs fp64 oracle (bf16x3 split products, fp32 accumulation): tests/test_full_size_parity_gpu.py",
    "row_stats": "2e-4 abs on m + log2(l) (log2 units)",
    "token_ids_argmax_fps": "bit-exact (rank ties only where fp64 scores differ < 3e-6 rel)",
    "backward_dq_dk": "<= 1.5e-2 of each gradient's max vs fp64 autograd, measured 3e-3 .. 7e-3 (gradient products dS^T.W / dS.K / "
                      "dS^T.Q' take single-bf16 operands, fp32 accumulation in TMEM; window sums of the logit gradient are added "
                      "in bf16 at L2; S recompute is bf16x3)",
    "context_gradient_e2e": "<= 1e-2 of max vs the live reference (tests/test_e2e_gpu.py)",
    "soft_argmax_keypoints": "0.5 px at 512^2 (measured 1e-3)",
}
DTYPE = ("f32 I/O and accumulation; forward Q.K^T as bf16x3 split products on tcgen05 (fp32-grade, rtol 1e-3); backward: logit "
         "recompute bf16x3, gradient products single bf16 with fp32 accumulation, the logit gradient's window sums in bf16 (<= 1.5e-2 "
         "of max, measured <= 7e-3); UNet/VAE trunk stock PyTorch fp32")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="all", choices=["all", "optimize", "extract", "cub9", "sdxl"])
    ap.add_argument("--model", default="", help="stand-in architecture instead of the workload's (e.g. sd15: what the reference loads, "
                                                 "8 heads of dim 160 / 80; BASELINE names sd21-base)")
    ap.add_argument("--images-per-pass", type=int, default=0, help="images per UNet pass on each GPU (0: the workload's default)")
    ap.add_argument("--precision", default="auto", choices=["auto", "tc_bf16x3", "tc_bf16", "simt_f32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=400.0, help="wall-clock cap for the --impl reference run")
    ap.add_argument("--tiny", action="store_true", help="tiny model/config (harness smoke, not a bench number)")
    ap.add_argument("--channels-last", action="store_true",
                    help="A/B of the layout lever: store the trunk's convolution weights in NHWC (measured SLOWER on B200 with fp32 "
                         "activations: 181.6 vs 176.3 ms per optimisation step, 127 vs 117 ms per extracted image; off by default)")
    ap.add_argument("--trunk-tf32", action="store_true",
                    help="INFORMATIONAL, not the reference's precision: let the stock UNet/VAE Linear layers use TF32 tensor cores "
                         "(torch.backends.cuda.matmul.allow_tf32).  The default keeps the reference's fp32 matmuls; a line produced "
                         "with this flag says so in `dtype` and `config.trunk_precision`.")
    return ap.parse_args()


def workload_cfg(name: str, args) -> dict:
    cfg = dict(WORKLOADS[name])
    if args.tiny:
        cfg.update(TINY, model="tiny-sd21")
        if name == "cub9":
            cfg["max_query_tokens"] = 16 ** 2
    if args.images_per_pass:
        cfg["images_per_pass"] = args.images_per_pass
    if args.model and not args.tiny:
        cfg["model"] = args.model
    return cfg


def config_block(cfg: dict, args) -> dict:
    """The `config` object of the JSON line: identical for the `ours` and the `reference` arm of a workload."""
    return {"workload": cfg["name"], "unet": cfg["model"] + " random-init stand-in", "image": cfg["image"], "num_tokens": cfg["T"],
            "feature_upsample_res": cfg["R"], "top_k": cfg["top_k"], "furthest_point_num_samples": cfg["n_cand"],
            "captured_layers": len(cfg["layers"]), "augmentations_per_image": cfg["n_aug"],
            "batch": (f"global batch {cfg['global_batch']} split over the ranks (strong scaling)" if "global_batch" in cfg
                      else f"batch {cfg.get('batch_size', cfg['images_per_pass'])} per GPU (weak scaling)"),
            "l2": "inputs larger than L2: each step streams >1 GB of activations", "tiny": bool(args.tiny),
            "trunk_precision": "tf32 matmuls (informational)" if args.trunk_tf32 else "fp32 matmuls, TF32 cuDNN convolutions (PyTorch / reference default)"}


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

    def __init__(self, images: torch.Tensor, world: int = 1):
        # every rank holds its own pool (seed 100 + rank) and `optimize_embedding` shards whatever dataset it is given by
        # rank: present the pool `world` times so that this rank's shard is the whole pool at any N
        self.images, self.world = images, world

    def __len__(self):
        return self.images.shape[0] * self.world

    def __getitem__(self, i):
        return {"img": self.images[i % self.images.shape[0]]}


def make_pool(cfg, rank: int) -> torch.Tensor:
    from stablekeypoints_b200.datasets import SyntheticKeypoints
    ds = SyntheticKeypoints(length=cfg["pool"], image_size=cfg["image"], seed=100 + rank)
    return torch.stack([ds[i]["img"] for i in range(len(ds))])


def ncu_dram_bytes(kernel: str, images: int):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` from the newest committed `ncu --set full`
    summary under profiles/ (captured at 8 images per launch; the traffic is proportional to the images).  None if absent."""
    pdir = os.path.join(ROOT, "profiles")
    for profile in ("r2_attn_tc_full.txt", "r1e_attn_tc_full.txt"):
        path = os.path.join(pdir, profile)
        if not os.path.exists(path):
            continue
        mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        total, inside = 0.0, False
        for ln in open(path):
            if ln.startswith("# kernel:"):
                inside = kernel in ln
            for key in ("dram__bytes_read.sum = ", "dram__bytes_write.sum = "):
                if inside and ln.startswith(key):
                    val, unit = ln[len(key):].split()[:2]
                    total += float(val) * mult[unit]
        if total:
            return total * images / 8.0, profile
    return None, None


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


def attn_exponentials(cfg_model: str, T: int, R: int, n_layers: int = 4) -> float:
    """exponentials per image-forward: R^2 * T * sum_l heads_l."""
    from stablekeypoints_b200.sd_standin import unet_config
    c = unet_config(cfg_model)
    order = []
    for lvl in reversed(range(len(c.block_channels))):
        if c.attn_levels[lvl]:
            order += [c.heads[lvl]] * (c.layers_per_block + 1) * c.transformer_depth[lvl]
    return float(R) * R * T * sum(order[:n_layers])


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
def setup_pipeline(args, cfg, local_rank):
    """(device, ldm, controllers, context) for a workload.  configs[4] goes through the SDXL controller API
    (`sdxl_monkey_patch.register_attention_control` + its keyed AttentionStore); everything else through `load_ldm`."""
    from stablekeypoints_b200 import _lib, optimize_token
    from stablekeypoints_b200.sd_standin import PipelineStandIn, _ModuleWrap, unet_config
    _lib.load()                                  # fail loudly if libskp_b200.so is missing: no fallback exists
    dev = f"cuda:{local_rank}"
    ucfg = unet_config(cfg["model"])
    vae_ch = (32, 32, 64, 64) if args.tiny else (128, 256, 512, 512)
    if cfg["model"] == "sdxl":
        from stablekeypoints_b200 import sdxl_monkey_patch
        ldm = PipelineStandIn(ucfg, vae_channels=vae_ch, seed=0).to(dev)
        store = sdxl_monkey_patch.AttentionStore(max_query_tokens=cfg["max_query_tokens"], max_captured_layers=len(cfg["layers"]))
        store.stop_after_capture = True
        sdxl_monkey_patch.register_attention_control(ldm, store, feature_upsample_res=cfg["R"])
        ldm.unet = _ModuleWrap(ldm.unet, [local_rank])
        ldm.vae = _ModuleWrap(ldm.vae, [local_rank])
        for m in (ldm.vae, ldm.text_encoder, ldm.unet):
            for p in m.parameters():
                p.requires_grad = False
        controllers = {torch.device(dev): store}
    else:
        pipe = PipelineStandIn(ucfg, vae_channels=vae_ch, seed=0).to(dev) if args.tiny else None
        ldm, controllers, _ = optimize_token.load_ldm(dev, type=cfg["model"], feature_upsample_res=cfg["R"],
                                                      images_per_pass=cfg["images_per_pass"], seed=0, ldm=pipe,
                                                      max_query_tokens=cfg["max_query_tokens"],
                                                      max_captured_layers=len(cfg["layers"]))
    if args.channels_last:
        optimize_token.to_channels_last(ldm)
    context = torch.randn(1, cfg["T"], ucfg.context_dim, generator=torch.Generator().manual_seed(1)).to(dev)
    return dev, ldm, controllers, context


def run_optimize(args, cfg, dev, ldm, controllers, context, dataset, global_batch, per_pass, log_fn, timer):
    from stablekeypoints_b200 import optimize
    return optimize.optimize_embedding(
        ldm, top_k_strategy="gaussian", wandb_log=False, context=context, device=dev, num_steps=args.warmup + args.steps,
        layers=cfg["layers"], lr=cfg["lr"], num_tokens=cfg["T"], top_k=cfg["top_k"], augment_degrees=cfg["degrees"],
        augment_scale=cfg["scale"], augment_translate=cfg["translate"], sigma=cfg["sigma"],
        sharpening_loss_weight=cfg["w_sharp"], equivariance_attn_loss_weight=cfg["w_equiv"],
        batch_size=global_batch, num_gpus=per_pass, furthest_point_num_samples=cfg["n_cand"], controllers=controllers,
        dataset=dataset, log_fn=log_fn, step_hook=timer)


def run_extract(args, cfg, dev, ldm, controllers, context, pool, indices, timer, read_back):
    """One step = `images_per_pass` images: their n_aug augmentations in ONE UNet pass, TTA ensemble, argmax readout
    (`eval.extract_keypoints`, the body of `evaluate` / `precompute_all_keypoints`)."""
    from stablekeypoints_b200 import eval as skp_eval
    ipp = cfg["images_per_pass"]
    out = []
    for step in range(1, args.warmup + args.steps + 1):
        ids = [((step - 1) * ipp + j) % pool.shape[0] for j in range(ipp)]
        samples = [{"img": pool[i]} for i in ids]
        kp = skp_eval.extract_keypoints(
            ldm, samples, context, indices, device=dev, from_where=None, layers=cfg["layers"], noise_level=-1,
            augmentation_iterations=cfg["n_aug"], augment_degrees=cfg["degrees"], augment_scale=cfg["scale"],
            augment_translate=cfg["translate"], controllers=controllers, num_gpus=cfg["n_aug"], upscale_size=cfg["image"])
        out.append(kp.cpu() if read_back else kp)
        timer(step)
    return out


class Probe:
    """launch counters + per-entry-point CUDA-event timing + clocks around one timed region"""

    def __init__(self, rank, local_rank):
        from stablekeypoints_b200 import _lib
        self._lib, self.rank, self.sampler, self.rec = _lib, rank, ClockSampler(local_rank), {}

    def on_start(self):
        self.rec["k0"], self.rec["c0"] = self._lib.load().skp_kernel_launches(), self._lib.launches
        self._lib.timing_start()
        if self.rank == 0:
            self.sampler.start()

    def on_end(self):
        self.rec["timing"] = self._lib.timing_summary()
        self.rec["k1"], self.rec["c1"] = self._lib.load().skp_kernel_launches(), self._lib.launches
        if self.rank == 0:
            self.rec["clocks"] = self.sampler.stop()


def rooflines(cfg, tm, steps, imgs_per_call, peaks, backward: bool):
    """(dominant, other) roofline objects of the fused attention-map entry points, from the live CUDA-event timing."""
    flops_fwd = attn_algorithmic_flops(cfg["model"], cfg["T"], cfg["R"], len(cfg["layers"])) * imgs_per_call
    exps = attn_exponentials(cfg["model"], cfg["T"], cfg["R"], len(cfg["layers"])) * imgs_per_call
    peak = peaks["bf16_tflops_sustained"]
    mufu_floor_ms = exps / (16.0 * 148 * 1.965e9) * 1e3     # MUFU ex2: 16 / clk / SM, 148 SMs, 1965 MHz

    def roof(symbol, label, flops_call, kernel, passes):
        # the optimisation calls the backward through its operand-reusing entry point (same kernels, ABI 107)
        t = tm.get(symbol + "_reuse") or tm.get(symbol, {"calls": 0, "ms": 0.0})
        ach = flops_call * t["calls"] / (t["ms"] / 1e3) / 1e12 if t["ms"] > 0 else None
        traffic, prof = ncu_dram_bytes(kernel, imgs_per_call)
        return {"kernel": label, "bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                "frac": (ach / peak) if ach else None, "peak_source": f"{peaks['source']} bf16 dense, sustained",
                "ceilings": {"bf16x3": 1.0 / 3.0, "mufu_ex2": (flops_call / (passes * mufu_floor_ms * 1e-3) / 1e12 / peak),
                             "note": "fractions of `peak` this formulation cannot exceed: three bf16 products per fp32-grade logit; "
                                     f"{passes} pass(es) over the exponentials at 16 ex2/clk/SM ({passes * mufu_floor_ms:.2f} ms per launch)"},
                "traffic": traffic, "traffic_source": (f"profiles/{prof} (ncu --set full, dram read+write of the main kernel at 8 images/"
                                                       "launch, scaled to this launch's images)") if prof else None,
                "ms_per_launch": t["ms"] / max(1, t["calls"]), "algorithmic_flops_per_launch": flops_call,
                "ms_per_step": t["ms"] / steps}

    r_fwd = roof("skp_attn_map_fwd", "skp_attn_map_fwd (native-resolution logit GEMM + attn_map_fwd_tc_kernel, bf16x3 on tcgen05)",
                 flops_fwd, "attn_map_fwd_tc_kernel", 1)
    if not backward:
        return r_fwd, None
    r_bwd = roof("skp_attn_map_bwd", "skp_attn_map_bwd[_reuse] (attn_map_bwd_tc_kernel: softmax recompute + dz = dS^T W on tcgen05, "
                 "native-resolution dq/dk GEMMs; logits / interpolation tables reused from the forward's workspace)", 2.0 * flops_fwd, "attn_map_bwd_tc_kernel", 1)
    return (r_bwd, r_fwd) if r_bwd["ms_per_step"] > r_fwd["ms_per_step"] else (r_fwd, r_bwd)


def measure_optimize(args, cfg, env, rank, local_rank, world, pools, *, global_batch, per_pass, e2e=True):
    """Token optimisation: `value` arm (+ `e2e` arm).  -> dict of raw measurements (rank 0 uses them)."""
    dev, ldm, controllers, context = env
    probe = Probe(rank, local_rank)
    timer = StepTimer(args.warmup, args.steps, probe.on_start, probe.on_end)
    run_optimize(args, cfg, dev, ldm, controllers, context, ImagePool(pools["dev"], world), global_batch, per_pass, False, timer)
    out = {"ms": max_over_ranks(timer.ms), "probe": probe.rec}
    if e2e:
        timer2 = StepTimer(args.warmup, args.steps)
        losses = []
        run_optimize(args, cfg, dev, ldm, controllers, context, ImagePool(pools["host"], world), global_batch, per_pass, losses.append, timer2)
        assert all(l["loss"] == l["loss"] for l in losses), "NaN loss in the e2e arm"
        out["ms_e2e"] = max_over_ranks(timer2.ms)
        out["h2d"] = (global_batch // world) * 3 * cfg["image"] ** 2 * 4
        out["d2h"] = 3 * 4
    return out


def measure_extract(args, cfg, env, rank, local_rank, world, pools):
    dev, ldm, controllers, context = env
    indices = torch.arange(cfg["top_k"]) * (cfg["T"] // cfg["top_k"])
    probe = Probe(rank, local_rank)
    timer = StepTimer(args.warmup, args.steps, probe.on_start, probe.on_end)
    run_extract(args, cfg, dev, ldm, controllers, context, pools["dev"], indices, timer, False)
    timer2 = StepTimer(args.warmup, args.steps)
    run_extract(args, cfg, dev, ldm, controllers, context, pools["host"], indices, timer2, True)
    return {"ms": max_over_ranks(timer.ms), "ms_e2e": max_over_ranks(timer2.ms), "probe": probe.rec,
            "h2d": cfg["images_per_pass"] * 3 * cfg["image"] ** 2 * 4, "d2h": cfg["images_per_pass"] * cfg["top_k"] * 2 * 4}


def block_from(args, cfg, m, units_per_step, metric, unit, world, peaks, imgs_per_call, backward, scaling):
    """JSON block of one workload from its raw measurements."""
    tm = m["probe"]["timing"]
    roofline, other = rooflines(cfg, tm, args.steps, imgs_per_call, peaks, backward)
    blk = {"metric": metric, "value": units_per_step * args.steps / (m["ms"] / 1e3), "unit": unit, "n_gpus": world,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": m["ms"] / args.steps, "higher_is_better": True,
           "scaling": scaling, "config": config_block(cfg, args),
           "gpu_launches": m["probe"]["k1"] - m["probe"]["k0"], "abi_calls": m["probe"]["c1"] - m["probe"]["c0"],
           "roofline": roofline, "kernels": {k: {"calls": v["calls"], "ms_per_step": v["ms"] / args.steps} for k, v in sorted(tm.items())},
           "own_kernel_ms_per_step": sum(v["ms"] for v in tm.values()) / args.steps, "clocks": m["probe"].get("clocks")}
    if other is not None:
        blk["roofline_second_kernel"] = other
    if "ms_e2e" in m:
        blk["e2e"] = {"value": units_per_step * args.steps / (m["ms_e2e"] / 1e3), "unit": unit, "h2d_bytes_per_step": m["h2d"],
                      "d2h_bytes_per_step": m["d2h"], "ms_per_step": m["ms_e2e"] / args.steps}
    return blk


def ours(args):
    from stablekeypoints_b200 import distributed as D
    rank, local_rank, world = D.init_from_env("nccl")
    torch.cuda.set_device(local_rank)
    torch.backends.cuda.matmul.allow_tf32 = bool(args.trunk_tf32)     # default False = the reference's fp32 Linear layers
    peaks = measured_peaks()
    main_name = "optimize" if args.workload == "all" else args.workload
    cfg = workload_cfg(main_name, args)
    env = setup_pipeline(args, cfg, local_rank)
    pool_host = make_pool(cfg, rank).pin_memory()
    pools = {"host": pool_host, "dev": pool_host.to(env[0])}
    extra = {}
    if main_name in ("optimize", "sdxl"):
        gb, pp = cfg["batch_size"] * world, cfg["images_per_pass"]
        m = measure_optimize(args, cfg, env, rank, local_rank, world, pools, global_batch=gb, per_pass=pp)
        # one step = one optimiser step of `batch_size` image-iterations on every GPU; value counts batch-`batch_size` step
        # equivalents over all GPUs (weak scaling)
        line_args = dict(units_per_step=1.0 * world, metric="token_opt_steps_per_sec", unit="steps/s", imgs_per_call=2 * pp,
                         backward=True, scaling="weak")
        step_def = (f"one optimiser step = {cfg['batch_size']} image-iterations per GPU (reference batch_size={cfg['batch_size']}); "
                    "value counts such step equivalents over all GPUs")
    elif main_name == "cub9":
        gb = cfg["global_batch"]
        if gb % world:
            raise SystemExit(f"cub9: global batch {gb} is not divisible by {world} ranks")
        pp = min(cfg["images_per_pass"], gb // world)
        m = measure_optimize(args, cfg, env, rank, local_rank, world, pools, global_batch=gb, per_pass=pp)
        line_args = dict(units_per_step=1.0, metric="token_opt_steps_per_sec", unit="steps/s", imgs_per_call=2 * pp,
                         backward=True, scaling="strong")
        step_def = f"one optimiser step over the global batch of {gb} images ({gb // world} per GPU, {pp} per UNet pass)"
    else:
        m = measure_extract(args, cfg, env, rank, local_rank, world, pools)
        line_args = dict(units_per_step=float(cfg["images_per_pass"] * world), metric="keypoint_extraction_images_per_sec",
                         unit="images/s", imgs_per_call=cfg["images_per_pass"] * cfg["n_aug"], backward=False, scaling="weak")
        step_def = f"one step = {cfg['images_per_pass']} images per GPU ({cfg['n_aug']} augmentations each in one UNet pass)"
    if args.workload == "all":
        # ---- strong scaling of the token optimisation: global batch 8 (configs[3]'s batch) split over the ranks
        gb = 8
        if gb % world == 0:
            pp = min(cfg["images_per_pass"], gb // world)
            extra["strong_raw"] = measure_optimize(args, cfg, env, rank, local_rank, world, pools, global_batch=gb, per_pass=pp, e2e=False)
            extra["strong_pp"] = pp
        # ---- the other half of the metric: keypoint extraction, images sharded over the ranks, no collective
        xcfg = workload_cfg("extract", args)
        extra["extract_cfg"] = xcfg
        extra["extract_raw"] = measure_extract(args, xcfg, env, rank, local_rank, world, pools)
    elif main_name == "sdxl":
        xcfg = dict(cfg, name=cfg["name"] + " — extraction")
        extra["extract_cfg"] = xcfg
        extra["extract_raw"] = measure_extract(args, xcfg, env, rank, local_rank, world, pools)
    if rank != 0:
        return
    line = block_from(args, cfg, m, world=world, peaks=peaks, **line_args)
    line.update({"vs_baseline": None, "dtype": DTYPE if not args.trunk_tf32 else DTYPE + " + TF32 matmuls in the trunk [reduced precision, informational run]",
                 "tolerance": TOLERANCE, "data": "synthetic", "impl": "ours", "step_definition": step_def,
                 "trunk_layout": "channels_last convolutions (A/B run)" if args.channels_last else "NCHW (PyTorch / reference default)"})
    if "strong_raw" in extra:
        s = extra["strong_raw"]
        line["strong"] = {"metric": "token_opt_steps_per_sec", "global_batch": 8, "images_per_gpu_per_step": 8 // world,
                          "images_per_unet_pass": 2 * extra["strong_pp"], "value": args.steps / (s["ms"] / 1e3), "unit": "steps/s",
                          "ms_per_step": s["ms"] / args.steps, "scaling": "strong",
                          "limit": "the stock fp32 trunk at a shrinking per-GPU batch (own kernels and the 2 MB NCCL all-reduce are < 6 % of a step)",
                          "own_kernel_ms_per_step": sum(v["ms"] for v in s["probe"]["timing"].values()) / args.steps}
    if "extract_raw" in extra:
        xcfg = extra["extract_cfg"]
        line["extract"] = block_from(args, xcfg, extra["extract_raw"], units_per_step=float(xcfg["images_per_pass"] * world),
                                     metric="keypoint_extraction_images_per_sec", unit="images/s", world=world, peaks=peaks,
                                     imgs_per_call=xcfg["images_per_pass"] * xcfg["n_aug"], backward=False, scaling="weak")
        line["extract"]["step_definition"] = (f"one step = {xcfg['images_per_pass']} images per GPU ({xcfg['n_aug']} augmentations each, "
                                              "all in one UNet pass)")
    if not args.no_cpu_baseline and world == 1 and main_name != "sdxl":
        kind = "extract" if main_name == "extract" else "optimize"
        line["cpu_baseline"] = cpu_leg(args, cfg, kind, n_samples=1, warmup=0)
        if "extract" in line and args.workload == "all":
            line["extract"]["cpu_baseline"] = cpu_leg(args, extra["extract_cfg"], "extract", n_samples=1, warmup=0)
    elif main_name == "sdxl":
        line["cpu_baseline"] = None
        line["cpu_baseline_note"] = "not timed: one SDXL image-iteration at 1024^2 on host cores takes minutes; configs[1] carries the CPU baseline"
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# CPU legs: the oracle port of the reference path on the host cores (oracle/cpu_pipeline.py)
# ----------------------------------------------------------------------------------------------
def cpu_leg(args, cfg, kind: str, n_samples: int, warmup: int, budget_s: float = 1e9) -> dict:
    """Times `warmup + n_samples` image-iterations (optimize) or single-augmentation extraction passes (extract) of the
    reference algorithm on CPU and converts to the metric's unit.  Stops early only when `budget_s` is exceeded."""
    from oracle import cpu_pipeline as P
    from oracle import keypoint_oracle as O
    from stablekeypoints_b200.datasets import SyntheticKeypoints
    from stablekeypoints_b200.sd_standin import PipelineStandIn, unet_config
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ucfg = unet_config(cfg["model"])
    vae_ch = (32, 32, 64, 64) if args.tiny else (128, 256, 512, 512)
    ldm = PipelineStandIn(ucfg, vae_channels=vae_ch, seed=0)
    for mod in (ldm.unet, ldm.vae):
        for p in mod.parameters():
            p.requires_grad = False
    store = P.CpuStore(max_query_tokens=cfg["max_query_tokens"], max_layers=len(cfg["layers"]))
    P.register_cpu_hook(ldm.unet, store, cfg["R"])
    g = torch.Generator().manual_seed(1)
    context = torch.randn(1, cfg["T"], ucfg.context_dim, generator=g).requires_grad_(True)
    ds = SyntheticKeypoints(length=cfg["pool"], image_size=cfg["image"], seed=100)
    lat = (1, 4, cfg["image"] // 8, cfg["image"] // 8)
    layers = tuple(cfg["layers"])
    times, t_begin = [], time.time()
    for i in range(warmup + n_samples):
        if len(times) > warmup and time.time() - t_begin + times[-1] > budget_s:
            break
        image = ds[i % len(ds)]["img"][None]
        theta = O.affine_theta(10.0, 0.9, 0.1, -0.1)
        t0 = time.perf_counter()
        if kind == "optimize":
            context.grad = None
            P.cpu_token_opt_iteration(ldm, store, image, theta, context, torch.randn(lat, generator=g),
                                      torch.randn(lat, generator=g), n_cand=cfg["n_cand"], top_k=cfg["top_k"],
                                      sigma=cfg["sigma"], w_sharp=cfg["w_sharp"], w_equiv=cfg["w_equiv"], layers=layers)
        else:
            idx = torch.arange(cfg["top_k"]) * (cfg["T"] // cfg["top_k"])
            P.cpu_extract_image(ldm, store, image, context.detach(), idx, [theta], [torch.randn(lat, generator=g)],
                                upscale=cfg["image"], layers=layers)
        times.append(time.perf_counter() - t0)
    timed = times[warmup:] if len(times) > warmup else times
    per_sample = sum(timed) / len(timed)
    if kind == "optimize":
        per_step = cfg.get("global_batch", cfg.get("batch_size", 4))
        value, unit = 1.0 / (per_sample * per_step), "steps/s"
        sample = (f"{len(timed)} image-iteration(s) (2 UNet fwd+bwd + hook + collect_maps + selection + losses) of the "
                  f"{per_step} per optimiser step, {per_sample:.2f} s each; steps/s = 1/({per_step} x that)")
    else:
        value, unit = 1.0 / (per_sample * cfg["n_aug"]), "images/s"
        sample = (f"{len(timed)} augmentation pass(es) of the {cfg['n_aug']} per image, {per_sample:.2f} s each; "
                  f"images/s = 1/({cfg['n_aug']} x that)")
    return {"value": value, "unit": unit, "cores": cores, "threads": torch.get_num_threads(), "kind": "port",
            "sample": sample, "samples_timed": len(timed), "warmup_samples": min(warmup, len(times) - len(timed)),
            "s_per_sample": per_sample}


def reference(args):
    """--impl reference: the reference algorithm's CPU path on this box's host cores (rank 0 only).

    One timed "step" = one bounded sample of the workload: ONE image-iteration of the token optimisation (the smallest unit
    the algorithm can be cut into: two UNet forward+backward passes at 512^2), i.e. 1/batch_size of an optimiser step;
    `steps` / `warmup` are the K / W asked for, `ms_per_step` the time of one such sample, and `value` converts to the
    metric's unit (steps/s = 1 / (batch_size x seconds per image-iteration))."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    main_name = "optimize" if args.workload == "all" else args.workload
    cfg = workload_cfg(main_name, args)
    kind = "extract" if main_name == "extract" else "optimize"
    if main_name == "sdxl":
        print(json.dumps({"impl": "reference", "unavailable": "configs[4]: one SDXL image-iteration at 1024^2 on host cores takes minutes; "
                          "the CPU arm is timed on configs[1]-[3]"}), flush=True)
        return
    cb = cpu_leg(args, cfg, kind, n_samples=max(1, args.steps), warmup=max(0, args.warmup), budget_s=args.cpu_budget_s)
    line = {
        "metric": "keypoint_extraction_images_per_sec" if kind == "extract" else "token_opt_steps_per_sec",
        "value": cb["value"], "unit": cb["unit"], "n_gpus": args.gpus, "steps": cb["samples_timed"], "warmup": cb["warmup_samples"],
        "ms_per_step": 1e3 * cb["s_per_sample"], "higher_is_better": True,
        "scaling": "strong" if main_name == "cub9" else "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "impl": "reference", "config": config_block(cfg, args),
        "step_definition": ("one timed step = one bounded sample = ONE image-iteration on all host cores (1/batch of an optimiser step); "
                            "value = 1 / (batch x s per sample)" if kind == "optimize" else
                            "one timed step = ONE augmentation pass (1/n_aug of an image); value = 1 / (n_aug x s per sample)"),
        "note": "CPU only: the reference (pure PyTorch + diffusers 0.8.0) cannot travel to the GPU box; this is its algorithm restated in "
                "oracle/cpu_pipeline.py (pinned to the live reference by tests/golden/e2e_tiny.npz), all host threads; one "
                "GPU-count-independent measurement",
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": cb["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if args.workload == "all":
        xcfg = workload_cfg("extract", args)
        xb = cpu_leg(args, xcfg, "extract", n_samples=max(1, min(args.steps, 5)), warmup=min(1, args.warmup), budget_s=60.0)
        line["extract"] = {"metric": "keypoint_extraction_images_per_sec", "value": xb["value"], "unit": xb["unit"],
                           "config": config_block(xcfg, args), "cpu_baseline": xb}
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
