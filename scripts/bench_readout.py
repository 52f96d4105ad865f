"""Achieved HBM bandwidth of the bandwidth-bound kernels (readout / selection statistics / resampling / TTA) on BATCHED
launches larger than L2: algorithmic bytes / CUDA-event time, against the measured copy bandwidth in
MEASURED_PEAKS.json.  Single-image launches are L2-resident and launch-bound; their latency is printed as well.
python scripts/bench_readout.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablekeypoints_b200 import ops

dev = "cuda:0"
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peak = 6456.5
try:
    peak = float(json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def timed(fn, n=10):
    for _ in range(3):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


g = torch.Generator(device=dev).manual_seed(0)
rows = []


def report(name, ms, nbytes):
    gbps = nbytes / ms / 1e6
    rows.append({"kernel": name, "ms": ms, "algorithmic_bytes": nbytes, "GB/s": gbps, "frac_of_measured_hbm": gbps / peak})
    print(f"{name:58s} {ms:8.3f} ms  {nbytes / 1e6:9.1f} MB  {gbps:7.0f} GB/s  {100 * gbps / peak:5.1f} % of {peak:.0f}")


# mode A statistics: 64 images x 500 planes x 128^2 = 2.1 GB
maps = torch.rand(64 * 500, 128, 128, device=dev, generator=g)
nb = maps.numel() * 4
report("plane_argmax          (32000 x 128^2)", timed(lambda: ops.plane_argmax(maps)), nb)
report("plane_stats KL        (32000 x 128^2)", timed(lambda: ops.plane_stats(maps, sigma=2.0, want_kl=True)), nb)
report("plane_stats entropy   (32000 x 128^2)", timed(lambda: ops.plane_stats(maps, want_kl=False, want_entropy=True)), nb)
report("find_k_max_pixels k=1 (32000 x 128^2)", timed(lambda: ops.find_k_max_pixels(maps, 1)), nb)
small = maps[:8000].contiguous()
report("bilinear_upsample 128->256 (8000 planes, read+write)", timed(lambda: ops.bilinear_upsample(small[None], 256)),
       small.numel() * 4 * 5)
del small
# mode C readout: 64 images x 10 planes x 512^2 = 0.67 GB
big = torch.rand(640, 512, 512, device=dev, generator=g)
report("plane_argmax          (640 x 512^2)", timed(lambda: ops.plane_argmax(big)), big.numel() * 4)
report("weighted_avg_readout  (640 x 512^2)", timed(lambda: ops.weighted_avg_readout(big)), big.numel() * 4)
del big
# TTA: 64 augmentations x 10 planes (128^2) accumulated into 10 x 512^2 sum / num (RMW = 4 passes of 10.5 MB per augmentation)
fm = torch.rand(64, 10, 128, 128, device=dev, generator=g)
th = torch.eye(2, 3, device=dev)[None].repeat(64, 1, 1)
th[:, 0, 2] = 0.01
s_, n_ = torch.zeros(10, 512, 512, device=dev), torch.zeros(10, 512, 512, device=dev)
# all augmentations are folded inside one launch: the feature planes are read once per augmentation, sum / num get ONE
# read-modify-write (the unfused reference order would be 4 passes of 10.5 MB per augmentation: 2.7 GB)
report("tta_accumulate (64 aug x 10 x 128^2 -> 10 x 512^2, fused)", timed(lambda: ops.tta_accumulate(fm, th, s_, n_)),
       64 * 10 * 128 * 128 * 4 + 4 * 10 * 512 * 512 * 4)
rows[-1]["samples_per_s"] = 64 * 10 * 512 * 512 / (rows[-1]["ms"] * 1e-3)
rows[-1]["note"] = ("gather-bound, not bandwidth-bound: every output pixel does 64 x (inverse warp + 4-tap un-warp, each tap a 4-tap bilinear "
                    "read of an L2-resident 128^2 plane) = 16 reads per (augmentation, plane, pixel); report samples/s, not bytes")
print(f"   = {rows[-1]['samples_per_s'] / 1e9:.2f} G (augmentation, plane, pixel) samples/s")
# losses of the token optimisation, batched over the images of a step (one launch each way): 64 images x 10 selected planes of
# the image and of its transformed copy (2 x 10 x 128^2 x 4 B read; backward reads the same and writes both gradients)
sel = torch.rand(2 * 256, 10, 128, 128, device=dev, generator=g).requires_grad_(True)
thl = torch.eye(2, 3, device=dev)[None].repeat(256, 1, 1)
thl[:, 0, 2] = 0.03
nbl = sel.numel() * 4
report("keypoint_losses_batched fwd (256 x 2 x 10 x 128^2)", timed(lambda: ops.keypoint_losses_batched(sel.detach(), thl, 2.0)), nbl)
out = ops.keypoint_losses_batched(sel, thl, 2.0)
gsum = torch.ones_like(out)
def _bwd():
    sel.grad = None
    out.backward(gsum, retain_graph=True)
report("keypoint_losses_batched bwd (read planes, write both gradients)", timed(_bwd), 2 * nbl)
del sel, out
img = torch.rand(64, 3, 512, 512, device=dev, generator=g)
tha = torch.eye(2, 3, device=dev)[None].repeat(64, 1, 1)
tha[:, 0, 1] = 0.1
report("affine_warp (64 x 3 x 512^2, read + write)", timed(lambda: ops.affine_warp(img, tha)), 2 * img.numel() * 4)
del img
# single-image latencies
one = maps[:500].contiguous()
for name, fn in [("plane_stats KL, one image (500 x 128^2)", lambda: ops.plane_stats(one, sigma=2.0, want_kl=True)),
                 ("plane_argmax, one image (500 x 128^2)", lambda: ops.plane_argmax(one))]:
    print(f"{name:58s} {timed(fn, 50) * 1e3:8.1f} us  (L2-resident, launch-bound)")
os.makedirs(os.path.join(root, "gpurun_out"), exist_ok=True)
json.dump({"peak_gbps": peak, "rows": rows}, open(os.path.join(root, "gpurun_out", "readout_bandwidth.json"), "w"), indent=1)
