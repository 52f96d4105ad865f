"""GPU-busy fraction of one token-optimisation image-iteration batch: sum of kernel time (kineto) vs wall clock.
python scripts/profile_step.py"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from stablekeypoints_b200 import optimize_token, optimize
from stablekeypoints_b200.invertable_transform import RandomAffineWithInverse

dev = "cuda:0"
ldm, controllers, _ = optimize_token.load_ldm(dev, type="sd21-base", feature_upsample_res=128)
ctx = torch.randn(1, 500, 1024, device=dev, requires_grad=True)
img = torch.rand(4, 3, 512, 512, device=dev)
tf = RandomAffineWithInverse(15, (0.8, 1.0), (0.25, 0.25))


def step():
    tr = tf(img)
    sh, eq, _ = optimize.token_opt_losses(ldm, img, tr, ctx, tf, controllers, layers=[0, 1, 2, 3], noise_level=-1, from_where=None,
                                          device=dev, top_k_strategy="gaussian", furthest_point_num_samples=25, top_k=10, sigma=2.0, num_subjects=1)
    (eq * 1000 + sh * 100).backward()


for _ in range(3):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3):
    step()
torch.cuda.synchronize()
wall = (time.perf_counter() - t0) / 3 * 1e3
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
ev = prof.key_averages()
busy = sum(e.self_device_time_total for e in ev) / 3 / 1e3
n = sum(e.count for e in ev) / 3
print(f"wall {wall:.1f} ms/step, kernel time {busy:.1f} ms/step ({100 * busy / wall:.0f} % busy), {n:.0f} kernels/step")
print(ev.table(sort_by="self_device_time_total", row_limit=22, max_name_column_width=70))
