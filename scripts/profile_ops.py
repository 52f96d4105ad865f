"""Which ATen ops of the stock-PyTorch trunk own the device time of one token-optimisation step (kineto, op level,
grouped by input shape).  python scripts/profile_ops.py"""
import sys, os
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
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
    step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="self_device_time_total", row_limit=25, max_name_column_width=50))
print(prof.key_averages(group_by_input_shape=True).table(sort_by="self_device_time_total", row_limit=40, max_name_column_width=40, max_shapes_column_width=70))
