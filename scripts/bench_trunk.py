"""Where does the stock-PyTorch trunk time go?  Times one token-opt image-iteration batch (8 images) under
different stock-PyTorch settings.  python scripts/bench_trunk.py"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablekeypoints_b200 import optimize_token, optimize
from stablekeypoints_b200.invertable_transform import RandomAffineWithInverse

dev = "cuda:0"


def step_time(ldm, controllers, n=4, B=4):
    ctx = torch.randn(1, 500, 1024, device=dev, requires_grad=True)
    img = torch.rand(B, 3, 512, 512, device=dev)
    tf = RandomAffineWithInverse(15, (0.8, 1.0), (0.25, 0.25))
    ts = []
    for i in range(n + 2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        tr = tf(img)
        sh, eq, _ = optimize.token_opt_losses(ldm, img, tr, ctx, tf, controllers, layers=[0, 1, 2, 3], noise_level=-1, from_where=None,
                                              device=dev, top_k_strategy="gaussian", furthest_point_num_samples=25, top_k=10, sigma=2.0, num_subjects=1)
        (eq * 1000 + sh * 100).backward()
        torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    return min(ts[2:]) * 1e3


def vae_time(ldm, B=8):
    img = torch.rand(B, 3, 512, 512, device=dev)
    with torch.no_grad():
        for _ in range(2):
            ldm.vae.encode(img)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ldm.vae.encode(img)
        torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3


ldm, controllers, _ = optimize_token.load_ldm(dev, type="sd21-base", feature_upsample_res=128)
print("default: step %.1f ms, vae encode(8) %.1f ms" % (step_time(ldm, controllers), vae_time(ldm)))
torch.backends.cudnn.benchmark = True
print("cudnn.benchmark: step %.1f ms, vae %.1f ms" % (step_time(ldm, controllers), vae_time(ldm)))
ldm.unet.module.to(memory_format=torch.channels_last); ldm.vae.module.to(memory_format=torch.channels_last)
print("+channels_last: step %.1f ms, vae %.1f ms" % (step_time(ldm, controllers), vae_time(ldm)))
torch.backends.cuda.matmul.allow_tf32 = True
print("+tf32 matmul (precision change, informational): step %.1f ms" % step_time(ldm, controllers))
from torch.profiler import profile, ProfilerActivity
torch.backends.cuda.matmul.allow_tf32 = False
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step_time(ldm, controllers, n=1)
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=60))
