"""GPU box: how much of a step is device time?  Sum of CUDA kernel durations (kineto) vs the wall time of the same steps, for
the extraction pass and the optimisation step of bench.py's default workloads.  If the device is busy > 95 % of the wall
time, CUDA-graph capture of the forward cannot move the step by more than the remainder."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

args = bench.parse()
args.steps, args.warmup = 2, 2
for name in ("extract", "optimize"):
    cfg = bench.workload_cfg(name, args)
    env = bench.setup_pipeline(args, cfg, 0)
    pool = bench.make_pool(cfg, 0).to(env[0])
    def run():
        if name == "extract":
            bench.run_extract(args, cfg, *env, pool, torch.arange(10) * 50, lambda s: None, False)
        else:
            bench.run_optimize(args, cfg, *env, bench.ImagePool(pool), cfg["batch_size"], cfg["images_per_pass"], False, lambda s: None)
    run()
    torch.cuda.synchronize()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA, torch.profiler.ProfilerActivity.CPU]) as prof:
        t0 = time.perf_counter()
        run()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    dev_ms = sum(e.device_time for e in ev) / 1e3 if ev else float("nan")
    n = args.steps + args.warmup
    print(f"{name}: wall {wall * 1e3 / n:.1f} ms/step, device kernels {dev_ms / n:.1f} ms/step ({100 * dev_ms / (wall * 1e3):.1f} % busy), "
          f"{len(ev) / n:.0f} device activities per step", flush=True)
    del env, pool
    torch.cuda.empty_cache()
