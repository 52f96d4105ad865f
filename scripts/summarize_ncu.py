"""Summarise ncu outputs into profiles/ (tracked).  Usage:
  python scripts/summarize_ncu.py launches gpurun_out/X_launches.csv > profiles/X_launches.txt
  python scripts/summarize_ncu.py full gpurun_out/X.ncu-rep > profiles/X_full.txt
"""
import collections
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tmem", "smsp__inst_executed.sum", "lts__t_sector_hit_rate.pct", "sm__cycles_active.avg"]


def own_kernel_names():
    """__global__ function names of stablekeypoints_b200/csrc/*.cu (ncu prints kernel names without the namespace)."""
    import glob, os, re
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "stablekeypoints_b200", "csrc")
    names = set()
    for f in glob.glob(os.path.join(root, "*.cu")):
        names.update(re.findall(r"(\w+_kernel)\s*\(", open(f).read()))
    return names


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = row["Kernel Name"].split("(")[0]
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(row["Metric Unit"], 1.0)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    ours = {k: a for k, a in agg.items() if any(n in k for n in own_kernel_names())}
    t_ours = sum(a[1] for a in ours.values())
    print(f"# {path}: per-kernel device time (ncu, cold-cache, serialised: compare SHARES)")
    print(f"# all launches: {sum(a[0] for a in agg.values())} kernels, {tot / 1e3:.1f} ms; this library's kernels: "
          f"{sum(a[0] for a in ours.values())} launches, {t_ours / 1e3:.1f} ms = {100 * t_ours / tot:.1f} % of the device time")
    print("# --- this library")
    for k, (n, t) in sorted(ours.items(), key=lambda x: -x[1][1]):
        print(f"{k[:72]:72s} n={n:5d} total={t / 1e3:10.3f} ms  avg={t / n:11.1f} us  share={t / tot * 100:5.1f}%")
    print("# --- everything else (stock PyTorch trunk), top 20")
    rest = [(k, a) for k, a in agg.items() if k not in ours]
    for k, (n, t) in sorted(rest, key=lambda x: -x[1][1])[:20]:
        print(f"{k[:72]:72s} n={n:5d} total={t / 1e3:10.3f} ms  avg={t / n:11.1f} us  share={t / tot * 100:5.1f}%")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print(f"# kernel: {r[hdr.index('Kernel Name')][:100]}")
        for i, n in enumerate(hdr):
            if any(n.startswith(k) for k in KEYS):
                print(f"{n} = {r[i]} {units[i]}")
        print()


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
