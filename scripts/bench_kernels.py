"""Kernel-only timing of the attention-map forward / backward C-ABI calls at BASELINE shapes
(SD-2.1-base capture set, T=500, R=128, B images).  python scripts/bench_kernels.py [B] [precision]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablekeypoints_b200 import ops, _lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
prec = sys.argv[2] if len(sys.argv) > 2 else "auto"
dev = "cuda:0"
SPECS = {"sd21": [(16, 1280, 20), (16, 1280, 20), (16, 1280, 20), (32, 640, 10)],          # BASELINE capture set
         "sd15": [(16, 1280, 8), (16, 1280, 8), (16, 1280, 8), (32, 640, 8)],              # what the reference loads: d = 160 / 80
         "sdxl": [(32, 1280, 20)] * 4,                                                     # configs[4]
         "cub9": [(16, 1280, 20)] * 3 + [(32, 640, 10)] * 3 + [(64, 320, 5)] * 3}          # configs[3]
spec = SPECS[os.environ.get("BK_MODEL", "sd21")]
g = torch.Generator().manual_seed(0)
qs = [(torch.randn(B, s * s, C, generator=g) * 1.5).to(dev).requires_grad_(True) for s, C, h in spec]
T = int(os.environ.get("BK_T", "500"))
ks = [torch.randn(1, T, C, generator=g).to(dev).requires_grad_(True) for s, C, h in spec]
heads = [h for _, _, h in spec]
sc = [(C // h) ** -0.5 for _, C, h in spec]
idx = torch.stack([torch.randperm(T, generator=g)[:10] for _ in range(B)]).to(dev)
w = torch.randn(B, 10, 128, 128, generator=g).to(dev)


def run():
    with torch.no_grad():
        full, stats, ws = ops.attention_maps([q.detach() for q in qs], [k.detach() for k in ks], heads, sc, 128, precision=prec,
                                             return_stats=True, return_workspace=True)
    sel = ops.selected_planes(full, idx, qs, ks, heads, sc, 128, precision=prec, row_stats=stats,
                              fwd_workspace=None if os.environ.get("BK_NO_REUSE") else ws)
    (sel * w).sum().backward()


if os.environ.get("BK_SELECT"):        # mode C only (extraction): forward restricted to 10 token ids shared by all images
    ids = idx[0].contiguous()

    def run():
        with torch.no_grad():
            ops.attention_maps([q.detach() for q in qs], [k.detach() for k in ks], heads, sc, 128, token_idx=ids, precision=prec)

for _ in range(3):
    run()
_lib.timing_start()
for _ in range(10):
    run()
t = _lib.timing_summary()
flops = 2 * 128 * 128 * T * sum(C for _, C, _ in spec) * B
for k, v in t.items():
    ms = v["ms"] / v["calls"]
    mult = 1 if "fwd" in k else 2
    print(f"{k}: {ms:.3f} ms/call  ({ms / B * 1e3:.0f} us/image)  {mult * flops / ms / 1e9:.1f} TFLOP/s algorithmic")
