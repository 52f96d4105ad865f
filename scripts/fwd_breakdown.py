"""Per-kernel device time of ONE skp_attn_map_fwd call at BASELINE shapes (run under
ncu --metrics gpu__time_duration.sum --clock-control none, or plain for a CUDA-event total).
python scripts/fwd_breakdown.py [B] [n_sel]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablekeypoints_b200 import ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
nsel = int(sys.argv[2]) if len(sys.argv) > 2 else 0
dev = "cuda:0"
spec = [(16, 1280, 20), (16, 1280, 20), (16, 1280, 20), (32, 640, 10)]
g = torch.Generator().manual_seed(0)
qs = [(torch.randn(B, s * s, C, generator=g) * 1.5).to(dev) for s, C, h in spec]
ks = [torch.randn(1, 500, C, generator=g).to(dev) for s, C, h in spec]
heads = [h for _, _, h in spec]
sc = [(C // h) ** -0.5 for _, C, h in spec]
idx = torch.randperm(500, generator=g)[:nsel].to(dev) if nsel else None
with torch.no_grad():
    for _ in range(3):
        m = ops.attention_maps(qs, ks, heads, sc, 128, token_idx=idx, precision="auto")
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(10):
        m = ops.attention_maps(qs, ks, heads, sc, 128, token_idx=idx, precision="auto")
    b.record()
    torch.cuda.synchronize()
print(f"attention_maps B={B} n_sel={nsel}: {a.elapsed_time(b) / 10:.3f} ms per call (incl. output allocation)")
