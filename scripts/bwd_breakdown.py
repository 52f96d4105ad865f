"""Per-kernel device time of ONE skp_attn_map_bwd call at BASELINE shapes (run under
ncu --metrics gpu__time_duration.sum --clock-control none).  python scripts/bwd_breakdown.py [B]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablekeypoints_b200 import ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dev = "cuda:0"
spec = [(16, 1280, 20), (16, 1280, 20), (16, 1280, 20), (32, 640, 10)]
g = torch.Generator().manual_seed(0)
qs = [(torch.randn(B, s * s, C, generator=g) * 1.5).to(dev).requires_grad_(True) for s, C, h in spec]
ks = [torch.randn(1, 500, C, generator=g).to(dev).requires_grad_(True) for s, C, h in spec]
heads = [h for _, _, h in spec]
sc = [(C // h) ** -0.5 for _, C, h in spec]
idx = torch.stack([torch.randperm(500, generator=g)[:10] for _ in range(B)]).to(dev)
w = torch.randn(B, 10, 128, 128, generator=g).to(dev)
with torch.no_grad():
    full, stats = ops.attention_maps([q.detach() for q in qs], [k.detach() for k in ks], heads, sc, 128, return_stats=True)
torch.cuda.synchronize()
sel = ops.selected_planes(full, idx, qs, ks, heads, sc, 128, row_stats=stats)
(sel * w).sum().backward()
torch.cuda.synchronize()
