"""GPU probe for the tcgen05 attention-map kernel: every template variant against the SIMT fp32
kernel, smallest shapes first, printing max errors (diagnostics for blind bring-up)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablekeypoints_b200 import ops

dev = "cuda:0"

def layers(spec, T, B, seed, shared=True, qscale=1.5):
    g = torch.Generator().manual_seed(seed)
    qs = [(torch.randn(B, s * s, C, generator=g) * qscale).to(dev) for s, C, h in spec]
    ks = [torch.randn(1 if shared else B, T, C, generator=g).to(dev) for s, C, h in spec]
    return qs, ks, [h for _, _, h in spec], [(C // h) ** -0.5 for _, C, h in spec]

cases = [
    ("1 layer 1 head d=64 T=100 s=R",   [(16, 64, 1)], 100, 16, 1),
    ("1 layer 1 head d=64 T=256",       [(4, 64, 1)], 256, 16, 1),
    ("1 layer 2 heads d=32 T=37",       [(4, 64, 2)], 37, 16, 1),
    ("3 heads d=16 (ring wrap)",        [(4, 48, 3)], 64, 16, 1),
    ("d=160 (3 k-blocks) 5 heads",      [(4, 800, 5)], 200, 16, 2),
    ("split T=300 1 head",              [(4, 64, 1)], 300, 16, 1),
    ("split T=500 4 layers",            [(4, 128, 2), (4, 128, 2), (8, 64, 4), (16, 32, 2)], 500, 16, 2),
    ("R=32 multi tile, T=500",          [(8, 128, 2), (16, 64, 1)], 500, 32, 2),
]
ok = True
for name, spec, T, R, B in cases:
    qs, ks, heads, sc = layers(spec, T, B, 1)
    ref = ops.attention_maps(qs, ks, heads, sc, R, precision="simt_f32")
    torch.cuda.synchronize()
    idx = torch.tensor([T - 1, 0, T // 2, 3 % T, 7 % T], device=dev)
    for prec, tol in (("tc_bf16x3", 1e-3), ("tc_bf16", 6e-2)):
        t0 = time.time()
        got = ops.attention_maps(qs, ks, heads, sc, R, precision=prec)
        torch.cuda.synchronize()
        rel = ((got - ref).abs() / (ref.abs() + 1e-7)).max().item()
        gsel = ops.attention_maps(qs, ks, heads, sc, R, token_idx=idx, precision=prec)
        torch.cuda.synchronize()
        rel2 = ((gsel - ref[:, idx]).abs() / (ref[:, idx].abs() + 1e-7)).max().item()
        good = rel < tol and rel2 < tol and torch.isfinite(got).all().item()
        ok &= good
        print(f"{'OK ' if good else 'BAD'} {name:32s} {prec:10s} all-T rel {rel:.2e}  selected rel {rel2:.2e}  sum-1 {abs(got.sum(1)-1).max().item():.1e}  ({time.time()-t0:.2f}s)", flush=True)
print("PROBE", "PASS" if ok else "FAIL")
