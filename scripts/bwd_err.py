"""Per-tensor error of the tcgen05 backward against the fp64 oracle autograd (max |err| / max |grad|).
SKP_BWD_LR=0/1 python scripts/bwd_err.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import keypoint_oracle as O
from stablekeypoints_b200 import ops

dev = "cuda:0"
for spec, T, R, B, shared in [([(4, 64, 4), (8, 32, 2)], 40, 16, 2, True), ([(8, 128, 2), (16, 64, 1)], 300, 32, 2, False),
                              ([(16, 128, 2), (32, 64, 1)], 500, 128, 1, True)]:
    g = torch.Generator().manual_seed(5)
    qs = [torch.randn(B, s * s, C, generator=g) for s, C, _ in spec]
    ks = [torch.randn(1 if shared else B, T, C, generator=g) for _, C, _ in spec]
    heads = [h for _, _, h in spec]
    sc = [(C // h) ** -0.5 for _, C, h in spec]
    g = torch.Generator().manual_seed(9)
    idx = torch.stack([torch.randperm(T, generator=g)[:6] for _ in range(B)])
    w = torch.randn(B, idx.shape[1], R, R, generator=g)
    q64 = [q.double().requires_grad_(True) for q in qs]
    k64 = [k.double().requires_grad_(True) for k in ks]
    loss = 0
    for b in range(B):
        m = O.attention_maps_from_qk([q[b] for q in q64], [k[0 if shared else b] for k in k64], heads, sc, R)
        loss = loss + (m[idx[b]] * w[b].double()).sum()
    grads = torch.autograd.grad(loss, q64 + k64)
    qc = [q.to(dev).requires_grad_(True) for q in qs]
    kc = [k.to(dev).requires_grad_(True) for k in ks]
    with torch.no_grad():
        full, stats = ops.attention_maps([q.detach() for q in qc], [k.detach() for k in kc], heads, sc, R, return_stats=True)
    sel = ops.selected_planes(full, idx.to(dev), qc, kc, heads, sc, R, row_stats=stats)
    (sel * w.to(dev)).sum().backward()
    names = [f"dq{l}" for l in range(len(spec))] + [f"dk{l}" for l in range(len(spec))]
    print(spec, T, R, B, shared, " ".join(f"{n}={float((t.grad.cpu().double() - wn).abs().max() / wn.abs().max()):.2e}"
                                          for n, t, wn in zip(names, qc + kc, grads)))
