"""The BENCHED problem pinned to the fp64 oracle (not to the repo's own CUDA-core kernel).

`bench.py`'s token-optimisation step calls `skp_attn_map_fwd` / `skp_attn_map_bwd` on 8 images (4 images + their warped
copies), SD-2.1-base capture set (3 x 16^2 with 20 heads + 1 x 32^2 with 10 heads, dim_head 64), T = 500, R = 128, one
shared k, per-image selected tokens; extraction calls the forward with 10 images and 10 selected tokens (mode C).  The
full maps of that problem are 8 x 500 x 128^2 values through 70 heads — hours on the CPU — but a pixel's maps depend on
no other pixel, so `oracle.attention_maps_at_pixels` evaluates >= 512 random (image, pixel) columns, all 500 tokens, in
fp64 in seconds (`ptp_utils.py:513-538` + `optimize.py:44-75` restated), and the gradient is linear in the upstream
gradient, so an upstream gradient that is zero outside those pixels makes fp64 autograd through the same restricted
oracle the exact reference for dq / dk of every layer.

Tolerances: forward rtol 1e-3 (north_star: fp32 maps within 1e-3 relative; bf16x3 products, fp32 accumulation);
row statistics: m + log2(l) == log2-sum-exp2 of the logits within 2e-4 absolute (log2 units);
backward: 1.5e-2 of each gradient's max (single-bf16 dS / W / k operands in the gradient products, fp32 accumulation:
the bar `tests/test_kernels_gpu.py` states for the upsample-after-logits formulation).
"""
import pytest
import torch

from oracle import keypoint_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SD21 = [(16, 1280, 20), (16, 1280, 20), (16, 1280, 20), (32, 640, 10)]
T, R = 500, 128


@pytest.fixture(scope="module")
def ops():
    from stablekeypoints_b200 import ops as _ops
    return _ops


def _problem(B, seed):
    g = torch.Generator().manual_seed(seed)
    qs = [torch.randn(B, s * s, c, generator=g) * 1.5 for s, c, _ in SD21]
    ks = [torch.randn(1, T, c, generator=g) for _, c, _ in SD21]
    heads = [h for _, _, h in SD21]
    scales = [(c // h) ** -0.5 for _, c, h in SD21]
    return g, qs, ks, heads, scales


def _pixels(g, B, per_image):
    """per_image random pixels for every image; always includes the corners and tile / row boundaries"""
    fixed = torch.tensor([0, R - 1, R, 127 * R, R * R - 1, 64 * R + 63, 64 * R + 64, 2 * R - 1])
    return [torch.cat([fixed, torch.randperm(R * R, generator=g)[: per_image - len(fixed)]]) for _ in range(B)]


def test_bench_forward_mode_a_and_row_stats_vs_fp64_oracle(ops):
    """Mode A (all 500 tokens, with row statistics) exactly as the token optimisation calls it: 8 images, 2-CTA cluster."""
    B = 8
    g, qs, ks, heads, sc = _problem(B, 101)
    maps, stats = ops.attention_maps([q.to(DEV) for q in qs], [k.to(DEV) for k in ks], heads, sc, R, precision="auto",
                                     return_stats=True)
    assert maps.shape == (B, T, R, R) and stats.shape == (B, sum(heads), R * R, 2)
    pix = _pixels(g, B, 72)                                   # 8 x 72 = 576 (image, pixel) columns
    maps = maps.reshape(B, T, R * R).cpu()
    stats = stats.cpu()
    for b in range(B):
        want, lse2 = O.attention_maps_at_pixels([q[b] for q in qs], [k[0] for k in ks], heads, sc, R, pix[b], return_lse2=True)
        got = maps[b][:, pix[b]].double()
        assert torch.allclose(got, want, rtol=1e-3, atol=1e-9), f"image {b}: max rel {((got - want).abs() / want).max():.2e}"
        st = stats[b][:, pix[b]].double()                     # (heads, n, 2) = (max, sum exp2) in log2 units
        got_lse = st[..., 0] + torch.log2(st[..., 1])
        assert (got_lse - lse2).abs().max().item() < 2e-4


def test_bench_forward_mode_c_vs_fp64_oracle(ops):
    """Mode C (10 selected tokens) exactly as extraction calls it: 10 augmented images per launch, ids shared."""
    B = 10
    g, qs, ks, heads, sc = _problem(B, 102)
    idx = torch.arange(10) * 50
    idx[3], idx[7] = 499, 255                                 # last token (second CTA's ragged end) and the half boundary
    maps = ops.attention_maps([q.to(DEV) for q in qs], [k.to(DEV) for k in ks], heads, sc, R, token_idx=idx.to(DEV),
                              precision="auto")
    assert maps.shape == (B, 10, R, R)
    pix = _pixels(g, B, 56)                                   # 560 columns
    maps = maps.reshape(B, 10, R * R).cpu()
    for b in range(B):
        want = O.attention_maps_at_pixels([q[b] for q in qs], [k[0] for k in ks], heads, sc, R, pix[b])[idx]
        got = maps[b][:, pix[b]].double()
        assert torch.allclose(got, want, rtol=1e-3, atol=1e-9), f"image {b}: max rel {((got - want).abs() / want).max():.2e}"


def test_bench_backward_vs_fp64_oracle_autograd(ops):
    """The tcgen05 backward of the benched call (8 images, per-image token ids, forward statistics handed over) against
    fp64 autograd through the oracle restricted to the pixels that carry upstream gradient: dq of every layer (16^2 ones
    on the upsample-after-logits path, the 32^2 one on its own path) and dk (summed over the 8 images: shared k)."""
    B, K = 8, 10
    g, qs, ks, heads, sc = _problem(B, 103)
    idx = torch.stack([torch.randperm(T, generator=g)[:K] for _ in range(B)])
    pix = _pixels(g, B, 64)                                   # 512 (image, pixel) columns carry gradient
    w = torch.zeros(B, K, R * R)
    for b in range(B):
        w[b][:, pix[b]] = torch.randn(K, len(pix[b]), generator=g)
    # fp64 reference
    q64 = [q.double().requires_grad_(True) for q in qs]
    k64 = [k.double().requires_grad_(True) for k in ks]
    loss = 0.0
    for b in range(B):
        m = O.attention_maps_at_pixels([q[b] for q in q64], [k[0] for k in k64], heads, sc, R, pix[b])      # (T, n)
        loss = loss + (m[idx[b]] * w[b][:, pix[b]].double()).sum()
    want = torch.autograd.grad(loss, q64 + k64)
    # CUDA path, as optimize.token_opt_losses drives it
    qc = [q.to(DEV).requires_grad_(True) for q in qs]
    kc = [k.to(DEV).requires_grad_(True) for k in ks]
    with torch.no_grad():
        full, stats, fwd_ws = ops.attention_maps([q.detach() for q in qc], [k.detach() for k in kc], heads, sc, R,
                                                 return_stats=True, return_workspace=True)
    # (i) as the optimisation drives it: the backward reuses the operands the forward staged in its workspace
    # (skp_attn_map_bwd_reuse); (ii) without the hand-over (skp_attn_map_bwd semantics): bit-identical gradients
    grads = []
    for ws in (fwd_ws, None):
        for t in qc + kc:
            t.grad = None
        sel = ops.selected_planes(full, idx.to(DEV), qc, kc, heads, sc, R, row_stats=stats, fwd_workspace=ws)
        (sel * w.reshape(B, K, R, R).to(DEV)).sum().backward()
        grads.append([t.grad.clone() for t in qc + kc])
    names = [f"dq[{i}]" for i in range(4)] + [f"dk[{i}]" for i in range(4)]
    for name, got, ref in zip(names, grads[0], want):
        scale = float(ref.abs().max())
        err = (got.cpu().double() - ref).abs().max().item()
        print(f"  {name}: max error {err / scale:.2e} of the gradient's max")      # pytest -s: the measured precision of this build
        assert torch.isfinite(got).all() and err <= 1.5e-2 * scale, f"{name}: {err / scale:.2e} of max"
    for name, a, b_ in zip(names, grads[0], grads[1]):
        # same operands, same kernels: only the order of the fp32 reductions into dz / dq may differ
        assert (a - b_).abs().max().item() <= 1e-5 * float(b_.abs().max()), name
