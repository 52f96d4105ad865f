"""Parity of the CUDA kernels (through the C ABI) against the CPU oracle and the committed
golden vectors produced by the live reference.  Run on the B200 box: pytest -m gpu.

Tolerances (stated per north_star): integer outputs (argmax pixels, ranked / FPS token ids)
bit-exact; attention maps rtol 1e-3 in the fp32-grade modes (simt_f32, tc_bf16x3), bf16 mode
rtol 3e-2; soft-argmax keypoints 0.5 px.
"""
import os

import numpy as np
import pytest
import torch

from oracle import keypoint_oracle as O

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")
DEV = "cuda:0"


def _load(name):
    return {k: v for k, v in np.load(os.path.join(G, name)).items()}


def _cu(x):
    return torch.as_tensor(x).to(DEV)


@pytest.fixture(scope="module")
def ops():
    from stablekeypoints_b200 import ops as _ops
    return _ops


def _assert_rank_equal_up_to_ties(got, want, score64, rel_tol=3e-6):
    """Ranked ids must be identical except where the fp64 scores of the two candidates for a
    rank differ by less than rel_tol (the reference's own fp32 sums are not reproducible below
    that: SURVEY.md 7.3 'bit-exact integer outputs are only well-defined away from ties')."""
    got, want = np.asarray(got), np.asarray(want)
    assert got.shape == want.shape
    for g_, w_ in zip(got, want):
        if g_ != w_:
            a, b = float(score64[g_]), float(score64[w_])
            assert abs(a - b) <= rel_tol * max(abs(a), abs(b)), f"rank mismatch {g_} vs {w_}: scores {a} {b}"


def _golden_layers():
    d = _load("attention_maps.npz")
    qs = [torch.from_numpy(d[f"q_low{i}"])[None] for i in range(4)]
    ks = [torch.from_numpy(d[f"k{i}"])[None] for i in range(4)]
    return d, qs, ks, [int(h) for h in d["heads"]], [float(s) for s in d["scale"]], int(d["R"])


@pytest.mark.parametrize("precision,rtol", [("simt_f32", 1e-3), ("auto", 1e-3)])
def test_attention_maps_golden(ops, precision, rtol):
    d, qs, ks, heads, sc, R = _golden_layers()
    A = ops.attention_maps([_cu(q) for q in qs], [_cu(k) for k in ks], heads, sc, R, precision=precision)[0]
    ref = torch.from_numpy(d["ref_maps_A"])
    assert torch.allclose(A.cpu(), ref, rtol=rtol, atol=1e-7)
    assert torch.allclose(A.sum(0).cpu(), torch.ones(R, R), atol=1e-5)
    Bm = ops.bilinear_upsample(A[:8].contiguous(), 128)
    assert torch.allclose(Bm.cpu(), torch.from_numpy(d["ref_maps_B"]), rtol=rtol, atol=1e-7)
    idx = torch.from_numpy(d["indices"])
    Cm = ops.attention_maps([_cu(q) for q in qs], [_cu(k) for k in ks], heads, sc, R, token_idx=_cu(idx), precision=precision)[0]
    assert torch.allclose(Cm.cpu(), ref[idx], rtol=rtol, atol=1e-7)
    assert torch.allclose(ops.bilinear_upsample(Cm, 128).cpu(), torch.from_numpy(d["ref_maps_C"]), rtol=rtol, atol=1e-7)


def _random_layers(spec, T, B, seed, shared_k=True):
    g = torch.Generator().manual_seed(seed)
    qs, ks, heads, scales = [], [], [], []
    for (s, C, h) in spec:
        qs.append(torch.randn(B, s * s, C, generator=g) * 1.5)
        ks.append(torch.randn(1 if shared_k else B, T, C, generator=g))
        heads.append(h)
        scales.append((C // h) ** -0.5)
    return qs, ks, heads, scales


@pytest.mark.parametrize("precision", ["simt_f32", "auto", "auto-staged"])
@pytest.mark.parametrize("spec,T,R,B,shared", [
    ([(4, 64, 4), (4, 64, 4), (8, 32, 2)], 37, 16, 2, True),          # unequal heads, ragged T
    ([(8, 128, 2), (16, 64, 4), (32, 32, 2)], 130, 32, 1, False),     # three scales incl. s == R
    ([(4, 96, 2)], 500, 16, 3, True),                                 # dim_head 48, full token count
    ([(16, 128, 2), (32, 64, 1)], 70, 128, 1, True),                  # BASELINE geometry: 64- and 128-tap windows
])
def test_attention_maps_vs_oracle_unequal_heads(ops, monkeypatch, precision, spec, T, R, B, shared):
    """"auto" = tcgen05 bf16x3, operands from the native-resolution logits where the windows allow it (attn_tc_lr.cu);
    "auto-staged" = the same kernel on the staged upsampled q' (SKP_FWD_LR=0).  Same 1e-3 bar for all."""
    monkeypatch.setenv("SKP_FWD_LR", "0" if precision == "auto-staged" else "1")
    precision = "auto" if precision == "auto-staged" else precision
    qs, ks, heads, sc = _random_layers(spec, T, B, 3, shared)
    got = ops.attention_maps([_cu(q) for q in qs], [_cu(k) for k in ks], heads, sc, R, precision=precision).cpu()
    for b in range(B):
        want = O.attention_maps_from_qk([q[b] for q in qs], [k[0 if shared else b] for k in ks], heads, sc, R)
        assert torch.allclose(got[b], want, rtol=1e-3, atol=1e-7)


@pytest.mark.parametrize("fast", ["1", "0"])
@pytest.mark.parametrize("qscale", [1.0, 8.0])
def test_attention_maps_bounded_and_unbounded_logit_paths(ops, monkeypatch, fast, qscale):
    """The forward skips the row-max pass when the range of the native-resolution logits of an (image, head) bounds
    exp2(S - centre) away from over/underflow (attn_tc.cu FW_FAST_HALFWIDTH), else subtracts the row max: both against the
    fp64 oracle, with small logits (bounded path unless SKP_FWD_FAST=0) and a mix of both in one call: the second layer's q
    is scaled so that its logits span more than the bound allows (range ~ 170 log2 units; much larger logits leave the 1e-3
    bar of ANY bf16x3 product: the split drops the lo x lo term, 2^-17 relative to |S|).  On the bounded path the shift
    is ONE constant per (image, head); on the other it is the row's own maximum."""
    monkeypatch.setenv("SKP_FWD_FAST", fast)
    spec, T, R, B = [(16, 128, 2), (32, 64, 1)], 300, 128, 2
    qs, ks, heads, sc = _random_layers(spec, T, B, 17, True)
    qs[1] = qs[1] * qscale
    qc, kc = [_cu(q) for q in qs], [_cu(k) for k in ks]
    got, stats = ops.attention_maps(qc, kc, heads, sc, R, precision="auto", return_stats=True)
    g = torch.Generator().manual_seed(3)
    pix = torch.randperm(R * R, generator=g)[:96]
    for b in range(B):
        want, lse2 = O.attention_maps_at_pixels([q[b] for q in qs], [k[0] for k in ks], heads, sc, R, pix, return_lse2=True)
        have = got[b].reshape(T, -1)[:, pix].cpu().double()
        assert torch.allclose(have, want, rtol=1e-3, atol=1e-9), f"max rel {((have - want).abs() / (want + 1e-9)).max():.2e}"
        st = stats[b][:, pix].cpu().double()
        assert (st[..., 0] + torch.log2(st[..., 1]) - lse2).abs().max().item() < 1e-3
        m_l0, m_l1 = stats[b][:heads[0], pix, 0], stats[b][heads[0]:, pix, 0]          # shifts of layer 0 / layer 1 rows
        if fast == "1":
            assert float((m_l0 - m_l0[:, :1]).abs().max()) == 0.0                       # one constant per (image, head)
        if qscale > 1.0 or fast == "0":
            assert float((m_l1 - m_l1[:, :1]).abs().max()) > 0.0                        # per-row maxima
    assert torch.allclose(got.sum(1), torch.ones(B, R, R, device=DEV), atol=2e-5)


def test_attention_maps_backward_matches_oracle_autograd(ops):
    spec, T, R, B = [(4, 64, 4), (8, 32, 2)], 40, 16, 2
    qs, ks, heads, sc = _random_layers(spec, T, B, 5, True)
    idx = torch.tensor([7, 0, 33])
    g = torch.Generator().manual_seed(9)
    w = torch.randn(B, len(idx), R, R, generator=g)
    # oracle in fp64 on CPU
    q64 = [q.double().requires_grad_(True) for q in qs]
    k64 = [k.double().requires_grad_(True) for k in ks]
    loss = 0
    for b in range(B):
        m = O.attention_maps_from_qk([q[b] for q in q64], [k[0] for k in k64], heads, sc, R)
        loss = loss + (m[idx] * w[b].double()).sum()
    grads = torch.autograd.grad(loss, q64 + k64)
    qc = [_cu(q).requires_grad_(True) for q in qs]
    kc = [_cu(k).requires_grad_(True) for k in ks]
    m = ops.attention_maps(qc, kc, heads, sc, R, precision="simt_f32")
    (m[:, _cu(idx)] * _cu(w)).sum().backward()
    for got, want in zip([t.grad for t in qc + kc], grads):
        scale = float(want.abs().max())
        assert torch.allclose(got.cpu().double(), want, rtol=2e-3, atol=2e-5 * scale)


def test_full_size_properties_sd21_shape(ops):
    """BASELINE.json config sizes (T=500, R=128, SD-2.1 heads): size-independent properties."""
    spec = [(16, 1280, 20), (16, 1280, 20), (16, 1280, 20), (32, 640, 10)]
    qs, ks, heads, sc = _random_layers(spec, 500, 1, 11, True)
    qc, kc = [_cu(q) for q in qs], [_cu(k) for k in ks]
    m = ops.attention_maps(qc, kc, heads, sc, 128, precision="auto")
    assert m.shape == (1, 500, 128, 128)
    assert torch.allclose(m.sum(1), torch.ones(1, 128, 128, device=DEV), atol=2e-5)   # softmax over tokens then means
    assert float(m.min()) >= 0.0
    idx = torch.tensor([499, 0, 250, 3, 77, 12, 498, 100, 9, 301])
    c = ops.attention_maps(qc, kc, heads, sc, 128, token_idx=_cu(idx), precision="auto")
    assert torch.allclose(c, m[:, _cu(idx)], rtol=1e-4, atol=1e-8)                  # gather commutes
    # a few pixels against the fp64 oracle restricted to those pixels is covered by the SIMT cross-check:
    m2 = ops.attention_maps(qc, kc, heads, sc, 128, precision="simt_f32")
    assert torch.allclose(m, m2, rtol=1e-3, atol=1e-7)
    am = ops.plane_argmax(m[0])
    assert torch.equal(am.cpu(), torch.argmax(m[0].reshape(500, -1), dim=-1).cpu())


def test_readout_and_selection_golden(ops):
    d = _load("selection_readout.npz")
    m = _cu(d["maps"])
    n, h, w = m.shape
    am = ops.plane_argmax(m)
    rc = torch.stack([am // w, am % w], -1).float() + 0.5
    assert np.array_equal(rc.cpu().numpy(), d["ref_max_pixel"])
    assert np.array_equal(ops.find_k_max_pixels(m, 3).cpu().numpy(), d["ref_k_max_pixels3"])
    wa = ops.weighted_avg_readout(m.clone()).cpu().numpy()
    assert np.abs(wa - d["ref_weighted_avg"]).max() < 0.5      # north_star: soft-argmax within 0.5 px
    assert np.allclose(wa, d["ref_weighted_avg"], atol=1e-3)
    for ns in (1, 2):
        kl, _, _ = ops.plane_stats(m, sigma=2.0, epsilon=1e-5, num_subjects=ns)
        kl64 = O.gaussian_kl(torch.from_numpy(d["maps"]), 2.0, 1e-5, ns, dtype=torch.float64)
        assert torch.allclose(kl.cpu().double(), kl64, rtol=1e-4)
        _assert_rank_equal_up_to_ties(ops.rank_ascending(kl, 25).cpu().numpy(), d[f"ref_topk_gaussian_ns{ns}"], kl64)
    _, ent, _ = ops.plane_stats(m, want_kl=False, want_entropy=True)
    ent64 = O.pixel_entropy(torch.from_numpy(d["maps"]), dtype=torch.float64)
    assert torch.allclose(ent.cpu().double(), ent64, rtol=1e-6)
    _assert_rank_equal_up_to_ties(ops.rank_ascending(ent, 25).cpu().numpy(), d["ref_entropy_sort"], ent64)
    # on identical float scores the ranking kernel itself is exact
    assert torch.equal(ops.rank_ascending(ent, 25).cpu(), torch.argsort(ent.cpu(), stable=True)[:25])
    cand = _cu(d["ref_topk_gaussian_ns1"])
    assert np.array_equal(ops.furthest_point_sampling(am, h, w, cand, 10).cpu().numpy(), d["ref_fps"])
    pos = ops.find_k_max_pixels(m[:6].contiguous(), 2) / 64
    assert np.allclose(ops.gaussian_circles(pos, 64, 2.0).cpu().numpy(), d["ref_gaussian_circles"], atol=2e-7)
    m48 = _cu(d["maps48"])
    am48 = ops.plane_argmax(m48)
    got = ops.furthest_point_sampling(am48, 48, 48, torch.arange(25, device=DEV), 10)
    assert np.array_equal(got.cpu().numpy(), d["ref_fps48"])


def test_plane_stats_register_path_matches_oracle_and_two_pass_kernel(ops):
    """Planes of 512*4*VPT pixels with one subject take the register-resident single-pass kernel; the same data at a
    16-byte-misaligned address takes the two-pass kernel.  Both against the fp64 oracle (ptp_utils.py:86-112, 165-187)."""
    g = torch.Generator().manual_seed(5)
    for hw in (32, 64, 128):        # 64x32 (VPT 1), 64^2 (VPT 2), 128^2 (VPT 8)
        h, w = (64, 32) if hw == 32 else (hw, hw)
        m = torch.rand(12, h, w, generator=g) * 4.0
        m[3, h - 1, w - 1] = 9.0      # peak in a corner: Gaussian window clipped by the border
        m[4, 0, 0] = 9.0
        kl64 = O.gaussian_kl(m, 2.0, 1e-5, 1, dtype=torch.float64) if h == w else None
        ent64 = O.pixel_entropy(m, dtype=torch.float64)
        flat = torch.empty(m.numel() + 1, device=DEV)
        flat[1:] = _cu(m).reshape(-1)
        for name, mm in (("register", _cu(m)), ("two-pass", flat[1:].view(12, h, w))):
            kl, ent, am = ops.plane_stats(mm, sigma=2.0, epsilon=1e-5, want_kl=h == w, want_entropy=True, want_argmax=True)
            assert torch.equal(am.cpu(), torch.argmax(m.reshape(12, -1), -1)), name
            assert torch.allclose(ent.cpu().double(), ent64, rtol=1e-6), name
            if kl64 is not None:
                assert torch.allclose(kl.cpu().double(), kl64, rtol=1e-4), name


def test_readout_edge_cases(ops):
    t = torch.zeros(3, 16, 16)
    t[0, 3, 4] = 1; t[0, 5, 1] = 1            # tie: first flat index wins
    t[1] = -1.0; t[1, 15, 15] = -0.5           # all negative
    t[2, 0, 0] = float("nan")                  # NaN counts as the maximum (torch.argmax)
    am = ops.plane_argmax(_cu(t)).cpu()
    assert am.tolist() == [3 * 16 + 4, 255, 0]
    big = torch.rand(10, 512, 512)
    assert torch.equal(ops.plane_argmax(_cu(big)).cpu(), torch.argmax(big.reshape(10, -1), -1))
    odd = torch.rand(7, 33, 35)
    assert torch.equal(ops.plane_argmax(_cu(odd)).cpu(), torch.argmax(odd.reshape(7, -1), -1))
    s = torch.tensor([0.5, -1.0, 0.5, float("nan"), -0.0, 0.0, 3.0])
    assert ops.rank_ascending(_cu(s), 7).cpu().tolist() == [1, 4, 5, 0, 2, 6, 3]
    with pytest.raises(NameError):
        ops.furthest_point_sampling(_cu(torch.zeros(4, dtype=torch.int64)), 8, 8, _cu(torch.tensor([1])), 2)


def test_losses_warps_tta_golden(ops):
    d = _load("losses_tta.npz")
    img, theta = _cu(d["img"]), _cu(d["theta"])
    wimg = ops.affine_warp(img, theta)
    assert np.allclose(wimg.cpu().numpy(), d["ref_warped"], atol=2e-5)
    assert np.allclose(ops.affine_warp(wimg, ops.invert_theta(theta)).cpu().numpy(), d["ref_unwarped"], atol=2e-5)
    a = _cu(d["a"]).requires_grad_(True)
    b = _cu(d["b"]).requires_grad_(True)
    idx = torch.arange(10, device=DEV)
    ls, le = ops.keypoint_losses(a, b, idx, ops.invert_theta(theta[0]), 2.0)
    assert abs(ls.item() - float(d["ref_sharpening"])) < 1e-5 * abs(float(d["ref_sharpening"]))
    assert abs(le.item() - float(d["ref_equivariance"])) < 1e-4 * abs(float(d["ref_equivariance"]))
    (ls * 100 + le * 1000).backward()
    for got, want in ((a.grad, d["ref_grad_a"]), (b.grad, d["ref_grad_b"])):
        assert np.allclose(got.cpu().numpy(), want, rtol=1e-3, atol=2e-5 * np.abs(want).max())
    # sharpening alone, on a subset of planes of a larger stack
    a2 = _cu(d["a"]).requires_grad_(True)
    sel = torch.tensor([4, 1, 7])
    ls2, _ = ops.keypoint_losses(a2, None, _cu(sel), None, 2.0)
    want = O.sharpening_loss(torch.from_numpy(d["a"])[sel], 2.0)
    assert abs(ls2.item() - want.item()) < 1e-5 * abs(want.item())
    # TTA ensemble
    maps, thetas = _cu(d["tta_maps"]), _cu(d["tta_thetas"])
    A, K, H, W = maps.shape
    s_ = torch.zeros(K, H, W, device=DEV); n_ = torch.zeros(K, H, W, device=DEV)
    ops.tta_accumulate(maps[:2].contiguous(), ops.invert_theta(thetas[:2]), s_, n_)
    ops.tta_accumulate(maps[2:].contiguous(), ops.invert_theta(thetas[2:]), s_, n_)
    ens = ops.tta_finalize(s_, n_)
    assert np.allclose(ens.cpu().numpy(), d["ref_tta"], rtol=1e-4, atol=2e-5)
    am = ops.plane_argmax(ens)
    rc = torch.stack([am // W, am % W], -1).float() + 0.5
    assert np.array_equal(rc.cpu().numpy(), d["ref_tta_keypoints"])


def test_tta_with_fused_upsample_matches_oracle(ops):
    g = torch.Generator().manual_seed(2)
    maps = torch.rand(3, 4, 32, 32, generator=g)
    thetas = torch.stack([O.affine_theta(10.0, 0.9, 0.1, -0.05), O.affine_theta(-7.0, 0.85, -0.2, 0.15),
                          O.affine_theta(0.0, 1.0, 0.0, 0.0)])
    up = [O.bilinear_upsample(m, 128) for m in maps]
    want = O.tta_ensemble(up, list(thetas))
    s_ = torch.zeros(4, 128, 128, device=DEV); n_ = torch.zeros(4, 128, 128, device=DEV)
    ops.tta_accumulate(_cu(maps), ops.invert_theta(_cu(thetas)), s_, n_)
    got = ops.tta_finalize(s_, n_).cpu()
    assert torch.allclose(got, want, rtol=1e-4, atol=2e-5)


def test_resampling_matches_oracle(ops):
    g = torch.Generator().manual_seed(4)
    x = torch.randn(2, 8, 8, 24, generator=g)
    want = O.bicubic_upsample(x.permute(0, 3, 1, 2).double(), 32).permute(0, 2, 3, 1)
    assert torch.allclose(ops.bicubic_upsample_nhwc(_cu(x), 32).cpu().double(), want, atol=1e-5)
    y = torch.randn(5, 16, 16, generator=g)
    assert torch.allclose(ops.bilinear_upsample(_cu(y), 64).cpu().double(), O.bilinear_upsample(y.double(), 64), atol=1e-6)
    assert torch.equal(ops.bilinear_upsample(_cu(y), 16).cpu(), y)


@pytest.mark.parametrize("lr", ["0", "1", "2"])
@pytest.mark.parametrize("spec,T,R,B,shared", [
    ([(4, 64, 4), (8, 32, 2)], 40, 16, 2, True),             # dim_head 16, several pixel rows per tile
    ([(8, 128, 2), (16, 64, 1)], 300, 32, 2, False),         # dim_head 64, two token halves, per-image k
    ([(16, 128, 2), (32, 64, 1)], 260, 128, 1, True),        # BASELINE geometry: one image row per tile, 8 tiles per window
])
def test_attention_maps_backward_tcgen05_matches_oracle_autograd(ops, monkeypatch, spec, T, R, B, shared, lr):
    """tcgen05 backward against fp64 autograd through the oracle.  Both gradient formulations: lr=0 staged q' with
    dQ'/dK products and the bicubic adjoint (bf16 dS, k, q' operands): 1e-2 of the gradient's max; lr=1 "upsample after
    the logits" on the layers with 64-tap windows (the interpolation table is a bf16 operand too): 1.5e-2 of the max
    (measured 1.04e-2 worst case on these problems, 0.2-0.4e-2 on unit-variance inputs, scripts/bwd_err.py)."""
    monkeypatch.setenv("SKP_BWD_LR", lr)      # 2: 128-tap windows (the 32^2 layer at R = 128) through the 128-token wide-window kernel too
    tol = 1e-2 if lr == "0" else 1.5e-2
    qs, ks, heads, sc = _random_layers(spec, T, B, 5, shared)
    g = torch.Generator().manual_seed(9)
    idx = torch.stack([torch.randperm(T, generator=g)[:6] for _ in range(B)])       # per-image token ids
    w = torch.randn(B, idx.shape[1], R, R, generator=g)
    q64 = [q.double().requires_grad_(True) for q in qs]
    k64 = [k.double().requires_grad_(True) for k in ks]
    loss = 0
    for b in range(B):
        m = O.attention_maps_from_qk([q[b] for q in q64], [k[0 if shared else b] for k in k64], heads, sc, R)
        loss = loss + (m[idx[b]] * w[b].double()).sum()
    grads = torch.autograd.grad(loss, q64 + k64)
    qc = [_cu(q).requires_grad_(True) for q in qs]
    kc = [_cu(k).requires_grad_(True) for k in ks]
    with torch.no_grad():
        full, stats = ops.attention_maps([q.detach() for q in qc], [k.detach() for k in kc], heads, sc, R, return_stats=True)
    sel = ops.selected_planes(full, _cu(idx), qc, kc, heads, sc, R, row_stats=stats)
    (sel * _cu(w)).sum().backward()
    for got, want in zip([t.grad for t in qc + kc], grads):
        scale = float(want.abs().max())
        assert torch.isfinite(got).all()
        assert (got.cpu().double() - want).abs().max().item() <= tol * scale
    # and the CUDA-core backward (no statistics passed) agrees on the same problem
    qd = [_cu(q).requires_grad_(True) for q in qs]
    kd = [_cu(k).requires_grad_(True) for k in ks]
    sel2 = ops.selected_planes(full, _cu(idx), qd, kd, heads, sc, R, precision="simt_f32")
    (sel2 * _cu(w)).sum().backward()
    for a, b_ in zip([t.grad for t in qc + kc], [t.grad for t in qd + kd]):
        assert (a - b_).abs().max().item() <= tol * float(b_.abs().max())


def test_backward_sd15_head_shapes_on_tensor_cores(ops):
    """SD-1.5 — the model the reference actually loads (`main.py:29`): 8 heads, dim_head 160 at 16^2 and 80 at 32^2.
    dim_head > 64 used to send the whole call to the fp32 CUDA-core kernel (260 ms per 8-image call); the
    upsample-after-logits kernels do not see dim_head, so the 16^2 layers take the 64-tap kernel and the 32^2 layer the
    128-tap one.  Against fp64 autograd through the oracle on a pixel subset (the gradient is linear in the upstream
    gradient), T = 300 (three token quarters, ragged), 1.5e-2 of each gradient's max; forward rtol 1e-3."""
    from stablekeypoints_b200 import _lib
    spec, T, R, B = [(16, 1280, 8), (32, 640, 8)], 300, 128, 2
    qs, ks, heads, sc = _random_layers(spec, T, B, 41, True)
    g = torch.Generator().manual_seed(8)
    idx = torch.stack([torch.randperm(T, generator=g)[:10] for _ in range(B)])
    pix = [torch.randperm(R * R, generator=g)[:48] for _ in range(B)]
    w = torch.zeros(B, 10, R * R)
    for b in range(B):
        w[b][:, pix[b]] = torch.randn(10, 48, generator=g)
    q64 = [q.double().requires_grad_(True) for q in qs]
    k64 = [k.double().requires_grad_(True) for k in ks]
    loss = 0.0
    for b in range(B):
        m = O.attention_maps_at_pixels([q[b] for q in q64], [k[0] for k in k64], heads, sc, R, pix[b])
        loss = loss + (m[idx[b]] * w[b][:, pix[b]].double()).sum()
    want = torch.autograd.grad(loss, q64 + k64)
    qc = [_cu(q).requires_grad_(True) for q in qs]
    kc = [_cu(k).requires_grad_(True) for k in ks]
    with torch.no_grad():
        full, stats = ops.attention_maps([q.detach() for q in qc], [k.detach() for k in kc], heads, sc, R, return_stats=True)
    for b in range(B):
        ref = O.attention_maps_at_pixels([q[b] for q in qs], [k[0] for k in ks], heads, sc, R, pix[b])
        assert torch.allclose(full[b].reshape(T, -1)[:, pix[b]].cpu().double(), ref, rtol=1e-3, atol=1e-9)
    before = _lib.load().skp_kernel_launches()
    sel = ops.selected_planes(full, _cu(idx), qc, kc, heads, sc, R, row_stats=stats)
    (sel * _cu(w.reshape(B, 10, R, R))).sum().backward()
    for name, got, ref in zip(["dq16", "dq32", "dk16", "dk32"], [t.grad for t in qc + kc], want):
        scale = float(ref.abs().max())
        err = (got.cpu().double() - ref).abs().max().item()
        assert torch.isfinite(got).all() and err <= 1.5e-2 * scale, f"{name}: {err / scale:.2e} of max"


def test_backward_full_size_tcgen05_vs_cuda_core(ops):
    """SD-2.1 head shape (d=64) at T=500, R=128 on a reduced channel count: tcgen05 vs fp32 CUDA cores."""
    spec = [(16, 256, 4), (32, 128, 2)]
    qs, ks, heads, sc = _random_layers(spec, 500, 2, 13, True)
    g = torch.Generator().manual_seed(2)
    idx = torch.stack([torch.randperm(500, generator=g)[:10] for _ in range(2)])
    w = torch.randn(2, 10, 128, 128, generator=g)
    grads = []
    for prec in ("auto", "simt_f32"):
        qc = [_cu(q).requires_grad_(True) for q in qs]
        kc = [_cu(k).requires_grad_(True) for k in ks]
        with torch.no_grad():
            full, stats = ops.attention_maps([q.detach() for q in qc], [k.detach() for k in kc], heads, sc, 128, return_stats=True)
        sel = ops.selected_planes(full, _cu(idx), qc, kc, heads, sc, 128, precision=prec, row_stats=stats if prec == "auto" else None)
        (sel * _cu(w)).sum().backward()
        grads.append([t.grad for t in qc + kc])
    for a, b_ in zip(*grads):
        assert torch.isfinite(a).all()
        assert (a - b_).abs().max().item() <= 1e-2 * float(b_.abs().max())


def test_config4_nine_layers_with_64sq(ops):
    """BASELINE.json configs[3]: capture set extended to 3 x 16^2 + 3 x 32^2 + 3 x 64^2 (SD-2.1 heads 20/10/5, d=64).
    The reference itself cannot stack unequal head counts (optimize.py:75), so the oracle defines the result
    (mean over layers of per-layer head means).  Oracle comparison at a reduced token count, size-independent
    properties and tcgen05 == CUDA-core at the full T=500, R=128; backward through the tcgen05 kernel."""
    spec = [(16, 1280, 20)] * 3 + [(32, 640, 10)] * 3 + [(64, 320, 5)] * 3
    # (i) oracle parity: same nine-layer structure, small T so the CPU oracle finishes in seconds
    small = [(16, 128, 2)] * 3 + [(32, 64, 1)] * 3 + [(64, 64, 1)] * 3        # d = 64 everywhere, unequal heads
    qs, ks, heads, sc = _random_layers(small, 24, 1, 21, True)
    sc = [64 ** -0.5] * len(qs)
    got = ops.attention_maps([_cu(q) for q in qs], [_cu(k) for k in ks], heads, sc, 128, precision="auto").cpu()
    want = O.attention_maps_from_qk([q[0] for q in qs], [k[0] for k in ks], heads, sc, 128)
    assert torch.allclose(got[0], want, rtol=1e-3, atol=1e-7)
    # (ii) full size
    qs, ks, heads, sc = _random_layers(spec, 500, 1, 22, True)
    qc, kc = [_cu(q) for q in qs], [_cu(k) for k in ks]
    with torch.no_grad():
        m, stats = ops.attention_maps(qc, kc, heads, sc, 128, precision="auto", return_stats=True)
    assert m.shape == (1, 500, 128, 128) and stats.shape == (1, sum(heads), 128 * 128, 2)
    assert torch.allclose(m.sum(1), torch.ones(1, 128, 128, device=DEV), atol=2e-5)
    m2 = ops.attention_maps(qc, kc, heads, sc, 128, precision="simt_f32")
    assert torch.allclose(m, m2, rtol=1e-3, atol=1e-7)
    # (iii) backward at full size: finite, and equal to the CUDA-core kernel on the 64^2 layers' gradients
    g = torch.Generator().manual_seed(4)
    idx = torch.randperm(500, generator=g)[:10][None]
    w = torch.randn(1, 10, 128, 128, generator=g)
    grads = []
    for prec in ("auto", "simt_f32"):
        qg = [q.clone().requires_grad_(True) for q in qc]
        kg = [k.clone().requires_grad_(True) for k in kc]
        sel = ops.selected_planes(m, _cu(idx), qg, kg, heads, sc, 128, precision=prec, row_stats=stats if prec == "auto" else None)
        (sel * _cu(w)).sum().backward()
        grads.append([t.grad for t in qg + kg])
    for a, b_ in zip(*grads):
        assert torch.isfinite(a).all()
        assert (a - b_).abs().max().item() <= 1e-2 * float(b_.abs().max())


def test_config5_sdxl_shapes_16_keypoints(ops):
    """BASELINE.json configs[4]: SDXL up path, first four captured cross-attention layers @32^2 (C=1280, h=20, d=64),
    K = 16 keypoints (the UMMA N of the selected-token product).  Parity unpinned by the reference (its SDXL patch is
    dead code): size-independent properties + tcgen05 == CUDA-core, forward and backward, per-image token ids."""
    spec = [(32, 1280, 20)] * 4
    B = 2
    qs, ks, heads, sc = _random_layers(spec, 500, B, 31, True)
    qc, kc = [_cu(q) for q in qs], [_cu(k) for k in ks]
    with torch.no_grad():
        m, stats = ops.attention_maps(qc, kc, heads, sc, 128, precision="auto", return_stats=True)
    assert torch.allclose(m.sum(1), torch.ones(B, 128, 128, device=DEV), atol=2e-5)
    m2 = ops.attention_maps(qc, kc, heads, sc, 128, precision="simt_f32")
    assert torch.allclose(m, m2, rtol=1e-3, atol=1e-7)
    g = torch.Generator().manual_seed(6)
    idx = torch.stack([torch.randperm(500, generator=g)[:16] for _ in range(B)])
    c = ops.attention_maps(qc, kc, heads, sc, 128, token_idx=_cu(idx), precision="auto")
    for b in range(B):
        assert torch.allclose(c[b], m[b, _cu(idx[b])], rtol=1e-4, atol=1e-8)
    w = torch.randn(B, 16, 128, 128, generator=g)
    grads = []
    for prec in ("auto", "simt_f32"):
        qg = [q.clone().requires_grad_(True) for q in qc]
        kg = [k.clone().requires_grad_(True) for k in kc]
        sel = ops.selected_planes(m, _cu(idx), qg, kg, heads, sc, 128, precision=prec, row_stats=stats if prec == "auto" else None)
        (sel * _cu(w)).sum().backward()
        grads.append([t.grad for t in qg + kg])
    for a, b_ in zip(*grads):
        assert torch.isfinite(a).all()
        assert (a - b_).abs().max().item() <= 1e-2 * float(b_.abs().max())
    # readout of the 16 selected planes: bit-exact argmax
    am = ops.plane_argmax(c[0])
    assert torch.equal(am.cpu(), torch.argmax(c[0].reshape(16, -1), dim=-1).cpu())


def test_torch_operator_library_calls_the_same_kernels(ops):
    """`torch.ops.skp_b200.*` == the `ops` functions they wrap, bit for bit."""
    import stablekeypoints_b200.torch_ops  # noqa: F401  (registers the library)
    g = torch.Generator().manual_seed(7)
    spec = [(8, 64, 4), (16, 32, 2)]
    qs = [torch.randn(2, s * s, c, generator=g).to(DEV) for s, c, _ in spec]
    ks = [torch.randn(1, 48, c, generator=g).to(DEV) for _, c, _ in spec]
    heads, sc = [h for _, _, h in spec], [(c // h) ** -0.5 for _, c, h in spec]
    a = torch.ops.skp_b200.attention_maps(qs, ks, heads, sc, 32, None)
    assert torch.equal(a, ops.attention_maps(qs, ks, heads, sc, 32))
    idx = torch.tensor([3, 40, 7], device=DEV)
    assert torch.equal(torch.ops.skp_b200.attention_maps(qs, ks, heads, sc, 32, idx), ops.attention_maps(qs, ks, heads, sc, 32, token_idx=idx))
    planes = a[0].contiguous()
    assert torch.equal(torch.ops.skp_b200.plane_argmax(planes), ops.plane_argmax(planes))
    kl, ent, am = torch.ops.skp_b200.plane_stats(planes, 2.0, 1e-5, 1)
    kl2, ent2, am2 = ops.plane_stats(planes, 2.0, 1e-5, 1, want_kl=True, want_entropy=True, want_argmax=True)
    assert torch.equal(kl, kl2) and torch.equal(ent, ent2) and torch.equal(am, am2)
    assert torch.equal(torch.ops.skp_b200.rank_ascending(kl, 10), ops.rank_ascending(kl, 10))
    assert torch.equal(torch.ops.skp_b200.bilinear_upsample(planes, 64), ops.bilinear_upsample(planes, 64))
