"""Oracle vs the committed golden vectors (generated from the live reference by
oracle/make_golden.py).  CPU only; runs in the container and on the GPU box."""
import os

import numpy as np
import torch

from oracle import keypoint_oracle as O

G = os.path.join(os.path.dirname(__file__), "golden")


def _load(name):
    return {k: v for k, v in np.load(os.path.join(G, name)).items()}


def test_attention_maps_golden():
    d = _load("attention_maps.npz")
    R = int(d["R"])
    heads = [int(h) for h in d["heads"]]
    qs = [torch.from_numpy(d[f"q_low{i}"]) for i in range(4)]
    ks = [torch.from_numpy(d[f"k{i}"]) for i in range(4)]
    sc = [float(s) for s in d["scale"]]
    idx = torch.from_numpy(d["indices"])
    A = O.attention_maps_from_qk(qs, ks, heads, sc, R)
    assert A.shape == (100, R, R)
    assert torch.allclose(A, torch.from_numpy(d["ref_maps_A"]), rtol=3e-4, atol=1e-7)
    # every pixel's map values sum to 1 over tokens (softmax over T then linear averaging)
    assert torch.allclose(A.sum(0), torch.ones(R, R), atol=1e-5)
    B = O.attention_maps_from_qk(qs, ks, heads, sc, R, upsample_res=128)
    assert torch.allclose(B[:8], torch.from_numpy(d["ref_maps_B"]), rtol=3e-4, atol=1e-7)
    C = O.attention_maps_from_qk(qs, ks, heads, sc, R, upsample_res=128, indices=idx)
    assert torch.allclose(C, torch.from_numpy(d["ref_maps_C"]), rtol=3e-4, atol=1e-7)


def test_selection_readout_golden():
    d = _load("selection_readout.npz")
    m = torch.from_numpy(d["maps"])
    assert np.array_equal(O.argmax_pixel(m).numpy(), d["ref_max_pixel"])
    assert np.array_equal(O.k_max_pixels(m, 3).numpy(), d["ref_k_max_pixels3"])
    assert np.allclose(O.weighted_avg_pixel(m).numpy(), d["ref_weighted_avg"], atol=1e-4)
    for ns in (1, 2):
        assert np.array_equal(O.gaussian_kl_rank(m, 25, 2.0, 1e-5, ns).numpy(), d[f"ref_topk_gaussian_ns{ns}"])
    assert np.array_equal(O.entropy_rank(m, 25).numpy(), d["ref_entropy_sort"])
    cand = torch.from_numpy(d["ref_topk_gaussian_ns1"])
    assert np.array_equal(O.furthest_points(m, 10, cand).numpy(), d["ref_fps"])
    pos = O.k_max_pixels(m[:6], 2) / 64
    assert np.array_equal(O.gaussian_targets(pos, 64, 2.0).numpy(), d["ref_gaussian_circles"])
    m48 = torch.from_numpy(d["maps48"])
    assert np.array_equal(O.furthest_points(m48, 10, torch.arange(25)).numpy(), d["ref_fps48"])


def test_losses_tta_golden():
    d = _load("losses_tta.npz")
    img, theta = torch.from_numpy(d["img"]), torch.from_numpy(d["theta"])
    w = O.affine_warp(img, theta)
    assert torch.allclose(w, torch.from_numpy(d["ref_warped"]), atol=2e-5)
    assert torch.allclose(O.affine_warp(w, O.invert_theta(theta)), torch.from_numpy(d["ref_unwarped"]), atol=2e-5)
    a = torch.from_numpy(d["a"]).requires_grad_(True)
    b = torch.from_numpy(d["b"]).requires_grad_(True)
    ls = O.sharpening_loss(a, 2.0)
    le = O.equivariance_loss(a, b, theta[0])
    assert abs(ls.item() - float(d["ref_sharpening"])) < 1e-6 * abs(float(d["ref_sharpening"])) + 1e-9
    assert abs(le.item() - float(d["ref_equivariance"])) < 1e-5 * abs(float(d["ref_equivariance"])) + 1e-9
    ga, gb = torch.autograd.grad(ls * 100 + le * 1000, [a, b])
    for got, want in ((ga, d["ref_grad_a"]), (gb, d["ref_grad_b"])):
        assert np.allclose(got.numpy(), want, rtol=1e-4, atol=1e-5 * np.abs(want).max())
    ens = O.tta_ensemble(list(torch.from_numpy(d["tta_maps"])), list(torch.from_numpy(d["tta_thetas"])))
    assert np.allclose(ens.numpy(), d["ref_tta"], rtol=1e-4, atol=1e-5)
    assert np.array_equal(O.argmax_pixel(ens).numpy(), d["ref_tta_keypoints"])


def test_known_answers():
    """Analytic pins derivable from the reference code."""
    m = torch.zeros(3, 16, 16)
    m[0, 2, 9] = 1; m[1, 15, 0] = 2; m[2, 7, 7] = 0.5
    assert O.argmax_pixel(m).tolist() == [[2.5, 9.5], [15.5, 0.5], [7.5, 7.5]]
    g = O.gaussian_targets((O.argmax_pixel(m) / 16)[None], 16, 2.0)
    assert g[0, 2, 9] == 1.0 and g[1, 15, 0] == 1.0
    eye = torch.tensor([[1.0, 0, 0], [0, 1.0, 0]])
    x = torch.rand(1, 2, 8, 8)
    assert torch.allclose(O.affine_warp(x, eye[None]), x, atol=1e-6)
    assert O.equivariance_loss(x[0], x[0], eye).item() < 1e-12
    # FPS on collinear candidates returns the two extremes first
    line = torch.zeros(5, 16, 16)
    for i, c in enumerate([3, 0, 8, 15, 5]):
        line[i, 4, c] = 1
    assert O.furthest_points(line, 3, torch.arange(5)).tolist()[:2] == [1, 3]


def test_pixel_restricted_oracle_equals_full_oracle():
    """`attention_maps_at_pixels` (the fp64 pin of the BASELINE-size GPU tests) == `attention_maps_from_qk` on those pixels,
    unequal heads per layer, and its log2-sum-exp2 equals the logits' own."""
    import torch
    from oracle import keypoint_oracle as O
    g = torch.Generator().manual_seed(0)
    spec = [(4, 64, 4), (8, 32, 2), (16, 32, 1)]
    qs = [torch.randn(s * s, c, generator=g).double() for s, c, _ in spec]
    ks = [torch.randn(23, c, generator=g).double() for _, c, _ in spec]
    heads = [h for _, _, h in spec]
    sc = [(c // h) ** -0.5 for _, c, h in spec]
    full = O.attention_maps_from_qk(qs, ks, heads, sc, 16).reshape(23, -1)
    pix = torch.tensor([0, 5, 17, 255, 128, 15, 16, 240])
    sub, lse2 = O.attention_maps_at_pixels(qs, ks, heads, sc, 16, pix, return_lse2=True)
    assert torch.allclose(sub, full[:, pix], rtol=1e-12, atol=1e-15)
    assert lse2.shape == (sum(heads), len(pix))
    probs = O.layer_probs_from_qk(qs[0][None], ks[0][None], heads[0], 16, sc[0])       # (h, R^2, T)
    up = O.bicubic_upsample(qs[0].reshape(1, 4, 4, 64).permute(0, 3, 1, 2), 16).permute(0, 2, 3, 1).reshape(1, 256, 64)
    sim = torch.einsum("bid,bjd->bij", O.split_heads(up, 4), O.split_heads(ks[0][None], 4)) * sc[0]
    assert torch.allclose(lse2[:4], torch.logsumexp(sim, -1)[:, pix] * 1.4426950408889634, rtol=1e-12)
    assert torch.allclose(torch.exp2(sim[:, pix] * 1.4426950408889634 - lse2[:4, :, None]), probs[:, pix], rtol=1e-10)
