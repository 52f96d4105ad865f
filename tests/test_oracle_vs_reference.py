"""Pins the oracle (oracle/keypoint_oracle.py) against the UNMODIFIED reference imported live.

Runs only where /root/reference exists (the build container).  The same comparisons are
frozen into tests/golden/*.npz by oracle/make_golden.py so they also hold on the GPU box.
"""
import math

import pytest
import torch
import torch.nn.functional as F

from oracle import keypoint_oracle as O

pytestmark = pytest.mark.reference


def _maps(n, size, seed, peaky=True):
    g = torch.Generator().manual_seed(seed)
    m = torch.rand(n, size, size, generator=g) * 0.2
    if peaky:
        for i in range(n):
            r, c = torch.randint(0, size, (2,), generator=g)
            rows = torch.arange(size)[:, None].float()
            cols = torch.arange(size)[None, :].float()
            m[i] += torch.exp(-((rows - r) ** 2 + (cols - c) ** 2) / (2 * (1.0 + 3 * torch.rand(1, generator=g)) ** 2))
    return m


@pytest.mark.parametrize("s,R", [(4, 32), (8, 32), (16, 128), (32, 128), (64, 128), (8, 8)])
def test_bicubic_matches_aten(ref, s, R):
    x = torch.randn(2, 3, s, s, dtype=torch.float64)
    want = F.interpolate(x, size=(R, R), mode="bicubic", align_corners=False)
    assert torch.allclose(O.bicubic_upsample(x, R), want, atol=1e-12)


@pytest.mark.parametrize("s,R", [(32, 64), (128, 256), (128, 512), (16, 16)])
def test_bilinear_matches_aten(ref, s, R):
    x = torch.randn(2, 3, s, s, dtype=torch.float64)
    want = F.interpolate(x, size=(R, R), mode="bilinear", align_corners=False)
    assert torch.allclose(O.bilinear_upsample(x, R), want, atol=1e-12)


def test_affine_warp_and_inverse_match_reference(ref):
    T = ref.invertable_transform.RandomAffineWithInverse(degrees=15, scale=[0.8, 1.0], translate=[0.25, 0.25])
    torch.manual_seed(3)
    img = torch.rand(3, 5, 32, 32)
    out = T(img)
    theta = T.last_params["theta"]
    assert torch.allclose(O.affine_warp(img, theta), out, atol=2e-6)
    back = T.inverse(out)
    assert torch.allclose(O.affine_warp(out, O.invert_theta(theta)), back, atol=2e-6)
    # create_affine_matrix restated
    th = T.create_affine_matrix(12.5, 0.9, (0.1, -0.2))[0]
    assert torch.equal(O.affine_theta(12.5, 0.9, 0.1, -0.2), th)


def test_captured_probs_and_collect_maps_match_reference_hook(ref):
    """Reference hook + collect_maps on the stand-in UNet vs the oracle fed with the same
    layer inputs (captured with forward hooks on the attention modules)."""
    from stablekeypoints_b200 import sd_standin as S
    cfg = S.unet_config("tiny-sd15")
    pipe = S.PipelineStandIn(cfg, vae_channels=(32, 32, 64, 64), seed=0)
    R = 32
    g = torch.Generator().manual_seed(1)
    ctx = torch.randn(1, 24, cfg.context_dim, generator=g)
    lat = torch.randn(1, 4, cfg.latent_side, cfg.latent_side, generator=g)
    ctl = ref.ptp_utils.AttentionStore()
    ref.ptp_utils.register_attention_control(pipe.unet, ctl, feature_upsample_res=R)
    seen = []
    mods = [m for n, m in pipe.unet.up_blocks.named_modules() if n.endswith("attn2")]
    orig = {}
    for m in mods:
        orig[m] = m.forward
        def wrap(x, context=None, mask=None, _m=m):
            seen.append((_m, x.detach().clone()))
            return orig[_m](x, context=context, mask=mask)
        m.forward = wrap
    with torch.no_grad():
        pipe.unet(lat, torch.tensor(0), ctx)
    stored = list(ctl.step_store["attn"])
    assert len(stored) == 4
    ours = []
    for (m, x), want in zip(seen[:4], stored):
        got = O.captured_layer_probs(x, ctx, m.to_q.weight, m.to_k.weight, m.heads, R)
        assert got.shape == want.shape
        assert torch.allclose(got, want, rtol=2e-4, atol=1e-7)
        # commuted form used by the CUDA path
        got2 = O.layer_probs_from_qk(m.to_q(x), m.to_k(ctx), m.heads, R, m.scale)
        assert torch.allclose(got2, want, rtol=2e-4, atol=1e-7)
        ours.append(got[: m.heads])
    for up, idx in [(-1, None), (64, None), (128, torch.tensor([3, 7, 11]))]:
        ctl.step_store["attn"] = [s.clone() for s in stored]
        want = ref.optimize.collect_maps(ctl, upsample_res=up, layers=[0, 1, 2, 3], indices=idx)
        got = O.collect_maps([s for s in stored], (0, 1, 2, 3), up, idx)
        assert torch.allclose(got, want, rtol=1e-5, atol=1e-8)
        got_pl = O.collect_maps([s for s in stored], (0, 1, 2, 3), up, idx, per_layer_head_mean=True)
        assert torch.allclose(got_pl, want, rtol=1e-5, atol=1e-8)


@pytest.mark.parametrize("size", [16, 128])
def test_readout_matches_reference(ref, size):
    m = _maps(12, size, 5)
    assert torch.equal(O.argmax_pixel(m), ref.eval.find_max_pixel(m))
    assert torch.equal(O.k_max_pixels(m, 3), ref.eval.find_k_max_pixels(m, 3))
    p = ref.eval.find_max_pixel(m)
    assert torch.equal(O.mask_disc(m, p, 0.05 * size), ref.eval.mask_radius(m, p, 0.05 * size))
    assert torch.allclose(O.weighted_avg_pixel(m), ref.eval.pixel_from_weighted_avg(m.clone()), atol=1e-4)
    assert torch.allclose(O.weighted_avg_pixel(m, -1), ref.eval.pixel_from_weighted_avg(m.clone(), -1), atol=1e-3)
    # ties: first index wins
    t = torch.zeros(2, size, size)
    t[0, 3, 4] = 1; t[0, 5, 1] = 1
    assert torch.equal(O.argmax_pixel(t), ref.eval.find_max_pixel(t))


def test_gaussians_and_selection_match_reference(ref):
    size = 64
    m = _maps(40, size, 7)
    pos = ref.eval.find_k_max_pixels(m, 2) / size
    want = ref.optimize_token.gaussian_circles(pos, size=size, sigma=2.0, device="cpu")
    assert torch.equal(O.gaussian_targets(pos, size, 2.0), want)
    for ns in (1, 2):
        want = ref.ptp_utils.find_top_k_gaussian(m, 25, sigma=2.0, num_subjects=ns)
        assert torch.equal(O.gaussian_kl_rank(m, 25, 2.0, 1e-5, ns), want)
    want = ref.ptp_utils.entropy_sort(m, 25)
    assert torch.equal(O.entropy_rank(m, 25), want)
    cand = ref.ptp_utils.find_top_k_gaussian(m, 25, sigma=2.0)
    want = ref.ptp_utils.furthest_point_sampling(m, 10, cand)
    assert torch.equal(O.furthest_points(m, 10, cand), want)
    # non-power-of-two size exercises the fp32 location rounding
    m2 = _maps(30, 48, 9)
    cand = torch.arange(25)
    assert torch.equal(O.furthest_points(m2, 10, cand), ref.ptp_utils.furthest_point_sampling(m2, 10, cand))


def test_losses_match_reference_with_gradients(ref):
    size = 32
    T = ref.invertable_transform.RandomAffineWithInverse(degrees=15, scale=[0.8, 1.0], translate=[0.25, 0.25])
    torch.manual_seed(11)
    T(torch.rand(1, 3, size, size))
    theta = T.last_params["theta"]
    a = _maps(10, size, 1).requires_grad_(True)
    b = _maps(10, size, 2).requires_grad_(True)
    want_s = ref.optimize.sharpening_loss(a, sigma=2.0, device="cpu")
    want_e = ref.optimize.equivariance_loss(a, b[None], T, 0)
    gw = torch.autograd.grad(want_s * 100 + want_e * 1000, [a, b])
    a2 = a.detach().clone().requires_grad_(True)
    b2 = b.detach().clone().requires_grad_(True)
    got_s = O.sharpening_loss(a2, 2.0)
    got_e = O.equivariance_loss(a2, b2, theta[0])
    assert torch.allclose(got_s, want_s, rtol=1e-6)
    assert torch.allclose(got_e, want_e, rtol=1e-5)
    gg = torch.autograd.grad(got_s * 100 + got_e * 1000, [a2, b2])
    for x, y in zip(gg, gw):
        assert torch.allclose(x, y, rtol=1e-4, atol=1e-5 * float(y.abs().max()))


def test_tta_ensemble_matches_reference_arithmetic(ref):
    size = 32
    T = ref.invertable_transform.RandomAffineWithInverse(degrees=15, scale=[0.8, 1.0], translate=[0.25, 0.25])
    torch.manual_seed(5)
    num = torch.zeros(4, size, size); tot = torch.zeros(4, size, size)
    maps, thetas = [], []
    for a in range(5):
        T(torch.rand(1, 3, size, size))
        m = _maps(4, size, 20 + a)
        maps.append(m); thetas.append(T.last_params["theta"][0])
        num += T.inverse(torch.ones_like(m)[None]).sum(0)   # eval.py:262-266
        tot += T.inverse(m[None]).sum(0)
    want = tot / num
    want[want != want] = 0
    assert torch.allclose(O.tta_ensemble(maps, thetas), want, rtol=1e-4, atol=1e-5)


def test_ddim_constants(ref):
    from stablekeypoints_b200.sd_standin import DDIMSchedulerStandIn
    ac = O.ddim_alphas_cumprod()
    assert abs(float(ac[0]) ** 0.5 - 0.99957) < 1e-4 and abs((1 - float(ac[0])) ** 0.5 - 0.02916) < 1e-4
    sch = DDIMSchedulerStandIn()
    assert int(sch.timesteps[-1]) == 0 and int(sch.timesteps[0]) == 980
