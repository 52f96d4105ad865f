"""CPU tests of the rows SURVEY.md §8 marks "next"/"lowest priority" that need no GPU: the linear regressors and
dataset metrics of `evaluate` (§8(f) rank 2), the SDXL controller API shape (§8(a) row 18) and the helpers the
reference defines but never calls (§8(a) row 19).  Compared with the live reference where it is present
(build container), with closed-form / known-answer checks everywhere."""
import numpy as np
import pytest
import torch

from stablekeypoints_b200 import keypoint_regressor as KR
from stablekeypoints_b200 import optimize as OPT
from stablekeypoints_b200 import sdxl_monkey_patch as SX
from stablekeypoints_b200.eval import keypoint_error, swap_points


def _problem(n=200, k=10, m=5, seed=0, noise=0.01):
    g = np.random.default_rng(seed)
    X = g.uniform(0.1, 0.9, size=(n, 2 * k))
    W = g.normal(size=(2 * k, 2 * m)) * 0.3
    Y = (X - 0.5) @ W + 0.5 + noise * g.normal(size=(n, 2 * m))
    return X, Y, W


def test_return_regressor_is_the_least_squares_solution():
    X, Y, W = _problem()
    got = KR.return_regressor(X, Y)
    want = np.linalg.lstsq(X - 0.5, Y - 0.5, rcond=None)[0]
    assert isinstance(got, np.ndarray) and got.dtype == np.float64
    assert np.allclose(got, want, atol=1e-9)
    assert np.abs(got - W).max() < 0.05
    # tensors in, tensor out, same numbers
    got_t = KR.return_regressor(torch.from_numpy(X), torch.from_numpy(Y))
    assert torch.is_tensor(got_t) and np.allclose(got_t.numpy(), got, atol=1e-12)


def test_return_regressor_visible_fits_each_column_on_its_visible_rows():
    X, Y, _ = _problem(seed=1)
    g = np.random.default_rng(2)
    vis = (g.uniform(size=Y.shape) > 0.3).astype(np.float64)
    Y_bad = Y.copy()
    Y_bad[vis == 0] = 17.0                            # garbage where invisible must not matter
    got = KR.return_regressor_visible(X, Y_bad, vis)
    for j in range(Y.shape[1]):
        rows = vis[:, j] == 1
        want = np.linalg.lstsq(X[rows] - 0.5, Y[rows, j] - 0.5, rcond=None)[0]
        assert np.allclose(got[:, j], want, atol=1e-8)


def test_human36m_regressor_undoes_label_swaps():
    X, Y, _ = _problem(n=300, k=16, m=32, seed=3, noise=0.0)
    Ys = Y.copy()
    flip = np.arange(0, 300, 4)                       # every 4th annotation mirrored
    Ys[flip] = swap_points(torch.from_numpy(Y[flip]).reshape(len(flip), -1, 2)).reshape(len(flip), -1).numpy()
    w_clean = KR.return_regressor_human36m(X, Y)
    w_fixed = KR.return_regressor_human36m(X, Ys)
    # the loop stops once at most 10 samples still prefer their mirrored labels (the reference's rule), so the
    # fit is not exact - but it must be far better than plain least squares on the mislabelled data
    t = (Y - 0.5).reshape(300, -1, 2)
    err = lambda w: np.linalg.norm(((X - 0.5) @ w).reshape(300, -1, 2) - t, axis=2).mean()
    assert err(w_fixed) < 0.25 * err(KR.return_regressor(X, Ys))
    assert np.allclose(np.abs((X - 0.5) @ w_clean - (Y - 0.5)).max(), 0, atol=1e-8)


def test_swap_points_table():
    p = torch.arange(32 * 2, dtype=torch.float32).reshape(1, 32, 2)
    s = swap_points(p)
    assert torch.equal(s[0, 1], p[0, 6]) and torch.equal(s[0, 6], p[0, 1])
    assert torch.equal(s[0, 17], p[0, 25]) and torch.equal(s[0, 0], p[0, 0])
    # the reference's table maps 20 <-> 28 and then 21 <-> 28: the later pair wins for 28
    assert torch.equal(s[0, 21], p[0, 28]) and torch.equal(s[0, 28], p[0, 21]) and torch.equal(s[0, 20], p[0, 28])


def test_keypoint_error_known_answers():
    gt = torch.tensor([[0.25, 0.25], [0.25, 0.75], [0.5, 0.5]])
    est = gt + torch.tensor([[0.0, 0.05], [0.0, 0.0], [0.03, 0.04]])
    eye = 0.5
    assert torch.isclose(keypoint_error(est, gt, "inter_eye_distance"), torch.tensor((0.05 + 0.0 + 0.05) / 3 / eye))
    vis = torch.tensor([1.0, 0.0, 1.0])
    assert torch.isclose(keypoint_error(est, gt, "visible", vis), torch.tensor(0.05))
    assert torch.isclose(keypoint_error(est, gt, "mean_average_error", vis), torch.tensor(0.1 * 256))
    assert torch.isclose(keypoint_error(est, gt, "pck"), torch.tensor(1 / 3))        # 12.8 px, 0 px, 12.8 px vs 6 px
    with pytest.raises(NotImplementedError):
        keypoint_error(est, gt, "nope")
    g = torch.Generator().manual_seed(0)
    gt32 = torch.rand(32, 2, generator=g)
    est32 = gt32 + 0.01 * torch.randn(32, 2, generator=g)
    assert torch.isclose(keypoint_error(gt32, gt32, "orientation_invariant"), torch.tensor(0.0))
    want = min((est32 - gt32).norm(dim=-1).mean(), (swap_points(est32[None])[0] - gt32).norm(dim=-1).mean()) * 128
    assert torch.isclose(keypoint_error(est32, gt32, "orientation_invariant"), want)


# ------------------------------------------------------------------ live reference (build container only)
@pytest.mark.reference
@pytest.mark.parametrize("method,n", [("inter_eye_distance", 5), ("visible", 7), ("mean_average_error", 7), ("pck", 9),
                                      ("orientation_invariant", 32)])
def test_keypoint_error_matches_oracle_restatement_of_evaluate(method, n):
    """`eval.keypoint_error` + the regressor application of `evaluate` against the oracle's line-by-line restatement of
    `eval.py:468-510`, random keypoints / regressors / visibility, all five metrics (32 keypoints: Human3.6M swap table)."""
    from oracle import keypoint_oracle as O
    g = torch.Generator().manual_seed(n)
    for trial in range(6):
        K = 10
        kp = torch.rand(K, 2, generator=g)
        W = torch.randn(2 * K, 2 * n, generator=g) * 0.2
        gt = torch.rand(n, 2, generator=g)
        vis = (torch.rand(n, generator=g) > 0.3).float() if method in ("visible", "mean_average_error") and trial % 2 else None
        if vis is not None and vis.sum() == 0:
            vis[0] = 1.0
        want = O.evaluate_image(kp, W, gt, method, vis)
        est = (((kp.reshape(1, -1) - 0.5) @ W) + 0.5).reshape(-1, 2)          # eval.evaluate's regressor application
        got = keypoint_error(est, gt, method, vis)
        assert torch.allclose(got, want, rtol=1e-6, atol=1e-7), (method, trial, float(got), float(want))


def test_oracle_swap_table_matches_reference(ref):
    from oracle import keypoint_oracle as O
    pts = torch.arange(32 * 2, dtype=torch.float32).reshape(1, 32, 2)
    assert torch.equal(O.swap_points_h36m(pts), ref.eval.swap_points(pts))
    assert torch.equal(swap_points(pts), ref.eval.swap_points(pts))


def test_regressors_match_reference(ref):
    X, Y, _ = _problem(seed=5)
    assert np.allclose(KR.return_regressor(X, Y), ref.keypoint_regressor.return_regressor(X, Y), atol=1e-10)
    vis = (np.random.default_rng(6).uniform(size=Y.shape) > 0.2).astype(np.float64)
    assert np.allclose(KR.return_regressor_visible(X, Y, vis), ref.keypoint_regressor.return_regressor_visible(X, Y, vis), atol=1e-10)


@pytest.mark.reference
def test_unused_helpers_match_reference(ref):
    g = torch.Generator().manual_seed(1)
    m = torch.rand(6, 16, 16, generator=g)
    assert torch.allclose(OPT.create_gaussian_kernel(5, 1.0), ref.optimize.create_gaussian_kernel(5, 1.0), atol=1e-7)
    assert torch.allclose(OPT.differentiable_argmax(m), ref.optimize.differentiable_argmax(m), atol=1e-5)
    assert torch.allclose(OPT.variance_loss(m), ref.optimize.variance_loss(m), atol=1e-6)
    a = m.clone().requires_grad_(True)
    b = m.clone().requires_grad_(True)
    la, lb = OPT.gaussian_loss(a, temperature=1e-2), ref.optimize.gaussian_loss(b, temperature=1e-2)
    assert torch.allclose(la, lb, atol=1e-7)
    la.backward(), lb.backward()
    assert torch.allclose(a.grad, b.grad, atol=1e-8)
    p = torch.rand(2, 32, 2, generator=g)
    assert torch.equal(swap_points(p), ref.eval.swap_points(p))


# ------------------------------------------------------------------ SDXL controller API shape
def test_sdxl_store_protocol_with_tensors():
    st = SX.AttentionStore()
    st.num_att_layers = 2
    a = torch.rand(4, 16, 7)                      # (2 x (uncond, cond) x heads, queries, tokens)
    a0 = a.clone()                                # the store keeps VIEWS of the caller's tensor and sums in place, like the reference
    big = torch.rand(4, 64 ** 2, 3)
    out = st(a, True, "up")
    assert out is a and len(st.step_store["up_cross"]) == 1 and st.step_store["up_cross"][0].shape[0] == 2
    assert st.cur_att_layer == 1 and st.cur_step == 0
    st(big, False, "down")                        # over the 32**2 gate: not kept, but counted; ends the step
    assert len(st.step_store["down_self"]) == 0
    assert st.cur_att_layer == 0 and st.cur_step == 1
    assert len(st.attention_store["up_cross"]) == 1            # between_steps moved the step store
    st(a, True, "up"); st(big, False, "down")
    avg = st.get_average_attention()
    assert torch.allclose(avg["up_cross"][0], a0[2:])            # (a + a) / 2 steps
    st.reset()
    assert st.cur_step == 0 and st.attention_store == {} and st.step_store["mid_cross"] == []


def test_sdxl_registration_captures_first_four_up_layers_and_feeds_collect_maps():
    from stablekeypoints_b200 import ptp_utils
    from stablekeypoints_b200.sd_standin import UNet2DConditionStandIn, unet_config
    torch.manual_seed(0)
    cfg = unet_config("tiny-sd21")
    unet = UNet2DConditionStandIn(cfg)
    st = SX.AttentionStore()
    st.stop_after_capture = True
    SX.register_attention_control(unet, st, feature_upsample_res=32)
    x = torch.randn(1, 4, cfg.latent_side, cfg.latent_side)
    ctx = torch.randn(1, 9, cfg.context_dim)
    with pytest.raises(ptp_utils.StopForward):
        unet(x, torch.zeros(1, dtype=torch.int64), ctx)
    assert len(st.step_store["up_cross"]) == 4 and st.step_store["down_cross"] == []
    assert st.step_store["attn"] is st.step_store["up_cross"]          # what collect_maps reads
    h = st.step_store["attn"][0]
    assert isinstance(h, ptp_utils.CapturedAttention) and h.shape == (h.heads, 32 ** 2, 9)
    with pytest.raises(AssertionError):
        SX.register_attention_control(torch.nn.Sequential(torch.nn.Linear(2, 2)), st)
