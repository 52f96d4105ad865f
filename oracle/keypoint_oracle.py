"""CPU oracle for the StableKeypoints cross-attention keypoint hot path.

TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import this package.  The product
path (`stablekeypoints_b200/`) never does; it fails loudly when the CUDA library is missing.

What this is: a restatement, in explicit formulas on torch-CPU tensors (fp32 or fp64), of
the reference algorithm for every row of SURVEY.md §8(a).  Each function cites the
reference `file:line` (relative to /root/reference) it follows.  It deliberately does NOT
call `F.interpolate`, `F.affine_grid` or `F.grid_sample`: the resampling arithmetic the
CUDA kernels implement is written out here, so that the oracle pins
the formulas and not a library call.

Parity pin: the reference ships no tests and no golden vectors for this path.
The oracle is pinned instead (a) against the UNMODIFIED reference imported live in the build
container (`tests/test_oracle_vs_reference.py`, via `oracle/ref_shim.py`), and (b) against
golden vectors generated from that same live reference by `oracle/make_golden.py` and
committed under `tests/golden/` (these travel to the GPU box).  For shapes the reference
cannot run (unequal head counts per layer = SD-2.1 / SDXL, 64x64 capture), parity is
"unpinned by the reference": the oracle restates the same algorithm with a per-layer head
mean, and says so in `collect_maps`.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import torch

Tensor = torch.Tensor


# ======================================================================================
# resampling primitives
# ======================================================================================
def _cubic_taps(t: Tensor, A: float = -0.75) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Keys cubic-convolution weights for fractional offset t (ATen upsample_bicubic2d)."""
    def near(u):   # |u| <= 1
        return ((A + 2.0) * u - (A + 3.0)) * u * u + 1.0

    def far(u):    # 1 < |u| < 2
        return ((A * u - 5.0 * A) * u + 8.0 * A) * u - 4.0 * A

    return far(t + 1.0), near(t), near(1.0 - t), far(2.0 - t)


def bicubic_matrix(n_in: int, n_out: int, dtype=torch.float64) -> Tensor:
    """(n_out, n_in) matrix M with up = M @ x along one axis.

    Follows `F.interpolate(mode="bicubic", align_corners=False)` as used at
    `unsupervised_keypoints/ptp_utils.py:520-526`: src = (dst+0.5)*(in/out)-0.5 (not
    clamped), taps floor(src)-1..+2 with indices clamped to the border.
    """
    M = torch.zeros(n_out, n_in, dtype=dtype)
    scale = n_in / n_out
    for o in range(n_out):
        src = (o + 0.5) * scale - 0.5
        i0 = math.floor(src)
        t = torch.tensor(src - i0, dtype=dtype)
        for tap, w in zip(range(i0 - 1, i0 + 3), _cubic_taps(t)):
            M[o, min(max(tap, 0), n_in - 1)] += w
    return M


def bilinear_matrix(n_in: int, n_out: int, dtype=torch.float64) -> Tensor:
    """(n_out, n_in) matrix for `F.interpolate(mode="bilinear", align_corners=False)`
    as used at `unsupervised_keypoints/optimize.py:63-70`: src = max((dst+0.5)*(in/out)-0.5, 0)."""
    M = torch.zeros(n_out, n_in, dtype=dtype)
    scale = n_in / n_out
    for o in range(n_out):
        src = max((o + 0.5) * scale - 0.5, 0.0)
        i0 = min(int(math.floor(src)), n_in - 1)
        i1 = min(i0 + 1, n_in - 1)
        lam = src - i0
        M[o, i0] += 1.0 - lam
        M[o, i1] += lam
    return M


def resize_separable(x: Tensor, My: Tensor, Mx: Tensor) -> Tensor:
    """x (..., h, w) -> (..., H, W) with row matrix My (H,h) and column matrix Mx (W,w)."""
    return torch.einsum("Hh,...hw,Ww->...HW", My.to(x.dtype), x, Mx.to(x.dtype))


def bicubic_upsample(x: Tensor, size: int) -> Tensor:
    h, w = x.shape[-2:]
    return resize_separable(x, bicubic_matrix(h, size, x.dtype), bicubic_matrix(w, size, x.dtype))


def bilinear_upsample(x: Tensor, size: int) -> Tensor:
    h, w = x.shape[-2:]
    return resize_separable(x, bilinear_matrix(h, size, x.dtype), bilinear_matrix(w, size, x.dtype))


def affine_theta(angle_deg: float, scale: float, tx: float, ty: float) -> Tensor:
    """`RandomAffineWithInverse.create_affine_matrix`, `invertable_transform.py:22-36`:
    built in fp32, rotation part scaled in fp32."""
    a = math.radians(angle_deg)
    th = torch.tensor([[math.cos(a), math.sin(a), tx], [-math.sin(a), math.cos(a), ty]], dtype=torch.float32)
    th[:, :2] = th[:, :2] * scale
    return th


def invert_theta(theta: Tensor) -> Tensor:
    """Inverse of the augmented 3x3 (`invertable_transform.py:77-84`). theta (..., 2, 3)."""
    a, b, c = theta[..., 0, 0], theta[..., 0, 1], theta[..., 0, 2]
    d, e, f = theta[..., 1, 0], theta[..., 1, 1], theta[..., 1, 2]
    det = a * e - b * d
    inv = torch.stack([
        torch.stack([e / det, -b / det, (b * f - c * e) / det], dim=-1),
        torch.stack([-d / det, a / det, (c * d - a * f) / det], dim=-1)], dim=-2)
    return inv


def affine_warp(img: Tensor, theta: Tensor) -> Tensor:
    """`F.grid_sample(img, F.affine_grid(theta, img.size(), align_corners=False),
    align_corners=False)` with the defaults bilinear / zeros
    (`invertable_transform.py:64-68, 87-90`).  img (N,C,H,W), theta (N,2,3)."""
    N, C, H, W = img.shape
    dt = img.dtype
    xs = (2.0 * torch.arange(W, dtype=dt) + 1.0) / W - 1.0
    ys = (2.0 * torch.arange(H, dtype=dt) + 1.0) / H - 1.0
    th = theta.to(dt)
    gx = th[:, 0, 0, None, None] * xs[None, None, :] + th[:, 0, 1, None, None] * ys[None, :, None] + th[:, 0, 2, None, None]
    gy = th[:, 1, 0, None, None] * xs[None, None, :] + th[:, 1, 1, None, None] * ys[None, :, None] + th[:, 1, 2, None, None]
    ix = ((gx + 1.0) * W - 1.0) / 2.0
    iy = ((gy + 1.0) * H - 1.0) / 2.0
    x0 = torch.floor(ix)
    y0 = torch.floor(iy)
    out = torch.zeros_like(img)
    flat = img.reshape(N, C, H * W)
    for dy in (0, 1):
        for dx in (0, 1):
            xi = x0 + dx
            yi = y0 + dy
            w = (1.0 - (ix - xi).abs()) * (1.0 - (iy - yi).abs())
            ok = (xi >= 0) & (xi <= W - 1) & (yi >= 0) & (yi <= H - 1)
            idx = (yi.clamp(0, H - 1) * W + xi.clamp(0, W - 1)).long().reshape(N, 1, H * W).expand(N, C, H * W)
            vals = torch.gather(flat, 2, idx).reshape(N, C, H, W)
            out = out + vals * (w * ok.to(dt))[:, None]
    return out


# ======================================================================================
# row 1: captured cross-attention probabilities  (ptp_utils.py:508-538)
# ======================================================================================
def split_heads(t: Tensor, heads: int) -> Tensor:
    """(B, N, h*d) -> (B*h, N, d), batch-major / head-minor (diffusers 0.8.0 contract)."""
    b, n, c = t.shape
    return t.reshape(b, n, heads, c // heads).permute(0, 2, 1, 3).reshape(b * heads, n, c // heads)


def captured_layer_probs(x: Tensor, context: Tensor, w_q: Tensor, w_k: Tensor, heads: int,
                         feature_res: int) -> Tensor:
    """The tensor the reference hands to the controller for one captured layer.

    x (B, s*s, C) hidden state entering the cross-attention; context (B, T, ctx);
    w_q (C, C) / w_k (C, ctx) the bias-free `to_q` / `to_k` weights.
    `ptp_utils.py:513-538`: x -> (B,C,s,s) -> bicubic to feature_res -> (B,R^2,C) -> to_q ->
    heads -> (q.k^T)*scale -> softmax over TOKENS.  Returns (B*h, R^2, T).
    """
    B, n, C = x.shape
    s = int(round(n ** 0.5))
    img = x.reshape(B, s, s, C).permute(0, 3, 1, 2)
    up = bicubic_upsample(img, feature_res).permute(0, 2, 3, 1).reshape(B, feature_res * feature_res, C)
    q = split_heads(up @ w_q.t(), heads)
    k = split_heads(context @ w_k.t(), heads)
    scale = (C // heads) ** -0.5
    sim = torch.einsum("bid,bjd->bij", q, k) * scale
    return torch.softmax(sim, dim=-1)


def layer_probs_from_qk(q_low: Tensor, k: Tensor, heads: int, feature_res: int, scale: float) -> Tensor:
    """Same quantity from the already-projected low-resolution q (B, s*s, C) and k (B, T, C).

    Uses the identity to_q(upsample(x)) == upsample(to_q(x)) (to_q is linear and bias-free,
    bicubic interpolation is linear).  This is the formulation the CUDA path
    consumes: its inputs are the q/k the module computes anyway at `ptp_utils.py:483-490`.
    """
    B, n, C = q_low.shape
    s = int(round(n ** 0.5))
    up = bicubic_upsample(q_low.reshape(B, s, s, C).permute(0, 3, 1, 2), feature_res)
    q = split_heads(up.permute(0, 2, 3, 1).reshape(B, feature_res * feature_res, C), heads)
    sim = torch.einsum("bid,bjd->bij", q, split_heads(k, heads)) * scale
    return torch.softmax(sim, dim=-1)


# ======================================================================================
# row 3: collect_maps  (optimize.py:27-79)
# ======================================================================================
def collect_maps(stored: Sequence[Tensor], layers: Sequence[int] = (0, 1, 2, 3), upsample_res: int = -1,
                 indices: Optional[Tensor] = None, per_layer_head_mean: Optional[bool] = None) -> Tensor:
    """stored[l] is (h_l, R^2, T) for ONE image (`optimize.py:52-54` averages dim 0, so the
    reference is only meaningful with one image per controller: SURVEY.md §0.4).

    Reference semantics (`optimize.py:44-75`): reshape to (h,R,R,T); optional token gather;
    permute to (h,T',R,R); bilinear upsample unless upsample_res == -1; stack over layers;
    mean over (layer, head) jointly.  With equal head counts that joint mean equals the mean
    over layers of per-layer head means, which is how unequal head counts (SD-2.1, SDXL —
    where the reference's `torch.stack` raises) are defined here: "parity unpinned by the
    reference" for those shapes.
    """
    picked = [stored[l] for l in range(len(stored)) if l in layers]
    heads = {p.shape[0] for p in picked}
    if per_layer_head_mean is None:
        per_layer_head_mean = len(heads) > 1
    outs = []
    for data in picked:
        h, n, T = data.shape
        R = int(round(n ** 0.5))
        d = data.reshape(h, R, R, T)
        if indices is not None:
            d = d[..., indices]
        d = d.permute(0, 3, 1, 2)
        if upsample_res != -1 and R != upsample_res:
            d = bilinear_upsample(d, upsample_res)
        outs.append(d)
    if per_layer_head_mean:
        return torch.stack([o.mean(dim=0) for o in outs], dim=0).mean(dim=0)
    return torch.stack(outs, dim=0).mean(dim=(0, 1))


# ======================================================================================
# rows 5, 6, 15: readout  (eval.py:39-155)
# ======================================================================================
def argmax_pixel(maps: Tensor) -> Tensor:
    """`find_max_pixel`, `eval.py:39-60`: first flat argmax -> (row+0.5, col+0.5), float (N,2)."""
    n, h, w = maps.shape
    flat = torch.argmax(maps.reshape(n, -1), dim=-1)
    return torch.stack([flat // w, flat % w], dim=-1).to(torch.float32) + 0.5


def mask_disc(maps: Tensor, centers: Tensor, radius: float) -> Tensor:
    """`mask_radius`, `eval.py:83-111`: keep pixels with (col-cc)^2+(row-cr)^2 > radius^2, where
    col/row are INTEGER pixel indices and centers carry the +0.5 offset."""
    n, h, w = maps.shape
    cols = torch.arange(w, dtype=torch.float32)[None, None, :]
    rows = torch.arange(h, dtype=torch.float32)[None, :, None]
    d2 = (cols - centers[:, 1, None, None]) ** 2 + (rows - centers[:, 0, None, None]) ** 2
    keep = d2 > torch.tensor(radius ** 2, dtype=torch.float32)
    return maps * keep.to(maps.dtype)


def k_max_pixels(maps: Tensor, num: int) -> Tensor:
    """`find_k_max_pixels`, `eval.py:62-81` -> (num, N, 2); disc radius 0.05*h between picks."""
    pts = []
    cur = maps
    for _ in range(num):
        p = argmax_pixel(cur)
        pts.append(p)
        cur = mask_disc(cur, p, 0.05 * maps.shape[1])
    return torch.stack(pts)


def weighted_avg_pixel(maps: Tensor, distance: float = 5.0) -> Tensor:
    """`pixel_from_weighted_avg`, `eval.py:113-155` (windowed soft-argmax).  Does not mutate."""
    n, m, w = maps.shape
    hm = maps.clone()
    rows = torch.arange(m, dtype=torch.float32)[None, :, None]
    cols = torch.arange(w, dtype=torch.float32)[None, None, :]
    if distance != -1:
        mx = argmax_pixel(hm)
        r0 = mx[:, 0].long().to(torch.float32)[:, None, None]
        c0 = mx[:, 1].long().to(torch.float32)[:, None, None]
        far = torch.sqrt((rows - r0) ** 2 + (cols - c0) ** 2) > distance
        hm = torch.where(far, torch.zeros_like(hm), hm)
    norm = hm / (hm.sum(dim=(1, 2), keepdim=True) + 1e-6)
    return torch.stack([(rows * norm).sum(dim=(1, 2)), (cols * norm).sum(dim=(1, 2))], dim=-1) + 0.5


# ======================================================================================
# row 9: gaussian targets  (optimize_token.py:204-242)
# ======================================================================================
def gaussian_targets(pos: Tensor, size: int, sigma: float) -> Tensor:
    """pos (S, N, 2) in [0,1] as (row, col) -> (N, size, size): mean over the S subjects of
    exp(-((col+0.5-pc*size)^2 + (row+0.5-pr*size)^2) / (2 sigma^2))."""
    centre = pos.to(torch.float32) * size
    rows = torch.arange(size, dtype=torch.float32)[None, None, :, None] + 0.5
    cols = torch.arange(size, dtype=torch.float32)[None, None, None, :] + 0.5
    d2 = (cols - centre[..., 1, None, None]) ** 2 + (rows - centre[..., 0, None, None]) ** 2
    g = torch.exp(-1 * d2 / (2.0 * sigma ** 2.0))
    return g.mean(dim=0)


# ======================================================================================
# rows 7, 7b, 8: token selection  (ptp_utils.py:86-187)
# ======================================================================================
def gaussian_kl(maps: Tensor, sigma: float = 3.0, epsilon: float = 1e-5, num_subjects: int = 1,
                dtype=torch.float32) -> Tensor:
    """KL(target || softmax_pixels(map)) per plane, `ptp_utils.py:97-108`."""
    n, h, w = maps.shape
    pos = k_max_pixels(maps, num_subjects) / h
    p = torch.softmax((maps.reshape(n, h * w) + epsilon).to(dtype), dim=-1)
    tgt = gaussian_targets(pos, h, sigma).reshape(n, h * w).to(dtype) + epsilon
    tgt = tgt / tgt.sum(dim=-1, keepdim=True)
    return torch.sum(tgt * (torch.log(tgt) - torch.log(p)), dim=-1)


def gaussian_kl_rank(maps: Tensor, top_k: int, sigma: float = 3.0, epsilon: float = 1e-5,
                     num_subjects: int = 1) -> Tensor:
    """`find_top_k_gaussian`, `ptp_utils.py:86-112`: ascending KL, first top_k token ids."""
    return torch.argsort(gaussian_kl(maps, sigma, epsilon, num_subjects), dim=-1, descending=False)[:top_k]


def pixel_entropy(maps: Tensor, dtype=torch.float32) -> Tensor:
    """Entropy of softmax over pixels, `ptp_utils.py:179-182` (`Categorical(probs).entropy()`
    clamps log-probabilities at the dtype's finfo.min before multiplying)."""
    n, h, w = maps.shape
    p = torch.softmax(maps.reshape(n, h * w).to(dtype), dim=-1)
    p = p / p.sum(dim=-1, keepdim=True)
    logp = torch.clamp(torch.log(p), min=torch.finfo(dtype).min)
    return -(p * logp).sum(dim=-1)


def entropy_rank(maps: Tensor, top_k: int) -> Tensor:
    """`entropy_sort`, `ptp_utils.py:165-187`."""
    return torch.argsort(pixel_entropy(maps), dim=-1, descending=False)[:top_k]


def furthest_points(maps: Tensor, top_k: int, candidates: Tensor) -> Tensor:
    """`furthest_point_sampling`, `ptp_utils.py:115-159`, restated on plain Python floats
    rounded through fp32 exactly as the tensor ops do: locations (k+0.5)/image_h, distance
    sqrt(dr^2+dc^2), strict '>' so the first candidate (list order) wins ties."""
    import numpy as np
    n, h, w = maps.shape
    loc = (argmax_pixel(maps) / h).numpy().astype(np.float32)
    cand = [int(c) for c in candidates]

    def dist(a: int, b: int) -> np.float32:
        d = loc[a] - loc[b]
        return np.sqrt(np.float32(d[0] * d[0]) + np.float32(d[1] * d[1]), dtype=np.float32)

    best, pair = np.float32(-1.0), None
    for i in range(len(cand)):
        for j in range(i + 1, len(cand)):
            dd = dist(cand[i], cand[j])
            if dd > best:
                best, pair = dd, (cand[i], cand[j])
    if pair is None:
        raise NameError("furthest_pair")   # the reference fails the same way with < 2 candidates
    chosen = [pair[0], pair[1]]
    for _ in range(top_k - 2):
        far, pick = np.float32(-1.0), None
        for c in cand:
            if c in chosen:
                continue
            m = min(dist(c, s) for s in chosen)
            if m > far:
                far, pick = m, c
        if pick is not None:
            chosen.append(pick)
    return torch.tensor(chosen, dtype=torch.int64)


# ======================================================================================
# rows 10, 11: losses  (optimize.py:157-206)
# ======================================================================================
def sharpening_loss(sel_maps: Tensor, sigma: float = 1.0, num_subjects: int = 1) -> Tensor:
    """`sharpening_loss` -> `find_gaussian_loss_at_point`, `optimize.py:166-206`:
    MSE(map, gaussian at the map's own argmax); the target carries no gradient."""
    with torch.no_grad():
        pos = k_max_pixels(sel_maps.detach().float(), num_subjects) / sel_maps.shape[-1]
        tgt = gaussian_targets(pos, sel_maps.shape[1], sigma).to(sel_maps.dtype)
    return torch.mean((sel_maps - tgt) ** 2)


def equivariance_loss(sel_maps: Tensor, sel_maps_t: Tensor, theta: Tensor) -> Tensor:
    """`equivariance_loss`, `optimize.py:157-163`, for one image: un-warp the maps of the
    transformed image with the inverse affine and take the MSE against the original maps.
    sel_maps, sel_maps_t (K,H,W); theta (2,3) the forward transform applied to the image."""
    inv = invert_theta(theta.to(torch.float32)[None])
    back = affine_warp(sel_maps_t[None], inv)[0]
    return torch.mean((sel_maps - back) ** 2)


# ======================================================================================
# row 14: test-time-augmentation ensemble  (eval.py:239-330)
# ======================================================================================
def tta_ensemble(maps_per_aug: Sequence[Tensor], thetas: Sequence[Tensor]) -> Tensor:
    """sum_a inv_a(maps_a) / sum_a inv_a(ones), NaN -> 0.  maps (K,H,W); theta (2,3)."""
    num = torch.zeros_like(maps_per_aug[0])
    tot = torch.zeros_like(maps_per_aug[0])
    for m, th in zip(maps_per_aug, thetas):
        inv = invert_theta(th.to(torch.float32)[None])
        num = num + affine_warp(torch.ones_like(m)[None], inv)[0]
        tot = tot + affine_warp(m[None], inv)[0]
    out = tot / num
    out[out != out] = 0
    return out


# ======================================================================================
# row 0: DDIM forward noising constants  (optimize_token.py:25-35, ptp_utils.py:219-223)
# ======================================================================================
def ddim_alphas_cumprod(beta_start=0.00085, beta_end=0.012, n=1000) -> Tensor:
    betas = torch.linspace(beta_start ** 0.5, beta_end ** 0.5, n, dtype=torch.float32) ** 2
    return torch.cumprod(1.0 - betas, dim=0)


# ======================================================================================
# composite: one image's maps from per-layer (q_low, k)
# ======================================================================================
def attention_maps_from_qk(q_lows: Sequence[Tensor], ks: Sequence[Tensor], heads: Sequence[int],
                           scales: Sequence[float], feature_res: int, upsample_res: int = -1,
                           indices: Optional[Tensor] = None) -> Tensor:
    """One image (B=1 slices): per captured layer softmax(q'k^T scale), then `collect_maps`."""
    stored = [layer_probs_from_qk(q[None] if q.dim() == 2 else q, k[None] if k.dim() == 2 else k, h,
                                  feature_res, sc)
              for q, k, h, sc in zip(q_lows, ks, heads, scales)]
    return collect_maps(stored, layers=range(len(stored)), upsample_res=upsample_res, indices=indices)


def attention_maps_at_pixels(q_lows: Sequence[Tensor], ks: Sequence[Tensor], heads: Sequence[int],
                             scales: Sequence[float], feature_res: int, pixels: Tensor, dtype=torch.float64,
                             return_lse2: bool = False):
    """`attention_maps_from_qk` restricted to a subset of the R x R feature pixels, so that BASELINE-size problems
    (T = 500, R = 128, 70 heads) can be pinned in fp64 in seconds: the maps of a pixel depend on no other pixel.

    q_lows[l] (s_l^2, C_l), ks[l] (T, C_l) for ONE image; `pixels` (n,) flat indices into R*R.
    Same arithmetic as `layer_probs_from_qk` + `collect_maps` (`ptp_utils.py:513-538`, `optimize.py:44-75`): the
    bicubically upsampled q' of pixel (y, x) is sum_{i,j} My[y,i] Mx[x,j] q[i,j,:]; logits q'.k * scale per head;
    softmax over tokens; mean over heads, then over layers.  -> (T, n); differentiable.
    return_lse2: also the per-(head, pixel) log2-sum-exp2 of the logits in log2 units (total heads, n), which is what
    the forward kernel's row statistics encode (m + log2(l))."""
    R = feature_res
    py, px = pixels // R, pixels % R
    out = 0.0
    lse = []
    for q, k, h, sc in zip(q_lows, ks, heads, scales):
        s = int(round(q.shape[0] ** 0.5))
        C = q.shape[1]
        M = bicubic_matrix(s, R, dtype)
        W = (M[py][:, :, None] * M[px][:, None, :]).reshape(len(pixels), s * s)        # (n, s^2)
        qp = (W @ q.to(dtype)).reshape(len(pixels), h, C // h).permute(1, 0, 2)        # (h, n, d)
        kh = k.to(dtype).reshape(-1, h, C // h).permute(1, 0, 2)                       # (h, T, d)
        sim = torch.einsum("hnd,htd->hnt", qp, kh) * sc
        out = out + torch.softmax(sim, dim=-1).mean(dim=0).t() / len(q_lows)           # (T, n)
        if return_lse2:
            lse.append(torch.logsumexp(sim, dim=-1) * 1.4426950408889634)
    return (out, torch.cat(lse, dim=0)) if return_lse2 else out


# ======================================================================================
# rows 16 / f2: regressor application and the five dataset metrics of `evaluate` (eval.py:468-510)
# ======================================================================================
H36M_SWAPS = ((1, 6), (2, 7), (3, 8), (4, 9), (5, 10), (17, 25), (18, 26), (19, 27), (20, 28), (21, 28), (22, 30), (23, 31))


def swap_points_h36m(points: Tensor) -> Tensor:
    """`swap_points`, `eval.py:360-371`: left/right relabelling of Human3.6M keypoints, points (B, N, D).  The table —
    including its (21, 28) entry, which overwrites the (20, 28) one — is the reference's."""
    perm = list(range(points.shape[1]))
    for a, b in H36M_SWAPS:
        perm[a] = b
        perm[b] = a
    return points[:, perm, :]


def evaluate_image(keypoints: Tensor, regressor: Tensor, gt_kpts: Tensor, evaluation_method: str = "inter_eye_distance",
                   visible: Optional[Tensor] = None) -> Tensor:
    """The per-image tail of `evaluate`, `eval.py:468-510`, restated line by line on CPU tensors.

    keypoints (K, 2) in [0, 1] = readout / 512 (`:463-466`); regressor (2K, 2n); gt_kpts (n, 2) in [0, 1].
    `:469`: estimated = ((keypoints.view(1,-1) - 0.5) @ regressor) + 0.5 -> (n, 2); `:475-477`: both scaled by 256 for
    `mean_average_error` / `pck`; `:480`: per-keypoint L2; then per method (`:482-509`)."""
    est = ((keypoints.reshape(1, -1) - 0.5) @ regressor + 0.5).reshape(-1, 2)
    gt = gt_kpts.clone()
    if evaluation_method in ("mean_average_error", "pck"):
        est = est * 256
        gt = gt * 256
    l2 = (est - gt).norm(dim=-1)
    if evaluation_method == "inter_eye_distance":
        eye = torch.sqrt(torch.sum((gt[0] - gt[1]) ** 2, dim=-1))
        return torch.mean(l2 / eye)
    if evaluation_method in ("visible", "mean_average_error"):
        vis = torch.ones_like(l2) if visible is None else visible.to(l2.dtype)
        tot = (l2 * vis).sum()
        return tot / vis.sum() if evaluation_method == "visible" else tot
    if evaluation_method == "pck":
        return (l2 < 6).to(torch.float32).mean()
    if evaluation_method == "orientation_invariant":
        m = l2.mean()
        sw = (swap_points_h36m(est[None])[0] - gt).norm(dim=-1).mean()
        return (sw if sw < m else m) * 128
    raise NotImplementedError(evaluation_method)
