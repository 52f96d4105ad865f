"""Host mirror logic that needs no GPU: controller protocol, capture gate, registration, stand-in shapes,
the CPU pipeline port against the reference golden."""
import os

import numpy as np
import pytest
import torch

from stablekeypoints_b200 import ptp_utils
from stablekeypoints_b200.sd_standin import PipelineStandIn, UNet2DConditionStandIn, unet_config

G = os.path.join(os.path.dirname(__file__), "golden")


def _tiny_unet(name="tiny-sd21"):
    torch.manual_seed(0)
    return UNet2DConditionStandIn(unet_config(name)), unet_config(name)


def _run(unet, cfg, controller, B=1, T=7):
    x = torch.randn(B, 4, cfg.latent_side, cfg.latent_side)
    ctx = torch.randn(1, T, cfg.context_dim)
    try:
        unet(x, torch.zeros(B, dtype=torch.int64), ctx.expand(B, -1, -1))
    except ptp_utils.StopForward:
        return True
    return False


def test_registration_counts_up_block_layers_and_gate_matches_reference():
    unet, cfg = _tiny_unet()
    ctl = ptp_utils.AttentionStore()
    ptp_utils.register_attention_control(unet, ctl, feature_upsample_res=32)
    assert ctl.num_att_layers == 18          # 3 levels x 3 transformer blocks x (attn1 + attn2), ptp_utils.py:555-570
    with torch.no_grad():
        stopped = _run(unet, cfg, ctl, B=2)
    assert not stopped
    store = ctl.step_store["attn"]
    # reference gate (ptp_utils.py:508-512): cross-attention only, <= 32^2 queries, first 4 -> 3 @ 4x4 then 1 @ 8x8
    assert [h.side for h in store] == [4, 4, 4, 8]
    assert [h.heads for h in store] == [8, 8, 8, 4]
    h0 = store[0]
    # the batch shares one context (context.expand): k is projected once; the handle still reports the reference's shape
    assert h0.q.shape == (2, 16, 128) and h0.k.shape == (1, 7, 128) and h0.k_shared and h0.shape == (2 * 8, 32 * 32, 7)
    ctl.reset()
    # a genuinely per-sample context through the same controller is never collapsed to sample 0's
    with torch.no_grad():
        unet(torch.randn(2, 4, cfg.latent_side, cfg.latent_side), torch.zeros(2, dtype=torch.int64), torch.randn(2, 7, cfg.context_dim))
    h1 = ctl.step_store["attn"][0]
    assert h1.k.shape == (2, 7, 128) and not h1.k_shared
    ctl.reset()
    assert ctl.step_store == {"attn": []}


def test_gate_is_parameterised_for_the_64x64_extension_and_early_stop():
    unet, cfg = _tiny_unet()
    ctl = ptp_utils.AttentionStore(max_query_tokens=16 ** 2, max_captured_layers=9)
    ptp_utils.register_attention_control(unet, ctl, feature_upsample_res=32)
    with torch.no_grad():
        _run(unet, cfg, ctl)
    assert [h.side for h in ctl.step_store["attn"]] == [4, 4, 4, 8, 8, 8, 16, 16, 16]
    ctl2 = ptp_utils.AttentionStore()
    ctl2.stop_after_capture = True
    ptp_utils.register_attention_control(unet, ctl2, feature_upsample_res=32)
    with torch.no_grad():
        assert _run(unet, cfg, ctl2, B=3)              # raised StopForward after the 4th captured layer
    assert len(ctl2.step_store["attn"]) == 4
    assert ctl2.step_store["attn"][0].k.shape[0] == 1 and ctl2.step_store["attn"][0].k_shared   # one k for the batch


def test_no_cross_attention_raises_like_the_reference():
    with pytest.raises(AssertionError):
        ptp_utils.register_attention_control(torch.nn.Sequential(torch.nn.Linear(2, 2)), ptp_utils.AttentionStore())


def test_patched_forward_equals_module_forward():
    """The trunk's own attention output is unchanged by the hook (SDPA vs einsum/softmax/matmul)."""
    unet, cfg = _tiny_unet("tiny-sd15")
    x = torch.randn(2, 4, cfg.latent_side, cfg.latent_side)
    ctx = torch.randn(2, 5, cfg.context_dim)
    t = torch.zeros(2, dtype=torch.int64)
    with torch.no_grad():
        want = unet(x, t, ctx)["sample"]
        ptp_utils.register_attention_control(unet, ptp_utils.AttentionStore(), feature_upsample_res=32)
        got = unet(x, t, ctx)["sample"]
    assert torch.allclose(got, want, rtol=1e-4, atol=1e-5)


def test_cpu_pipeline_port_matches_reference_golden():
    """oracle/cpu_pipeline.py (bench's CPU baseline) reproduces the live reference's iteration exactly."""
    from oracle import cpu_pipeline as P
    d = np.load(os.path.join(G, "e2e_tiny.npz"))
    ldm = PipelineStandIn(unet_config("tiny-sd15"), vae_channels=(32, 32, 64, 64), seed=int(d["cfg_seed"]))
    for p in list(ldm.unet.parameters()) + list(ldm.vae.parameters()):
        p.requires_grad = False
    st = P.CpuStore()
    P.register_cpu_hook(ldm.unet, st, int(d["cfg_R"]))
    ctx = torch.from_numpy(d["context"]).requires_grad_(True)
    sh, eq, idx = P.cpu_token_opt_iteration(
        ldm, st, torch.from_numpy(d["image"]), torch.from_numpy(d["theta"])[0], ctx, torch.from_numpy(d["noise_a"]),
        torch.from_numpy(d["noise_b"]), n_cand=int(d["cfg_n_cand"]), top_k=int(d["cfg_top_k"]), sigma=float(d["cfg_sigma"]))
    assert idx.tolist() == d["ref_idx"].tolist()
    assert abs(sh.item() - float(d["ref_sharp"])) < 1e-6 and abs(eq.item() - float(d["ref_equiv"])) < 1e-7
    want = torch.from_numpy(d["ref_context_grad"])
    assert (ctx.grad - want).abs().max().item() <= 1e-4 * want.abs().max().item()


def test_optimize_embedding_names_a_dataset_smaller_than_one_unet_pass():
    """A shard with fewer images than one UNet pass used to surface as a bare StopIteration from the DataLoader (bench.py
    --workload sdxl at 8 ranks); the argument check happens before anything touches the GPU."""
    from stablekeypoints_b200 import optimize
    from stablekeypoints_b200.datasets import SyntheticKeypoints
    ds = SyntheticKeypoints(length=1, image_size=32)
    with pytest.raises(ValueError, match="fewer than the 2 of one UNet pass"):
        optimize.optimize_embedding(None, wandb_log=False, context=torch.zeros(1, 4, 8), device="cpu", num_steps=1, batch_size=2,
                                    num_gpus=2, dataset=ds)
    with pytest.raises(ValueError, match="multiple of num_gpus"):
        optimize.optimize_embedding(None, wandb_log=False, context=torch.zeros(1, 4, 8), device="cpu", num_steps=1, batch_size=3,
                                    num_gpus=2, dataset=ds)
