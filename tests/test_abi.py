"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports exactly the
symbols include/skp_b200.h declares (no compute calls: there is no GPU here)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "skp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(skp_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from stablekeypoints_b200 import _lib
    assert os.path.exists(_lib.LIB_PATH), "run ./build.sh (or __graft_entry__.build()) first"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    declared = _header_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in skp_b200.h but not exported"
    assert sorted(_lib.SIGNATURES) == declared, "ctypes signatures out of sync with the header"
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = sorted(set(re.findall(r" T (skp_[a-z0-9_]+)", out)))
    assert exported == declared, "library exports symbols the header does not declare"


def test_library_binds_and_reports_version():
    from stablekeypoints_b200 import _lib
    lib = _lib.load()
    assert lib.skp_version() >= 104
    assert isinstance(lib.skp_last_error(), bytes)


def test_workspace_query_per_call_is_host_only_and_consistent():
    """skp_attn_map_workspace_bytes[_for] are pure host arithmetic: BASELINE capture set (SD-2.1-base, T=500, R=128, 8 images)."""
    from stablekeypoints_b200 import _lib
    lib = _lib.load()
    spec = [(16, 1280, 20)] * 3 + [(32, 640, 10)]
    arr = (_lib.AttnLayer * len(spec))()
    for i, (s, c, h) in enumerate(spec):
        arr[i].s, arr[i].C, arr[i].heads, arr[i].k_shared = s, c, h, 1
        arr[i].scale, arr[i].weight = (c // h) ** -0.5, 1.0 / len(spec)
    q = lambda prec, call: lib.skp_attn_map_workspace_bytes_for(arr, len(spec), 8, 500, 128, prec, call)
    for prec in (_lib.PREC_AUTO, _lib.PREC_TC_BF16X3):
        fwd, bs, bn = q(prec, _lib.WS_FWD), q(prec, _lib.WS_BWD_STATS), q(prec, _lib.WS_BWD_NOSTATS)
        assert q(prec, _lib.WS_ANY) == max(fwd, bs, bn) == lib.skp_attn_map_workspace_bytes(arr, len(spec), 8, 500, 128, prec)
        assert 0 < fwd < 600e6 and 0 < bs < 1.2e9          # tcgen05 paths: operand copies at native resolution
        assert bn == 4 * 8 * 128 * 128 * (3 * 1280 + 640) + 4 * 8 * 128 * 16 * 1280 > 2e9   # CUDA-core backward: upsampled dq + tmp
    assert q(_lib.PREC_SIMT_F32, _lib.WS_FWD) == 0
    assert q(_lib.PREC_SIMT_F32, _lib.WS_BWD_STATS) == q(_lib.PREC_SIMT_F32, _lib.WS_BWD_NOSTATS)
    assert lib.skp_attn_map_workspace_bytes_for(None, 0, 8, 500, 128, 0, 0) == 0


def test_cpu_tensors_are_refused():
    import torch
    from stablekeypoints_b200 import ops
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.plane_argmax(torch.rand(2, 8, 8))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.attention_maps([torch.rand(1, 16, 32)], [torch.rand(1, 5, 32)], [2], [0.25], 8)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "stablekeypoints_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports oracle/"


def test_torch_operator_library_is_registered_with_meta_kernels_and_no_cpu_kernel():
    """the entry points are also `torch.ops.skp_b200.*` operators (CUDA key only; shapes from Meta kernels)."""
    import torch
    import stablekeypoints_b200.torch_ops as T
    for name in T.SCHEMAS:
        assert hasattr(torch.ops.skp_b200, name), name
    qs = [torch.empty(2, 16, 32, device="meta"), torch.empty(2, 64, 16, device="meta")]
    ks = [torch.empty(1, 40, 32, device="meta"), torch.empty(1, 40, 16, device="meta")]
    m = torch.ops.skp_b200.attention_maps(qs, ks, [2, 1], [0.25, 0.25], 32, None)
    assert m.shape == (2, 40, 32, 32) and m.dtype == torch.float32
    m = torch.ops.skp_b200.attention_maps(qs, ks, [2, 1], [0.25, 0.25], 32, torch.empty(5, dtype=torch.int64, device="meta"))
    assert m.shape == (2, 5, 32, 32)
    planes = torch.empty(7, 32, 32, device="meta")
    assert torch.ops.skp_b200.plane_argmax(planes).shape == (7,)
    kl, ent, am = torch.ops.skp_b200.plane_stats(planes, 2.0, 1e-5, 1)
    assert kl.shape == ent.shape == am.shape == (7,) and am.dtype == torch.int64
    assert torch.ops.skp_b200.find_k_max_pixels(planes, 3).shape == (3, 7, 2)
    assert torch.ops.skp_b200.bilinear_upsample(planes, 128).shape == (7, 128, 128)
    import pytest
    with pytest.raises((NotImplementedError, RuntimeError)):        # no CPU kernel: PyTorch's dispatcher says so
        torch.ops.skp_b200.plane_argmax(torch.zeros(2, 8, 8))


def test_torch_operators_trace_under_fake_tensors():
    """The Meta kernels are enough for `make_fx` to trace a graph through the registered operators on a machine without a GPU."""
    import torch
    from torch.fx.experimental.proxy_tensor import make_fx
    import stablekeypoints_b200.torch_ops  # noqa: F401

    def readout(planes):
        kl, ent, am = torch.ops.skp_b200.plane_stats(planes, 2.0, 1e-5, 1)
        return torch.ops.skp_b200.rank_ascending(kl, 4), torch.ops.skp_b200.bilinear_upsample(planes, 64)

    gm = make_fx(readout, tracing_mode="fake")(torch.empty(9, 16, 16, device="meta"))
    targets = [str(n.target) for n in gm.graph.nodes if n.op == "call_function"]
    assert any("skp_b200.plane_stats" in t for t in targets) and any("skp_b200.rank_ascending" in t for t in targets)
    assert any("skp_b200.bilinear_upsample" in t for t in targets)
