"""Import shim for the UNMODIFIED reference at /root/reference (container only).

TEST INFRASTRUCTURE.  Used by `oracle/make_golden.py` and by the `-m "not gpu"` tests
that pin the oracle against the live reference.  `/root/reference` does not exist on the
GPU box, so nothing on the product path, in `bench.py` or in the `-m gpu` tests may call
this module.

The reference imports third-party packages that are absent offline (diffusers, h5py,
matplotlib) and a namespace package `datasets/` that Hugging Face `datasets` shadows
(SURVEY.md §8(c)).  We register inert stand-ins for those names and then import the
reference's own modules unmodified; every arithmetic function of the hot path
(`ptp_utils`, `optimize`, `eval`, `optimize_token`, `invertable_transform`) is then the
reference's real code running on CPU.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("SKP_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "unsupervised_keypoints"))


def _stub(name: str, **attrs):
    mod = types.ModuleType(name)
    mod.__dict__.update(attrs)
    sys.modules[name] = mod
    return mod


_loaded = None


def load():
    """Returns a namespace with the reference modules (ptp_utils, optimize, eval, ...)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")

    class _Missing:  # any attribute access on a stubbed class raises loudly
        def __init__(self, *a, **k):
            raise RuntimeError("stubbed third-party class instantiated")

    if "diffusers" not in sys.modules:
        _stub("diffusers", StableDiffusionPipeline=_Missing, DDIMScheduler=_Missing)
    for name in ("h5py", "matplotlib", "matplotlib.colors", "matplotlib.pyplot", "wandb"):
        if name not in sys.modules:
            try:
                importlib.import_module(name)
            except Exception:
                _stub(name)
    # the reference's namespace package `datasets/` must win over Hugging Face `datasets`
    saved = {k: v for k, v in sys.modules.items() if k == "datasets" or k.startswith("datasets.")}
    for k in saved:
        del sys.modules[k]
    pkg = types.ModuleType("datasets")
    pkg.__path__ = [os.path.join(REFERENCE_ROOT, "datasets")]
    sys.modules["datasets"] = pkg
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        # same order as the reference's main.py:7 (circular imports at ptp_utils.py:22-29)
        optimize_token = importlib.import_module("unsupervised_keypoints.optimize_token")
        ptp_utils = importlib.import_module("unsupervised_keypoints.ptp_utils")
        optimize = importlib.import_module("unsupervised_keypoints.optimize")
        ref_eval = importlib.import_module("unsupervised_keypoints.eval")
        inv = importlib.import_module("unsupervised_keypoints.invertable_transform")
        kreg = importlib.import_module("unsupervised_keypoints.keypoint_regressor")
    finally:
        sys.path.remove(REFERENCE_ROOT)
    _loaded = types.SimpleNamespace(
        optimize_token=optimize_token, ptp_utils=ptp_utils, optimize=optimize,
        eval=ref_eval, invertable_transform=inv, keypoint_regressor=kreg,
    )
    return _loaded
