"""SDXL variant of the controller API (`unsupervised_keypoints/sdxl_monkey_patch.py:8-214`).

In the reference this module is dead code: its registration looks for `AttnProcessor2_0` objects
among `nn.Module.children()` (processors are not modules) and calls a 5-argument closure with
two arguments (`sdxl_monkey_patch.py:164-167`), so nothing is ever patched (SURVEY.md §2 #8,
§8(a) row 18).  What a user of the reference can rely on is the API SHAPE, kept here:

* `AttentionControl.__call__(attn, is_cross, place_in_unet)` — hands the second half of the batch
  (the conditional half under classifier-free guidance) to `forward`, counts layers and steps,
  calls `between_steps()` after the last layer of a step (`:25-35`);
* `AttentionStore` — `step_store` keyed `{down,mid,up}_{cross,self}`, keeps only entries with at
  most 32**2 query positions (`:55-60`), `between_steps` accumulates into `attention_store`,
  `get_average_attention()` divides by `cur_step` (`:62-75`);
* `register_attention_control(model, controller)` with `model.unet` the UNet (`:89, 203-214`).

Behaviour of the registration is DEFINED here (parity unpinned — there is nothing to pin against)
as the algorithm of the SD path (`ptp_utils.py:508-538` + `optimize.py:27-79`) on SDXL-shaped
modules: every `CrossAttention` under the `down*`/`mid*`/`up*` children is patched; cross-attention
layers of the up path within the capture gate hand the controller a `CapturedAttention` handle
(own q / k projections; the fused kernel upsamples q), stored under `up_cross`.  `step_store["attn"]`
aliases `up_cross`, so `optimize.collect_maps(_batched)` and every stage driver work unchanged.
"""
import abc

import torch
import torch.nn.functional as F

from .ptp_utils import CapturedAttention, StopForward


class AttentionControl(abc.ABC):
    def step_callback(self, x_t):
        return x_t

    def between_steps(self):
        return

    @property
    def num_uncond_att_layers(self):
        return 0

    @abc.abstractmethod
    def forward(self, attn, is_cross: bool, place_in_unet: str):
        raise NotImplementedError

    def __call__(self, attn, is_cross: bool, place_in_unet: str):
        if self.cur_att_layer >= self.num_uncond_att_layers:
            if isinstance(attn, torch.Tensor):
                half = attn.shape[0] // 2          # conditional half of a classifier-free-guidance batch
                attn[half:] = self.forward(attn[half:], is_cross, place_in_unet)
            else:                                  # a CapturedAttention handle: no guidance split to undo
                attn = self.forward(attn, is_cross, place_in_unet)
        self.cur_att_layer += 1
        if self.cur_att_layer == self.num_att_layers + self.num_uncond_att_layers:
            self.cur_att_layer = 0
            self.cur_step += 1
            self.between_steps()
        return attn

    def reset(self):
        self.cur_step = 0
        self.cur_att_layer = 0

    def __init__(self):
        self.cur_step = 0
        self.num_att_layers = -1
        self.cur_att_layer = 0


class _KeyedStore(dict):
    """`{place}_{cross|self}` -> list; "attn" is an alias of "up_cross" (what `collect_maps` reads)."""

    def __missing__(self, key):
        if key == "attn":
            return self["up_cross"]
        raise KeyError(key)


class AttentionStore(AttentionControl):
    KEYS = ("down_cross", "mid_cross", "up_cross", "down_self", "mid_self", "up_self")

    @staticmethod
    def get_empty_store():
        return _KeyedStore({k: [] for k in AttentionStore.KEYS})

    def forward(self, attn, is_cross: bool, place_in_unet: str):
        key = f"{place_in_unet}_{'cross' if is_cross else 'self'}"
        # the reference's memory gate (`:58`) is on the QUERY positions of the layer; a handle reports the upsampled
        # (B*h, R^2, T) shape of the tensor it stands for, so its native side is what the gate sees
        n_query = attn.side ** 2 if isinstance(attn, CapturedAttention) else attn.shape[1]
        if n_query <= self.max_query_tokens:
            self.step_store[key].append(attn)
        return attn

    def between_steps(self):
        """Accumulates tensors over diffusion steps (`:62-69`).  Handles are not summable: with the
        single-step use of the keypoint path (`t = timesteps[-1]`) the step store is simply kept until
        `collect_maps` / `reset()` consumes it."""
        if any(isinstance(x, CapturedAttention) for v in self.step_store.values() for x in v):
            return
        if len(self.attention_store) == 0:
            self.attention_store = self.step_store
        else:
            for key in self.attention_store:
                for i in range(len(self.attention_store[key])):
                    self.attention_store[key][i] += self.step_store[key][i]
        self.step_store = self.get_empty_store()

    def get_average_attention(self):
        return {key: [item / self.cur_step for item in self.attention_store[key]] for key in self.attention_store}

    def reset(self):
        super().reset()
        self.step_store = self.get_empty_store()
        self.attention_store = {}

    def __init__(self, max_query_tokens: int = 32 ** 2, max_captured_layers: int = 4):
        super().__init__()
        self.step_store = self.get_empty_store()
        self.attention_store = {}
        self.max_query_tokens = max_query_tokens
        self.max_captured_layers = max_captured_layers     # first N up-path cross-attention layers (`ptp_utils.py:508-512`)
        self.stop_after_capture = False


def register_attention_control(model, controller, feature_upsample_res=128):
    """Patches every `CrossAttention` of `model.unet` (or of `model` itself when it has no `.unet`).

    Returns nothing; sets `controller.num_att_layers` to the number of patched modules the
    controller will be called from per UNet pass (the captured ones)."""
    unet = getattr(model, "unet", model)
    unet = getattr(unet, "module", unet)

    def patched_forward(self, place_in_unet):
        to_out = self.to_out[0] if isinstance(self.to_out, torch.nn.ModuleList) else self.to_out

        def forward(x, context=None, mask=None):
            b, n, _ = x.shape
            h = self.heads
            is_cross = context is not None
            store = getattr(controller, "step_store", None)
            capture = (is_cross and place_in_unet == "up" and store is not None
                       and n <= getattr(controller, "max_query_tokens", 32 ** 2)
                       and len(store["up_cross"]) < getattr(controller, "max_captured_layers", 4))
            q = self.to_q(x)
            shared = is_cross and b > 1 and context.shape[0] == b and context.stride(0) == 0   # context.expand(B, ...)
            ctx = (context[:1] if shared else context) if is_cross else x
            k = self.to_k(ctx)
            if capture:
                controller(CapturedAttention(q, k, h, self.scale, int(n ** 0.5), feature_upsample_res, shared),
                           is_cross, place_in_unet)
                if getattr(controller, "stop_after_capture", False) and \
                        len(store["up_cross"]) >= getattr(controller, "max_captured_layers", 4):
                    raise StopForward()
            v = self.to_v(ctx)
            d = q.shape[-1] // h
            qh = q.reshape(b, -1, h, d).transpose(1, 2)
            kh = k.reshape(k.shape[0], -1, h, d).transpose(1, 2).expand(b, -1, -1, -1)
            vh = v.reshape(v.shape[0], -1, h, d).transpose(1, 2).expand(b, -1, -1, -1)
            am = None if mask is None else mask.reshape(b, 1, 1, -1)
            out = F.scaled_dot_product_attention(qh, kh, vh, attn_mask=am, scale=self.scale)
            return to_out(out.transpose(1, 2).reshape(b, -1, h * d))

        return forward

    def walk(net, place):
        if net.__class__.__name__ == "CrossAttention":
            net.forward = patched_forward(net, place)
            return 1
        return sum(walk(c, place) for c in net.children())

    patched = 0
    for name, child in unet.named_children():
        for place in ("down", "up", "mid"):
            if place in name:
                patched += walk(child, place)
                break
    assert patched != 0, "No CrossAttention modules found under the UNet's down/mid/up blocks."
    if controller is not None:
        controller.num_att_layers = getattr(controller, "max_captured_layers", 4)
