"""Shared helpers for golden fixtures (TEST INFRASTRUCTURE): seeded inputs, subsampling, error metrics."""
from __future__ import annotations

from typing import Dict

import numpy as np
import torch

GOLDEN_CASES = {
    # name: (config name or builder args, batch, out_gain, seed)
    "tiny_base": dict(cfg="tiny_base", batch=2, out_gain=8.0),
    "tiny_refine": dict(cfg="tiny_refine", batch=2, out_gain=8.0),
    "tiny_alt_qknorm": dict(cfg="tiny_alt_qknorm", batch=2, out_gain=8.0),
    "ufm_base_assumed": dict(cfg="ufm_base_assumed", batch=1, out_gain=8.0),
    "ufm_refine_assumed": dict(cfg="ufm_refine_assumed", batch=1, out_gain=8.0),
    # uncertainty head with Covariance2DAdaptor + ConfidenceAdaptor + MaskAdaptor; gain 1 keeps exp() sane
    "tiny_cov": dict(cfg="tiny_cov", batch=2, out_gain=1.0),
    # per-view cross-attention info sharing, Refine model, qk-norm on
    "tiny_cross": dict(cfg="tiny_cross", batch=2, out_gain=8.0),
    # cross-attention info sharing with norm_cross_tokens=False (norm_y = Identity)
    "tiny_cross_raw": dict(cfg="tiny_cross_raw", batch=2, out_gain=8.0),
    # DPT fusion blocks with (eval-mode) BatchNorm
    "tiny_bn": dict(cfg="tiny_bn", batch=2, out_gain=8.0),
    # RoPE2D on the global-attention info sharing (custom_positional_encoding="rope"), with qk-norm
    "tiny_rope": dict(cfg="tiny_rope", batch=2, out_gain=8.0),
    # FlowAdaptor.output_normalized_coordinate on the Refine model (the head predicts normalised landing positions);
    # a small head gain keeps the predicted positions inside the image
    "tiny_coord": dict(cfg="tiny_coord", batch=2, out_gain=0.05),
    # norm_intermediate=False on encoder and info sharing, encoder data_norm_type "dust3r" (Refine model)
    "tiny_nonorm": dict(cfg="tiny_nonorm", batch=2, out_gain=8.0),
    # random view index for the non-reference view's positional row: torch.manual_seed(rng_seed) right before the call
    "tiny_randidx": dict(cfg="tiny_randidx", batch=2, out_gain=8.0, rng_seed=1234),
    # DPT heads hooked onto permuted layers (hooks != [0,1,2,3])
    "tiny_hooks": dict(cfg="tiny_hooks", batch=2, out_gain=8.0),
    # DINOv2 with 4 register tokens (with_registers=True), Refine model
    "tiny_reg": dict(cfg="tiny_reg", batch=2, out_gain=8.0),
    # UNet fine-feature branch of UFM-Refine, both combine methods
    "tiny_unet": dict(cfg="tiny_unet", batch=2, out_gain=8.0),
    "tiny_unet_mod": dict(cfg="tiny_unet_mod", batch=1, out_gain=8.0),
}


def case_config(name: str):
    from ufm_b200.configs import get_config, tiny

    c = GOLDEN_CASES[name]["cfg"]
    if c == "tiny_alt_qknorm":
        return tiny(True, info_sharing_str="alternating_attention", qk_norm=True)
    return get_config(c)


def make_inputs(batch: int, H: int, W: int, seed: int = 1):
    """Synthetic uint8 pair batch, BHWC: iid uniform bytes."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    src = torch.randint(0, 256, (batch, H, W, 3), dtype=torch.uint8, generator=g)
    tgt = torch.randint(0, 256, (batch, H, W, 3), dtype=torch.uint8, generator=g)
    return src, tgt


# API-shell cases with foreign and DIFFERENT source / target sizes (base.py:253-334: AutomaticShapeSelection resize in,
# un-map out).  The first four are the shapes of the reference's own example pairs (examples/image_pairs: bike 1080x1080 /
# 607x1080, building 810x1080, cook 607x1080 / 1080x1080, fire_academy 580x1080) at one fifth of the size; the last one
# mixes aspect ratios on both sides.  (h0, w0, h1, w1); model = tiny_base (168 x 224 trained resolution), head gain 8.
SHAPE_CASES = [(216, 216, 121, 216), (162, 216, 162, 216), (121, 216, 216, 216), (116, 216, 116, 216), (162, 216, 100, 300)]
SHAPE_GOLDEN = "tiny_shapes"


def make_shape_inputs(i: int):
    """Seeded uint8 BHWC pair (batch 1) of SHAPE_CASES[i]."""
    h0, w0, h1, w1 = SHAPE_CASES[i]
    g = torch.Generator(device="cpu").manual_seed(100 + i)
    src = torch.randint(0, 256, (1, h0, w0, 3), dtype=torch.uint8, generator=g)
    tgt = torch.randint(0, 256, (1, h1, w1, 3), dtype=torch.uint8, generator=g)
    return src, tgt


def subsample(t: torch.Tensor, max_elems: int = 40000) -> torch.Tensor:
    """Deterministic strided subsample of the two trailing spatial dims so a fixture stays small."""
    t = t.detach().float()
    s = 1
    while t[..., ::s, ::s].numel() > max_elems and t.dim() >= 2:
        s += 1
    return t[..., ::s, ::s].contiguous()


def stats(t: torch.Tensor) -> np.ndarray:
    t = t.detach().double()
    return np.array([t.mean().item(), t.abs().mean().item(), t.norm().item(), t.abs().max().item()])


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


def max_abs(a: torch.Tensor, b: torch.Tensor) -> float:
    return (a.double() - b.double()).abs().max().item()


def pack_stages(stages: Dict[str, torch.Tensor]) -> Dict[str, np.ndarray]:
    out = {}
    for k, v in stages.items():
        out["sub:" + k] = subsample(v).numpy()
        out["stat:" + k] = stats(v)
    return out
