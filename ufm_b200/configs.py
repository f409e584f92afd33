"""Model configurations for the UFM dense-correspondence forward pass.

The production ``config.json`` of UFM-Base / UFM-Refine lives on the HuggingFace hub
(reference ``uniflowmatch/models/ufm.py:1254-1256``) and is not reachable offline, so the
"assumed" configs below pin the architecture that SURVEY.md Appendix A derives from the
constraints in the reference code.  Every dict here is a valid ``**kwargs`` for the
reference constructors ``UniFlowMatchConfidence`` (``ufm.py:483-509``) and
``UniFlowMatchClassificationRefinement`` (``ufm.py:720-754``) and for the engine classes of
the same name in :mod:`ufm_b200.models`.
"""
from __future__ import annotations

import copy
from typing import Any, Dict

# DINOv2 hub architectures (third-party ``facebookresearch/dinov2`` ``hub/backbones.py``; see
# SURVEY.md Appendix B).  (embed_dim, depth, heads)
DINOV2_ARCH = {
    "small": (384, 12, 6),
    "base": (768, 12, 12),
    "large": (1024, 24, 16),
}


def _dpt(input_dims, out_dim, layer_dims=(96, 192, 384, 768), feature_dim=256):
    return {
        "dpt_feature": {
            "patch_size": 14,
            "hooks": [0, 1, 2, 3],
            "input_feature_dims": list(input_dims),
            "layer_dims": list(layer_dims),
            "feature_dim": feature_dim,
            "use_bn": False,
        },
        "dpt_processor": {"input_feature_dim": feature_dim, "output_dim": out_dim},
    }


def ufm_base_assumed(inference_resolution=(560, 420)) -> Dict[str, Any]:
    """SURVEY.md Appendix A "UFM-Base-arch": DINOv2 ViT-L/14 + 12x768 global attention + 2 DPT heads."""
    W, H = inference_resolution
    return {
        "encoder_str": "dinov2",
        "encoder_kwargs": {
            "name": "dinov2_large", "size": "large", "data_norm_type": "dinov2", "patch_size": 14,
            "with_registers": False, "indices": [11, 23], "norm_intermediate": True,
            "keep_first_n_layers": None,
        },
        "info_sharing_str": "global_attention",
        "info_sharing_kwargs": {
            "name": "gat", "input_embed_dim": 1024, "max_num_views": 2,
            "use_rand_idx_pe_for_non_reference_views": False, "depth": 12, "dim": 768, "num_heads": 12,
            "mlp_ratio": 4.0, "qkv_bias": True, "qk_norm": False, "init_values": None,
            "custom_positional_encoding": None, "indices": [5, 8], "norm_intermediate": True,
        },
        "head_type": "dpt",
        "feature_head_kwargs": _dpt([1024, 768, 768, 768], 2),
        "adaptors_kwargs": {
            "flow": {"class": "FlowAdaptor", "kwargs": {
                "name": "flow", "flow_mean": [0.0, 0.0], "flow_std": [1.0, 1.0],
                "base_shape": [H, W], "scale_strategy": "none"}},
        },
        "uncertainty_head_type": "dpt",
        "uncertainty_head_kwargs": _dpt([1024, 768, 768, 768], 1),
        "uncertainty_adaptors_kwargs": {
            "non_occluded_mask": {"class": "MaskAdaptor", "kwargs": {"name": "non_occluded_mask"}},
        },
        "inference_resolution": [W, H],
    }


def ufm_refine_assumed(inference_resolution=(560, 420)) -> Dict[str, Any]:
    """SURVEY.md Appendix A Refine additions (``ufm.py:736-746`` defaults)."""
    cfg = ufm_base_assumed(inference_resolution)
    cfg.update({
        "classification_head_type": "patch_mlp",
        "classification_head_kwargs": {"input_feature_dim": 1792, "patch_size": 14, "output_dim": 16,
                                       "mlp_ratio": 4},
        "temperature": 4,
        "refinement_range": 5,
        "use_unet_feature": False,
    })
    return cfg


def tiny(refine: bool = False, inference_resolution=(224, 168), info_sharing_str="global_attention",
         qk_norm: bool = False, unc_adaptors: str = "mask", conf_type: str = "sigmoid",
         unet: str = "", rope: bool = False) -> Dict[str, Any]:
    """A small architecture with the same topology, for fast CPU goldens and GPU parity at every
    stage boundary: ViT-S/14 truncated to 4 layers, 4x128 info sharing, slim DPT heads."""
    W, H = inference_resolution
    cfg = ufm_refine_assumed(inference_resolution) if refine else ufm_base_assumed(inference_resolution)
    cfg = copy.deepcopy(cfg)
    cfg["encoder_kwargs"].update({"name": "dinov2_small", "size": "small", "indices": [1, 3],
                                  "keep_first_n_layers": 4})
    cfg["info_sharing_str"] = info_sharing_str
    cfg["info_sharing_kwargs"].update({"input_embed_dim": 384, "depth": 4, "dim": 128, "num_heads": 2,
                                       "indices": [1, 2], "qk_norm": qk_norm})
    if rope:
        cfg["info_sharing_kwargs"]["custom_positional_encoding"] = "rope"  # global_attention_transformer.py:110-113
    if info_sharing_str == "cross_attention":
        # MultiViewCrossAttentionTransformerIFR kwargs (cross_attention_transformer.py:264-292); SURVEY.md 8f N2
        ik = cfg["info_sharing_kwargs"]
        for k in ("max_num_views", "use_rand_idx_pe_for_non_reference_views"):
            ik.pop(k)
        ik.update({"num_views": 2, "norm_cross_tokens": True})
    if info_sharing_str == "alternating_attention":
        cfg["info_sharing_kwargs"].pop("max_num_views")
        cfg["info_sharing_kwargs"].update({"use_pe_for_non_reference_views": False, "max_num_views_for_pe": 1000})
    cfg["feature_head_kwargs"] = _dpt([384, 128, 128, 128], 2, layer_dims=(64, 64, 128, 128), feature_dim=64)
    cfg["uncertainty_head_kwargs"] = _dpt([384, 128, 128, 128], 1, layer_dims=(64, 64, 128, 128), feature_dim=64)
    if unc_adaptors == "cov_conf_mask":
        # flow covariance + keypoint confidence + covisibility mask from one 5-channel uncertainty head
        # (names as consumed by ufm.py:644-660)
        vmin, vmax = (1.0, 50.0) if conf_type == "exp" else (0.0, 1.0)
        cfg["uncertainty_head_kwargs"] = _dpt([384, 128, 128, 128], 5, layer_dims=(64, 64, 128, 128), feature_dim=64)
        cfg["uncertainty_adaptors_kwargs"] = {
            "flow_cov": {"class": "Covariance2DAdaptor", "kwargs": {"name": "flow_cov", "parametrization": "exp_tanh"}},
            "keypoint_confidence": {"class": "ConfidenceAdaptor", "kwargs": {
                "name": "keypoint_confidence", "confidence_type": conf_type, "vmin": vmin, "vmax": vmax}},
            "non_occluded_mask": {"class": "MaskAdaptor", "kwargs": {"name": "non_occluded_mask"}},
        }
    if refine:
        cfg["classification_head_kwargs"].update({"input_feature_dim": 384 + 128})
        if unet:
            # UNet fine-feature branch (ufm.py:817-825, 967-982); the UNet itself is not configurable
            cfg.update({"use_unet_feature": True, "feature_combine_method": unet})
    return cfg


def _raw_cross(cfg):
    """Cross-attention info sharing without the norm on the cross tokens (norm_cross_tokens=False)."""
    cfg["info_sharing_kwargs"]["norm_cross_tokens"] = False
    return cfg


def _use_bn(cfg):
    """DPT fusion blocks with BatchNorm (use_bn=True, dpt_block.py:134-149); inference = eval-mode statistics."""
    for k in ("feature_head_kwargs", "uncertainty_head_kwargs"):
        cfg[k]["dpt_feature"]["use_bn"] = True
    return cfg


def _nonorm(cfg):
    """norm_intermediate=False on the encoder (dinov2.py:259-261: every returned index stays un-normed) and on the info
    sharing (intermediates only, global_attention_transformer.py:549), with a non-default encoder data_norm_type."""
    cfg["encoder_kwargs"].update({"norm_intermediate": False, "data_norm_type": "dust3r"})
    cfg["info_sharing_kwargs"]["norm_intermediate"] = False
    return cfg


def _randidx(cfg):
    """use_rand_idx_pe_for_non_reference_views (global_attention_transformer.py:520-524): the second view's row of the
    view_pos_table is drawn per forward with torch.randint(1, max_num_views)."""
    cfg["info_sharing_kwargs"].update({"max_num_views": 6, "use_rand_idx_pe_for_non_reference_views": True})
    return cfg


def _hooks(cfg):
    """DPT heads hooked onto permuted layers (dpt.py:47,198-219): the flow head reads [inter0, enc_last, final, inter1], the
    uncertainty head [final, inter1, inter0, enc_last]; input_feature_dims follow the selected features."""
    for key, hooks in (("feature_head_kwargs", [1, 0, 3, 2]), ("uncertainty_head_kwargs", [3, 2, 1, 0])):
        f = cfg[key]["dpt_feature"]
        dims = [f["input_feature_dims"][0]] + [f["input_feature_dims"][1]] * 3  # dims of [enc_last, inter0, inter1, final]
        f["hooks"] = hooks
        f["input_feature_dims"] = [dims[h] for h in hooks]
    return cfg


def _registers(cfg):
    """DINOv2 with register tokens (with_registers=True -> hub dinov2_vit*14_reg, dinov2.py:24,49,74-79)."""
    cfg["encoder_kwargs"].update({"with_registers": True, "name": "dinov2_small_reg"})
    return cfg


def _coord(cfg):
    """FlowAdaptor predicting the landing position (output_normalized_coordinate, adaptors.py:110-118)."""
    cfg["adaptors_kwargs"]["flow"]["kwargs"]["output_normalized_coordinate"] = True
    return cfg


CONFIGS = {
    "ufm_base_assumed": ufm_base_assumed,
    "ufm_refine_assumed": ufm_refine_assumed,
    "ufm_base_hires": lambda: ufm_base_assumed((1120, 840)),
    "tiny_base": lambda: tiny(False),
    "tiny_refine": lambda: tiny(True),
    "tiny_cov": lambda: tiny(False, unc_adaptors="cov_conf_mask"),
    "tiny_cross": lambda: tiny(True, info_sharing_str="cross_attention", qk_norm=True),
    "tiny_cross_raw": lambda: _raw_cross(tiny(False, info_sharing_str="cross_attention")),
    "tiny_rope": lambda: tiny(False, qk_norm=True, rope=True),
    "tiny_bn": lambda: _use_bn(tiny(False)),
    "tiny_coord": lambda: _coord(tiny(True)),
    "tiny_nonorm": lambda: _nonorm(tiny(True)),
    "tiny_randidx": lambda: _randidx(tiny(False)),
    "tiny_hooks": lambda: _hooks(tiny(False)),
    "tiny_reg": lambda: _registers(tiny(True)),
    "tiny_unet": lambda: tiny(True, unet="conv"),
    "tiny_unet_mod": lambda: tiny(True, unet="modulate"),
    "ufm_refine_unet": lambda: dict(ufm_refine_assumed(), use_unet_feature=True, feature_combine_method="conv"),
}


def get_config(name: str) -> Dict[str, Any]:
    return CONFIGS[name]()
