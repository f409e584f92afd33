"""Deterministic random-init ``state_dict`` for a UFM config, with the reference's key names.

There are no checkpoints offline, so every parity and bench run uses random-init weights of the
named architecture.  The reference initialises through module constructors and the global RNG
(``global_attention_transformer.py:180-194``, PyTorch defaults for convs), which cannot be reproduced
without importing it; instead each tensor here is drawn from its own generator seeded by
``(seed, key)`` so the same weights can be produced on any box and loaded into the reference
(``load_state_dict``, strict), the oracle and the engine alike.  Key names follow the reference
modules: ``encoder.model.*`` (hub DINOv2), ``info_sharing.*``
(``global_attention_transformer.py:103-150``), ``head1.*`` / ``uncertainty_head.*``
(``dpt.py:94-178``, ``dpt_block.py:21-80,128-218``), ``classification_head.*``
(``mlp_feature.py:55-70``), ``classification_bias`` (``ufm.py:827-828``).

LayerNorm affine terms, biases and LayerScale are drawn non-trivially (not the 1/0 constants of
a fresh reference model) so that a kernel that drops one of them fails parity.
"""
from __future__ import annotations

import hashlib
import math
from collections import OrderedDict
from typing import Any, Dict, List, Tuple

import numpy as np
import torch

from .configs import DINOV2_ARCH

Spec = List[Tuple[str, Tuple[int, ...], str]]  # (key, shape, kind)


def _block_spec(prefix: str, dim: int, hidden: int, layerscale: bool, qk_norm: bool, head_dim: int,
                qkv_bias: bool = True) -> Spec:
    s: Spec = [
        (prefix + "norm1.weight", (dim,), "ln_w"), (prefix + "norm1.bias", (dim,), "ln_b"),
        (prefix + "attn.qkv.weight", (3 * dim, dim), "w"),
    ]
    if qkv_bias:
        s.append((prefix + "attn.qkv.bias", (3 * dim,), "b"))
    if qk_norm:
        s += [(prefix + "attn.q_norm.weight", (head_dim,), "ln_w"), (prefix + "attn.q_norm.bias", (head_dim,), "ln_b"),
              (prefix + "attn.k_norm.weight", (head_dim,), "ln_w"), (prefix + "attn.k_norm.bias", (head_dim,), "ln_b")]
    s += [(prefix + "attn.proj.weight", (dim, dim), "w"), (prefix + "attn.proj.bias", (dim,), "b")]
    if layerscale:
        s.append((prefix + "ls1.gamma", (dim,), "gamma"))
    s += [(prefix + "norm2.weight", (dim,), "ln_w"), (prefix + "norm2.bias", (dim,), "ln_b"),
          (prefix + "mlp.fc1.weight", (hidden, dim), "w"), (prefix + "mlp.fc1.bias", (hidden,), "b"),
          (prefix + "mlp.fc2.weight", (dim, hidden), "w"), (prefix + "mlp.fc2.bias", (dim,), "b")]
    if layerscale:
        s.append((prefix + "ls2.gamma", (dim,), "gamma"))
    return s


def encoder_spec(cfg: Dict[str, Any]) -> Spec:
    ek = cfg["encoder_kwargs"]
    dim, depth, heads = DINOV2_ARCH[ek["size"]]
    if ek.get("keep_first_n_layers") is not None:
        depth = min(depth, ek["keep_first_n_layers"])
    p = "encoder.model."
    patch = ek.get("patch_size", 14)
    s: Spec = [
        (p + "cls_token", (1, 1, dim), "cls"),
        (p + "pos_embed", (1, 1 + 37 * 37, dim), "pos"),
        (p + "patch_embed.proj.weight", (dim, 3, patch, patch), "w"),
        (p + "patch_embed.proj.bias", (dim,), "b"),
    ]
    if ek.get("with_registers"):
        s.append((p + "register_tokens", (1, 4, dim), "cls"))
    for i in range(depth):
        s += _block_spec(f"{p}blocks.{i}.", dim, 4 * dim, True, False, dim // heads)
    s += [(p + "norm.weight", (dim,), "ln_w"), (p + "norm.bias", (dim,), "ln_b")]
    return s


def info_sharing_spec(cfg: Dict[str, Any]) -> Spec:
    ik = cfg["info_sharing_kwargs"]
    dim, depth, heads = ik["dim"], ik["depth"], ik["num_heads"]
    hidden = int(dim * ik.get("mlp_ratio", 4.0))
    p = "info_sharing."
    s: Spec = []
    if ik["input_embed_dim"] != dim:
        s += [(p + "proj_embed.weight", (dim, ik["input_embed_dim"]), "w"), (p + "proj_embed.bias", (dim,), "b")]
    if cfg["info_sharing_str"] == "cross_attention":
        # MultiViewCrossAttentionTransformer (cross_attention_transformer.py:100-135): one CrossAttentionBlock stack per
        # view (transformer_blocks.py:406-519), no view embedding
        ls, qkn, hd, bias = bool(ik.get("init_values")), bool(ik.get("qk_norm")), dim // heads, ik.get("qkv_bias", True)
        for v in range(ik.get("num_views", 2)):
            for i in range(depth):
                q = f"{p}multi_view_branches.{v}.{i}."
                blk = _block_spec(q, dim, hidden, ls, qkn, hd, bias)
                # _block_spec = norm1, attn, [ls1], norm2, mlp, [ls2]; the cross block's MLP hangs off norm3 / ls3
                first = [e for e in blk if ".norm2." not in e[0] and ".mlp." not in e[0] and ".ls2." not in e[0]]
                s += first
                if ik.get("norm_cross_tokens", True):
                    s += [(q + "norm_y.weight", (dim,), "ln_w"), (q + "norm_y.bias", (dim,), "ln_b")]
                s += [(q + "norm2.weight", (dim,), "ln_w"), (q + "norm2.bias", (dim,), "ln_b")]
                for nm in ("projq", "projk", "projv"):
                    s.append((f"{q}cross_attn.{nm}.weight", (dim, dim), "w"))
                    if bias:
                        s.append((f"{q}cross_attn.{nm}.bias", (dim,), "b"))
                if qkn:
                    s += [(q + "cross_attn.q_norm.weight", (hd,), "ln_w"), (q + "cross_attn.q_norm.bias", (hd,), "ln_b"),
                          (q + "cross_attn.k_norm.weight", (hd,), "ln_w"), (q + "cross_attn.k_norm.bias", (hd,), "ln_b")]
                s += [(q + "cross_attn.proj.weight", (dim, dim), "w"), (q + "cross_attn.proj.bias", (dim,), "b")]
                if ls:
                    s.append((q + "ls2.gamma", (dim,), "gamma"))
                s += [(q + "norm3.weight", (dim,), "ln_w"), (q + "norm3.bias", (dim,), "ln_b"),
                      (q + "mlp.fc1.weight", (hidden, dim), "w"), (q + "mlp.fc1.bias", (hidden,), "b"),
                      (q + "mlp.fc2.weight", (dim, hidden), "w"), (q + "mlp.fc2.bias", (dim,), "b")]
                if ls:
                    s.append((q + "ls3.gamma", (dim,), "gamma"))
        s += [(p + "norm.weight", (dim,), "ln_w"), (p + "norm.bias", (dim,), "ln_b")]
        return s
    for i in range(depth):
        s += _block_spec(f"{p}self_attention_blocks.{i}.", dim, hidden, bool(ik.get("init_values")),
                         bool(ik.get("qk_norm")), dim // heads, ik.get("qkv_bias", True))
    s += [(p + "norm.weight", (dim,), "ln_w"), (p + "norm.bias", (dim,), "ln_b")]
    if cfg["info_sharing_str"] == "alternating_attention":
        n_pos = ik.get("max_num_views_for_pe", 1000) if ik.get("use_pe_for_non_reference_views") else 1
    else:
        n_pos = ik["max_num_views"]
    s.append((p + "view_pos_table", (n_pos, dim), "view_pos"))
    return s


def dpt_head_spec(prefix: str, hk: Dict[str, Any]) -> Spec:
    """``nn.Sequential(nn.Sequential(DPTFeature, DPTRegressionProcessor), AdaptorMap)`` (``ufm.py:261-273``).

    The reference registers the four ``layer_rn`` convs under three aliases
    (``scratch.layerK_rn``, ``scratch.layer_rn.K-1``, ``input_process.K-1.1``); all three keys are
    emitted and share one tensor."""
    f = hk["dpt_feature"]
    in_dims, ld, fd = f["input_feature_dims"], f["layer_dims"], f["feature_dim"]
    p = prefix + "0.0."
    s: Spec = []
    for k in range(4):
        s.append((f"{p}scratch.layer{k + 1}_rn.weight", (fd, ld[k], 3, 3), "w"))
    for k in range(4):
        s.append((f"{p}scratch.layer_rn.{k}.weight", (fd, ld[k], 3, 3), f"alias:{p}scratch.layer{k + 1}_rn.weight"))
    for r in (1, 2, 3, 4):
        q = f"{p}scratch.refinenet{r}."
        s += [(q + "out_conv.weight", (fd, fd, 1, 1), "w"), (q + "out_conv.bias", (fd,), "b")]
        for u in ((1, 2) if r != 4 else (2,)):
            for c in (1, 2):
                s += [(f"{q}resConfUnit{u}.conv{c}.weight", (fd, fd, 3, 3), "w")]
                if not f.get("use_bn"):
                    s += [(f"{q}resConfUnit{u}.conv{c}.bias", (fd,), "b")]
            if f.get("use_bn"):  # dpt_block.py:134-149: bias-free convs, each followed by a BatchNorm2d
                for c in (1, 2):
                    b = f"{q}resConfUnit{u}.bn{c}."
                    s += [(b + "weight", (fd,), "ln_w"), (b + "bias", (fd,), "ln_b"), (b + "running_mean", (fd,), "bn_mean"),
                          (b + "running_var", (fd,), "bn_var"), (b + "num_batches_tracked", (), "bn_count")]
    # input_process.k = Sequential(act_k_postprocess, layer_rn[k])
    ip = p + "input_process."
    s += [(ip + "0.0.0.weight", (ld[0], in_dims[0], 1, 1), "w"), (ip + "0.0.0.bias", (ld[0],), "b"),
          (ip + "0.0.1.weight", (ld[0], ld[0], 4, 4), "w_convt"), (ip + "0.0.1.bias", (ld[0],), "b"),
          (ip + "0.1.weight", (fd, ld[0], 3, 3), f"alias:{p}scratch.layer1_rn.weight"),
          (ip + "1.0.0.weight", (ld[1], in_dims[1], 1, 1), "w"), (ip + "1.0.0.bias", (ld[1],), "b"),
          (ip + "1.0.1.weight", (ld[1], ld[1], 2, 2), "w_convt"), (ip + "1.0.1.bias", (ld[1],), "b"),
          (ip + "1.1.weight", (fd, ld[1], 3, 3), f"alias:{p}scratch.layer2_rn.weight"),
          (ip + "2.0.0.weight", (ld[2], in_dims[2], 1, 1), "w"), (ip + "2.0.0.bias", (ld[2],), "b"),
          (ip + "2.1.weight", (fd, ld[2], 3, 3), f"alias:{p}scratch.layer3_rn.weight"),
          (ip + "3.0.0.weight", (ld[3], in_dims[3], 1, 1), "w"), (ip + "3.0.0.bias", (ld[3],), "b"),
          (ip + "3.0.1.weight", (ld[3], ld[3], 3, 3), "w"), (ip + "3.0.1.bias", (ld[3],), "b"),
          (ip + "3.1.weight", (fd, ld[3], 3, 3), f"alias:{p}scratch.layer4_rn.weight")]
    pr = hk["dpt_processor"]
    cin, cout = pr["input_feature_dim"], pr["output_dim"]
    hid = pr.get("hidden_dims") or [cin // 2, cin // 2]
    q = prefix + "0.1."
    s += [(q + "conv1.weight", (hid[0], cin, 3, 3), "w"), (q + "conv1.bias", (hid[0],), "b"),
          (q + "conv2.0.weight", (hid[1], hid[0], 3, 3), "w"), (q + "conv2.0.bias", (hid[1],), "b"),
          (q + "conv2.2.weight", (cout, hid[1], 1, 1), "w_out"), (q + "conv2.2.bias", (cout,), "b")]
    return s


def adaptor_spec(prefix: str, adaptors_kwargs: Dict[str, Any]) -> Spec:
    s: Spec = []
    for name, a in adaptors_kwargs.items():
        if a["class"] == "FlowAdaptor":
            s += [(f"{prefix}1.adaptors.{name}.flow_mean", (1, 2, 1, 1), "flow_mean:" + name),
                  (f"{prefix}1.adaptors.{name}.flow_std", (1, 2, 1, 1), "flow_std:" + name)]
    return s


def model_spec(cfg: Dict[str, Any]) -> Spec:
    s = encoder_spec(cfg) + info_sharing_spec(cfg)
    s += dpt_head_spec("head1.", cfg["feature_head_kwargs"]) + adaptor_spec("head1.", cfg["adaptors_kwargs"])
    if cfg.get("uncertainty_head_kwargs"):
        s += dpt_head_spec("uncertainty_head.", cfg["uncertainty_head_kwargs"])
        s += adaptor_spec("uncertainty_head.", cfg.get("uncertainty_adaptors_kwargs", {}))
        # ufm.py:553/839 also registers the adaptor dict itself; it holds no tensors for MaskAdaptor
    if "classification_head_kwargs" in cfg:
        ck = cfg["classification_head_kwargs"]
        d = ck["input_feature_dim"]
        hidden = int(ck.get("mlp_ratio", 4) * d)
        p = "classification_head."
        s += [(p + "mlp.fc1.weight", (hidden, d), "w"), (p + "mlp.fc1.bias", (hidden,), "b"),
              (p + "mlp.fc2.weight", (d, hidden), "w"), (p + "mlp.fc2.bias", (d,), "b"),
              (p + "linear.weight", (ck["output_dim"] * ck["patch_size"] ** 2, d, 1, 1), "w_desc"),
              (p + "linear.bias", (ck["output_dim"] * ck["patch_size"] ** 2,), "b")]
        s = [("classification_bias", (cfg.get("refinement_range", 5) ** 2,), "cls_bias")] + s
        if cfg.get("use_unet_feature"):
            s += unet_spec(cfg)
    return s


UNET_FEATURES = (64, 128, 256, 512)  # hard-coded in the reference (``ufm.py:818``)


def unet_spec(cfg: Dict[str, Any]) -> Spec:
    """``UNet(3, 16, [64,128,256,512])`` (``uniflowmatch/models/unet_encoder.py:26-72``) plus the two 1x1 combine convs
    (``ufm.py:817-825``)."""
    s: Spec = []

    def double_conv(p, cin, cout):
        return [(p + "conv.0.weight", (cout, cin, 3, 3), "w_relu"), (p + "conv.0.bias", (cout,), "b"),
                (p + "conv.2.weight", (cout, cout, 3, 3), "w_relu"), (p + "conv.2.bias", (cout,), "b")]

    u = "unet_feature."
    cin = 3
    for k, f in enumerate(UNET_FEATURES):
        s += double_conv(f"{u}downs.{k}.", cin, f)
        cin = f
    s += double_conv(u + "bottleneck.", UNET_FEATURES[-1], 2 * UNET_FEATURES[-1])
    for k, f in enumerate(reversed(UNET_FEATURES)):
        s += [(f"{u}ups.{2 * k}.weight", (2 * f, f, 2, 2), "w_convt_up"), (f"{u}ups.{2 * k}.bias", (f,), "b")]
        s += double_conv(f"{u}ups.{2 * k + 1}.", 2 * f, f)
    od = cfg["classification_head_kwargs"]["output_dim"]
    s += [(u + "final_conv.weight", (od, UNET_FEATURES[0], 1, 1), "w"), (u + "final_conv.bias", (od,), "b")]
    method = cfg.get("feature_combine_method", "conv")
    s += [("conv1.weight", (32, 32, 1, 1), "w_relu"), ("conv1.bias", (32,), "b")]
    c2in = {"conv": 32, "modulate": 16}[method]
    s += [("conv2.weight", (16, c2in, 1, 1), "w"), ("conv2.bias", (16,), "b")]
    return s


def sinusoid_table(n_position: int, d_hid: int, base: float = 10000.0) -> torch.Tensor:
    """Restates ``_get_sinusoid_encoding_table`` (``global_attention_transformer.py:168-178``)."""
    pos = np.arange(n_position, dtype=np.float64)[:, None]
    j = np.arange(d_hid)[None, :]
    tab = pos / np.power(base, 2 * (j // 2) / d_hid)
    tab[:, 0::2] = np.sin(tab[:, 0::2])
    tab[:, 1::2] = np.cos(tab[:, 1::2])
    return torch.tensor(tab, dtype=torch.float32)


def _gen(seed: int, key: str) -> torch.Generator:
    h = hashlib.sha256(f"{seed}:{key}".encode()).digest()
    g = torch.Generator(device="cpu")
    g.manual_seed(int.from_bytes(h[:8], "little") & ((1 << 63) - 1))
    return g


def make_state_dict(cfg: Dict[str, Any], seed: int = 0, out_gain: float = 1.0) -> "OrderedDict[str, torch.Tensor]":
    """Build the fp32 CPU ``state_dict``.  ``out_gain`` scales the last 1x1 conv of each DPT head so the
    outputs have a realistic dynamic range (random-init flow spans only +-0.07 px)."""
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for key, shape, kind in model_spec(cfg):
        if kind.startswith("alias:"):
            sd[key] = sd[kind[6:]]
            continue
        g = _gen(seed, key)
        if kind in ("w", "w_convt", "w_out", "w_desc", "w_relu", "w_convt_up"):
            if kind == "w_convt":  # ConvTranspose2d weight is (in, out, kh, kw)
                fan_in, fan_out = shape[0], shape[1]
            elif kind == "w_convt_up":  # k == stride: every output pixel sees one tap of every input channel
                fan_in = fan_out = shape[0]
            else:
                rf = int(np.prod(shape[2:])) if len(shape) > 2 else 1
                fan_in, fan_out = shape[1] * rf, shape[0] * rf
            a = math.sqrt(6.0 / (fan_in + fan_out))
            if kind == "w_relu":  # He-uniform: keeps the activation scale through the 18-conv ReLU stack of the UNet
                a = math.sqrt(6.0 / fan_in)
            t = (torch.rand(shape, generator=g) * 2 - 1) * a
            if kind == "w_out":
                t = t * out_gain
            if kind == "w_desc":
                t = t * 4.0  # descriptors with O(1) correlation so the refinement softmax is not flat
        elif kind == "b":
            t = torch.randn(shape, generator=g) * 0.02
        elif kind == "ln_w":
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif kind == "ln_b":
            t = 0.05 * torch.randn(shape, generator=g)
        elif kind == "gamma":
            t = 0.5 + 0.1 * torch.randn(shape, generator=g)
        elif kind == "bn_mean":
            t = 0.1 * torch.randn(shape, generator=g)
        elif kind == "bn_var":
            t = 0.5 + torch.rand(shape, generator=g)
        elif kind == "bn_count":
            t = torch.tensor(100, dtype=torch.long)
        elif kind == "cls":
            t = 0.02 * torch.randn(shape, generator=g)
        elif kind == "pos":
            t = 0.02 * torch.randn(shape, generator=g)
        elif kind == "cls_bias":
            t = 0.3 * torch.randn(shape, generator=g)
        elif kind == "view_pos":
            t = sinusoid_table(shape[0], shape[1])
        elif kind.startswith("flow_mean:"):
            t = torch.tensor(cfg["adaptors_kwargs"][kind[10:]]["kwargs"]["flow_mean"], dtype=torch.float32).view(1, 2, 1, 1)
        elif kind.startswith("flow_std:"):
            t = torch.tensor(cfg["adaptors_kwargs"][kind[9:]]["kwargs"]["flow_std"], dtype=torch.float32).view(1, 2, 1, 1)
        else:
            raise KeyError(kind)
        sd[key] = t.contiguous()
    return sd
