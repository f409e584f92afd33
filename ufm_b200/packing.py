"""Host-side weight packing: reference ``state_dict`` (SURVEY.md 8b key names) -> engine packed tensors.

One-time layout work at model load, not on the hot path: bf16 casts, conv weights to tap-major K layout with the
input channels padded to 64, the 1x1 -> ConvTranspose(k == stride) pairs of the DPT input stages folded into one
GEMM (``dpt.py:107-145``), the classification 1x1 conv rows permuted so the GEMM output is pixel-shuffle ordered
(``mlp_feature.py:62-70,99``), and the DINOv2 positional embedding interpolated for the inference resolution
(hub ``interpolate_pos_encoding``).
"""
from __future__ import annotations

import math
from typing import Any, Dict, Tuple

import torch
import torch.nn.functional as F

from . import _lib
from .configs import DINOV2_ARCH

PATCH_KPAD = 64


def _norm_indices(indices, depth):
    if isinstance(indices, int):
        return list(range(depth - indices, depth))
    return [depth + i if i < 0 else i for i in indices]


def inference_hw(cfg) -> Tuple[int, int]:
    res = cfg.get("inference_resolution") or [560, 420]
    if not isinstance(res[0], int):
        res = res[0]  # a list of trained resolutions (base.py:86-98): one engine per entry, this is the first
    return int(res[1]), int(res[0])  # (H, W)


def unc_adaptor_kinds(ua):
    """uncertainty_adaptors_kwargs -> [(engine adaptor kind, vmin, vmax)] in AdaptorMap order (adaptors.py:899-985)."""
    want = {"non_occluded_mask": "MaskAdaptor", "flow_cov": "Covariance2DAdaptor", "keypoint_confidence": "ConfidenceAdaptor"}
    out = []
    for name, a in ua.items():
        if want.get(name) != a["class"]:
            raise NotImplementedError(f"uncertainty adaptor {name!r} of class {a['class']!r}: UFM consumes {want} (ufm.py:644-660)")
        kw = a["kwargs"]
        if a["class"] == "MaskAdaptor":
            out.append((_lib.ADAPT_MASK, 0.0, 0.0))
        elif a["class"] == "Covariance2DAdaptor":
            if kw.get("parametrization", "exp_tanh") != "exp_tanh":
                raise ValueError(f"Invalid parametrization: {kw.get('parametrization')}")  # adaptors.py:963
            out.append((_lib.ADAPT_COV_EXP_TANH, 0.0, 0.0))
        else:
            kind = {"exp": _lib.ADAPT_CONF_EXP, "sigmoid": _lib.ADAPT_CONF_SIGMOID, "softmax": _lib.ADAPT_CONF_SOFTMAX}[kw["confidence_type"]]
            assert kw["vmin"] < kw["vmax"], "vmin must be less than vmax"  # adaptors.py:893
            out.append((kind, float(kw["vmin"]), float(kw["vmax"])))
    if not out or len(out) > 4:
        raise NotImplementedError("uncertainty head needs 1..4 adaptors")
    return out


def flow_affine(cfg, H, W, state_dict=None):
    """(scale, shift) of the FlowAdaptor.  flow_mean / flow_std are registered BUFFERS of the reference adaptor
    (adaptors.py:70-71, used at :104-112), so a checkpoint's values win over the constructor kwargs."""
    a = cfg["adaptors_kwargs"]
    if list(a.keys()) != ["flow"] or a["flow"]["class"] != "FlowAdaptor":
        raise NotImplementedError("engine flow head supports a single FlowAdaptor named 'flow' (ufm.py:645)")
    kw = a["flow"]["kwargs"]
    if kw.get("output_normalized_coordinate", False):
        # adaptors.py:110-118: 0.5 (x + 1) (W, H) + 0.5 - (px + 0.5, py + 0.5) = x W/2 + W/2 - px; the tail epilogue
        # subtracts the pixel index (ufm_config.flow_coord)
        return (0.5 * W, 0.5 * H), (0.5 * W, 0.5 * H)
    bh, bw = kw["base_shape"]
    sx, sy = {"none": (1.0, 1.0), "scale_width": (W / bw, W / bw), "scale_height": (H / bh, H / bh),
              "scale_both": (W / bw, H / bh)}[kw.get("scale_strategy", "none")]
    std, mean = kw["flow_std"], kw["flow_mean"]
    if state_dict is not None:
        bm, bs = (state_dict.get(f"head1.1.adaptors.flow.{k}") for k in ("flow_mean", "flow_std"))
        if bm is not None:
            mean = [float(v) for v in bm.detach().float().flatten().cpu()]
        if bs is not None:
            std = [float(v) for v in bs.detach().float().flatten().cpu()]
    return (std[0] * sx, std[1] * sy), (mean[0] * sx, mean[1] * sy)


def view_index_range(cfg: Dict[str, Any]):
    """How the non-reference view picks its row of ``view_pos_table``: ("fixed", row) or ("rand", high) where the reference
    draws ``torch.randint(low=1, high=high, size=(1,))`` on every forward (global_attention_transformer.py:520-524,
    alternating_attention_transformer.py:281-286); ("none", 0) = no positional encoding on the non-reference view."""
    ik = cfg["info_sharing_kwargs"]
    kind = cfg["info_sharing_str"]
    if kind == "global_attention":
        if ik.get("use_rand_idx_pe_for_non_reference_views", False):
            return "rand", int(ik["max_num_views"])
        return "fixed", 1
    if kind == "alternating_attention":
        if not ik.get("use_pe_for_non_reference_views", False):
            return "none", 0
        if ik.get("use_rand_idx_pe_for_non_reference_views", True):
            return "rand", int(ik.get("max_num_views_for_pe", 1000))
        return "fixed", 1
    return "none", 0  # cross attention: no view table


def engine_config(cfg: Dict[str, Any], state_dict=None) -> "_lib.UfmConfig":
    H, W = inference_hw(cfg)
    ek, ik = cfg["encoder_kwargs"], cfg["info_sharing_kwargs"]
    if cfg.get("encoder_str", "dinov2") != "dinov2":
        raise NotImplementedError("engine encoder is DINOv2 ViT-S/B/L (with or without registers)")
    if cfg["info_sharing_str"] not in ("global_attention", "alternating_attention", "cross_attention"):
        raise NotImplementedError(f"info_sharing_str={cfg['info_sharing_str']}")
    if cfg["info_sharing_str"] == "cross_attention":
        if ik.get("num_views", 2) != 2:
            raise AssertionError("Expected 2 views")  # cross_attention_transformer.py:382-384 with UFM's two views
    cpe = ik.get("custom_positional_encoding")
    if cpe is not None:
        # only the global-attention transformer maps the string "rope" to RoPE2D (global_attention_transformer.py:110-115);
        # the other two take a callable, which a config.json cannot carry
        if cfg["info_sharing_str"] != "global_attention" or not isinstance(cpe, str):
            raise NotImplementedError("custom_positional_encoding must be the string 'rope' with global_attention")
        if cpe != "rope":
            raise ValueError(f"Unknown custom positional encoding: {cpe}")  # global_attention_transformer.py:115
    if cfg.get("head_type", "dpt") != "dpt":
        raise NotImplementedError("head_type must be 'dpt'")
    if cfg.get("use_unet_feature") and "classification_head_kwargs" not in cfg:
        raise ValueError("use_unet_feature belongs to UniFlowMatchClassificationRefinement (ufm.py:742)")
    dim, depth, heads = DINOV2_ARCH[ek["size"]]
    if ek.get("keep_first_n_layers") is not None:
        depth = min(depth, ek["keep_first_n_layers"])
    take = _norm_indices(ek["indices"], depth)
    c = _lib.UfmConfig()
    c.img_h, c.img_w, c.patch = H, W, ek.get("patch_size", 14)
    c.enc_dim, c.enc_heads = dim, heads
    c.enc_take_first, c.enc_take_last = take[0], take[-1]
    c.enc_depth = take[-1] + 1  # blocks after the last returned index do not influence any output
    c.info_dim, c.info_depth, c.info_heads = ik["dim"], ik["depth"], ik["num_heads"]
    c.info_hidden = int(ik["dim"] * ik.get("mlp_ratio", 4.0))
    it = _norm_indices(ik["indices"], ik["depth"])
    if len(it) < 2:
        raise ValueError("info_sharing indices must name at least two layers (ufm.py:605-606)")
    c.info_take0, c.info_take1 = it[0], it[1]
    c.info_alternating = {"global_attention": 0, "alternating_attention": 1, "cross_attention": 2}[cfg["info_sharing_str"]]
    c.info_qk_norm = int(bool(ik.get("qk_norm")))
    c.info_rope = int(cpe == "rope")
    c.info_layerscale = int(bool(ik.get("init_values")))
    c.info_has_proj = 1
    c.enc_registers = 4 if ek.get("with_registers") else 0                    # hub dinov2_vit*14_reg (dinov2.py:24,49,74-79)
    c.enc_norm_intermediate = int(bool(ek.get("norm_intermediate", True)))    # dinov2.py:259-261
    c.info_norm_intermediate = int(bool(ik.get("norm_intermediate", True)))   # global_attention_transformer.py:549
    from .prepost import IMAGE_NORMALIZATION

    norm_name = ek.get("data_norm_type", "dinov2")
    if norm_name not in IMAGE_NORMALIZATION:
        raise ValueError(f"encoder data_norm_type {norm_name!r} is not in IMAGE_NORMALIZATION_DICT")
    mean, std = IMAGE_NORMALIZATION[norm_name]
    for i in range(3):
        c.img_mean[i], c.img_std[i] = float(mean[i]), float(std[i])
    f = cfg["feature_head_kwargs"]["dpt_feature"]
    pr = cfg["feature_head_kwargs"]["dpt_processor"]
    # DPTFeature.hooks (dpt.py:47,198-219): input stage k reads list_features[hooks[k]] of [enc_last, inter0, inter1, final]
    feat_dims = [dim, ik["dim"], ik["dim"], ik["dim"]]
    c.hooks_set = 1
    for name, dst in (("feature_head_kwargs", c.dpt_hooks), ("uncertainty_head_kwargs", c.unc_hooks)):
        hk = (cfg.get(name) or {}).get("dpt_feature", {})
        hooks = list(hk.get("hooks", [0, 1, 2, 3]))
        if len(hooks) != 4 or any(h not in (0, 1, 2, 3) for h in hooks):
            raise ValueError(f"{name}.dpt_feature.hooks={hooks}: four indices into the 4 layered features (dpt.py:198-205)")
        want = [feat_dims[h] for h in hooks]
        if hk and list(hk.get("input_feature_dims", want)) != want:
            raise AssertionError(f"Input feature dimension mismatch at hook: hooks {hooks} select dims {want}, "
                                 f"input_feature_dims is {hk.get('input_feature_dims')}")  # dpt.py:199-202
        for k in range(4):
            dst[k] = hooks[k]
    for i in range(4):
        c.dpt_layer_dims[i] = f["layer_dims"][i]
    c.dpt_feature_dim = f["feature_dim"]
    for hk in ("feature_head_kwargs", "uncertainty_head_kwargs"):
        for part in ("dpt_feature", "dpt_processor"):
            nl = (cfg.get(hk) or {}).get(part, {}).get("nonlinearity", "relu")
            if nl != "relu":
                raise NotImplementedError(f"{hk}.{part}.nonlinearity={nl!r}: the engine's DPT epilogues implement ReLU "
                                          "(dpt.py:55,246; dpt_block.py make_nonlinearity)")
    hd = pr.get("hidden_dims") or [pr["input_feature_dim"] // 2] * 2
    if hd[0] != hd[1]:
        raise NotImplementedError("DPT processor hidden dims must match")
    c.dpt_hidden = hd[0]
    c.flow_out_ch = pr["output_dim"]
    (s0, s1), (b0, b1) = flow_affine(cfg, H, W, state_dict)
    c.flow_scale[0], c.flow_scale[1], c.flow_shift[0], c.flow_shift[1] = s0, s1, b0, b1
    c.flow_coord = int(bool(cfg["adaptors_kwargs"]["flow"]["kwargs"].get("output_normalized_coordinate", False)))
    c.unc_out_ch = 0
    if cfg.get("uncertainty_head_kwargs"):
        # AdaptorMap order (ufm.py:553-556); names are the ones ufm.py:644-660 consumes
        ua = cfg.get("uncertainty_adaptors_kwargs", {})
        kinds = unc_adaptor_kinds(ua)
        c.unc_out_ch = sum(3 if k == _lib.ADAPT_COV_EXP_TANH else 1 for k, _, _ in kinds)
        if c.unc_out_ch != cfg["uncertainty_head_kwargs"]["dpt_processor"]["output_dim"]:
            raise ValueError("uncertainty head output_dim does not match the channels its adaptors require")
        if [k for k, _, _ in kinds] != [_lib.ADAPT_MASK]:
            c.unc_n_adaptors = len(kinds)
            for i, (k, vmin, vmax) in enumerate(kinds):
                c.unc_kind[i], c.unc_vmin[i], c.unc_vmax[i] = k, vmin, vmax
    c.refine = int("classification_head_kwargs" in cfg)
    if c.refine:
        ck = cfg["classification_head_kwargs"]
        c.cls_in = ck["input_feature_dim"]
        c.cls_hidden = int(ck.get("mlp_ratio", 4) * ck["input_feature_dim"])
        c.cls_out, c.cls_patch = ck["output_dim"], ck["patch_size"]
        c.refine_P = cfg.get("refinement_range", 5)
        c.temperature = float(cfg.get("temperature", 4))
        if cfg.get("use_unet_feature"):
            method = cfg.get("feature_combine_method", "conv")
            if method not in ("conv", "modulate"):
                raise ValueError(f"feature_combine_method={method!r}")  # ufm.py:821-824 leaves conv2 undefined otherwise
            if c.cls_out != 16:
                raise ValueError("the UNet branch emits 16 channels (ufm.py:818); classification output_dim must be 16")
            c.unet = {"conv": 1, "modulate": 2}[method]
    return c


def interpolate_pos_embed(pos_embed: torch.Tensor, h0: int, w0: int, offset: float = 0.1, antialias: bool = False) -> torch.Tensor:
    """Hub DINOv2 ``interpolate_pos_encoding``: bicubic, scale_factor=(n+0.1)/M, no antialias -> (1+h0*w0, D)."""
    N = pos_embed.shape[1] - 1
    M = int(math.sqrt(N))
    D = pos_embed.shape[-1]
    pe = pos_embed.float()
    if h0 * w0 == N and h0 == w0:
        return pe[0]
    patch = pe[:, 1:].reshape(1, M, M, D).permute(0, 3, 1, 2)
    patch = F.interpolate(patch, mode="bicubic", antialias=antialias,
                          scale_factor=(float(h0 + offset) / M, float(w0 + offset) / M))
    assert patch.shape[-2:] == (h0, w0)
    return torch.cat([pe[0, :1], patch.permute(0, 2, 3, 1).reshape(-1, D)], dim=0)


def _pad64(c):
    return (c + 63) // 64 * 64


def _conv3_pack(w: torch.Tensor) -> torch.Tensor:
    """(Cout, Cin, 3, 3) -> bf16 (Cout, 9 * pad64(Cin)), K index = (ky*3+kx) * Cpad + ci."""
    co, ci = w.shape[:2]
    t = torch.zeros(co, 9, _pad64(ci), dtype=torch.float32)
    t[:, :, :ci] = w.permute(0, 2, 3, 1).reshape(co, 9, ci)
    return t.reshape(co, -1).to(torch.bfloat16)


def _fold_convt(w1, b1, wt, bt):
    """1x1 conv (m<-c) followed by ConvTranspose2d(k == stride) (o<-m): one GEMM with rows ordered (kh, kw, o)."""
    w1 = w1.reshape(w1.shape[0], -1).double()
    wt = wt.double()
    k = wt.shape[-1]
    wf = torch.einsum("mokl,mc->kloc", wt, w1).reshape(k * k * wt.shape[1], w1.shape[1])
    bf = (torch.einsum("mokl,m->klo", wt, b1.double()) + bt.double()).reshape(-1)
    return wf.float().to(torch.bfloat16), bf.float()


def pack_weights(cfg: Dict[str, Any], sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Returns {engine name: CPU tensor (fp32 or bf16, contiguous)}."""
    sd = {k: v.detach().float().cpu() for k, v in sd.items()}
    ec = engine_config(cfg)
    out: Dict[str, torch.Tensor] = {}
    bf = lambda t: t.to(torch.bfloat16)
    hp, wp = ec.img_h // ec.patch, ec.img_w // ec.patch

    p = "encoder.model."
    D = ec.enc_dim
    wpe = sd[p + "patch_embed.proj.weight"].reshape(D, -1)
    kpad = (wpe.shape[1] + PATCH_KPAD - 1) // PATCH_KPAD * PATCH_KPAD
    wpad = torch.zeros(D, kpad)
    wpad[:, : wpe.shape[1]] = wpe
    out["enc.patch.w"] = bf(wpad)
    out["enc.patch.b"] = sd[p + "patch_embed.proj.bias"]
    reg = ec.enc_registers > 0  # hub *_reg models interpolate with antialias and without the 0.1 offset
    out["enc.pos"] = interpolate_pos_embed(sd[p + "pos_embed"], hp, wp, 0.0 if reg else 0.1, reg)
    # the rows in front of the patch tokens: cls + its positional row, then the register tokens (no positional embedding)
    prefix = [sd[p + "cls_token"].reshape(1, -1).float() + out["enc.pos"][:1]]
    if reg:
        prefix.append(sd[p + "register_tokens"].reshape(-1, D).float())
    out["enc.cls"] = torch.cat(prefix, dim=0).contiguous()

    def block(dst, src, layerscale, qk_norm):
        out[dst + ".ln1.w"], out[dst + ".ln1.b"] = sd[src + "norm1.weight"], sd[src + "norm1.bias"]
        out[dst + ".qkv.w"] = bf(sd[src + "attn.qkv.weight"])
        if src + "attn.qkv.bias" in sd:
            out[dst + ".qkv.b"] = sd[src + "attn.qkv.bias"]
        if qk_norm:
            out[dst + ".qn.w"], out[dst + ".qn.b"] = sd[src + "attn.q_norm.weight"], sd[src + "attn.q_norm.bias"]
            out[dst + ".kn.w"], out[dst + ".kn.b"] = sd[src + "attn.k_norm.weight"], sd[src + "attn.k_norm.bias"]
        out[dst + ".proj.w"], out[dst + ".proj.b"] = bf(sd[src + "attn.proj.weight"]), sd[src + "attn.proj.bias"]
        out[dst + ".ln2.w"], out[dst + ".ln2.b"] = sd[src + "norm2.weight"], sd[src + "norm2.bias"]
        out[dst + ".fc1.w"], out[dst + ".fc1.b"] = bf(sd[src + "mlp.fc1.weight"]), sd[src + "mlp.fc1.bias"]
        out[dst + ".fc2.w"], out[dst + ".fc2.b"] = bf(sd[src + "mlp.fc2.weight"]), sd[src + "mlp.fc2.bias"]
        if layerscale:
            out[dst + ".ls1"], out[dst + ".ls2"] = sd[src + "ls1.gamma"], sd[src + "ls2.gamma"]

    for i in range(ec.enc_depth):
        block(f"enc.{i}", f"{p}blocks.{i}.", True, False)
    out["enc.norm.w"], out["enc.norm.b"] = sd[p + "norm.weight"], sd[p + "norm.bias"]

    p = "info_sharing."
    d = ec.info_dim
    if p + "proj_embed.weight" in sd:
        out["info.proj.w"], out["info.proj.b"] = bf(sd[p + "proj_embed.weight"]), sd[p + "proj_embed.bias"]
    else:  # nn.Identity (global_attention_transformer.py:104-107) as an exact identity GEMM
        out["info.proj.w"], out["info.proj.b"] = bf(torch.eye(d)), torch.zeros(d)
    ik = cfg["info_sharing_kwargs"]
    if ec.info_alternating == 2:
        # per-view CrossAttentionBlock stacks (transformer_blocks.py:406-519): block() packs norm1 / attn / norm2 / mlp;
        # here the MLP hangs off norm3 / ls3 and the cross-attention off norm2 / ls2, projk and projv are fused
        ls, qkn = bool(ec.info_layerscale), bool(ec.info_qk_norm)
        for v in range(2):
            for i in range(ec.info_depth):
                src, dst = f"{p}multi_view_branches.{v}.{i}.", f"info.v{v}.{i}"
                g = lambda k: sd[src + k]
                out[dst + ".ln1.w"], out[dst + ".ln1.b"] = g("norm1.weight"), g("norm1.bias")
                out[dst + ".qkv.w"] = bf(g("attn.qkv.weight"))
                if src + "attn.qkv.bias" in sd:
                    out[dst + ".qkv.b"] = g("attn.qkv.bias")
                out[dst + ".proj.w"], out[dst + ".proj.b"] = bf(g("attn.proj.weight")), g("attn.proj.bias")
                if src + "norm_y.weight" in sd:  # nn.Identity() when norm_cross_tokens=False (transformer_blocks.py:463)
                    out[dst + ".lny.w"], out[dst + ".lny.b"] = g("norm_y.weight"), g("norm_y.bias")
                out[dst + ".ln2.w"], out[dst + ".ln2.b"] = g("norm2.weight"), g("norm2.bias")
                out[dst + ".cq.w"] = bf(g("cross_attn.projq.weight"))
                out[dst + ".ckv.w"] = bf(torch.cat([g("cross_attn.projk.weight"), g("cross_attn.projv.weight")], 0))
                if src + "cross_attn.projq.bias" in sd:
                    out[dst + ".cq.b"] = g("cross_attn.projq.bias")
                    out[dst + ".ckv.b"] = torch.cat([g("cross_attn.projk.bias"), g("cross_attn.projv.bias")], 0)
                out[dst + ".cproj.w"], out[dst + ".cproj.b"] = bf(g("cross_attn.proj.weight")), g("cross_attn.proj.bias")
                out[dst + ".ln3.w"], out[dst + ".ln3.b"] = g("norm3.weight"), g("norm3.bias")
                out[dst + ".fc1.w"], out[dst + ".fc1.b"] = bf(g("mlp.fc1.weight")), g("mlp.fc1.bias")
                out[dst + ".fc2.w"], out[dst + ".fc2.b"] = bf(g("mlp.fc2.weight")), g("mlp.fc2.bias")
                if qkn:
                    out[dst + ".qn.w"], out[dst + ".qn.b"] = g("attn.q_norm.weight"), g("attn.q_norm.bias")
                    out[dst + ".kn.w"], out[dst + ".kn.b"] = g("attn.k_norm.weight"), g("attn.k_norm.bias")
                    out[dst + ".cqn.w"], out[dst + ".cqn.b"] = g("cross_attn.q_norm.weight"), g("cross_attn.q_norm.bias")
                    out[dst + ".ckn.w"], out[dst + ".ckn.b"] = g("cross_attn.k_norm.weight"), g("cross_attn.k_norm.bias")
                if ls:
                    out[dst + ".ls1"], out[dst + ".ls2"], out[dst + ".ls3"] = g("ls1.gamma"), g("ls2.gamma"), g("ls3.gamma")
    else:
        # row 0 = the reference view; the other view adds table row `view_index` (1 by default; drawn per forward by the
        # host when use_rand_idx_pe_for_non_reference_views, see view_index_range / ufm_set_view_index), or nothing at all
        # (alternating attention without use_pe_for_non_reference_views): the engine then reads the appended zero row
        tab = sd[p + "view_pos_table"].float()
        out["info.viewpos"] = torch.cat([tab, torch.zeros(1, d)], dim=0).contiguous()
        for i in range(ec.info_depth):
            block(f"info.{i}", f"{p}self_attention_blocks.{i}.", bool(ec.info_layerscale), bool(ec.info_qk_norm))
    out["info.norm.w"], out["info.norm.b"] = sd[p + "norm.weight"], sd[p + "norm.bias"]

    def head(dst, src):
        ip = src + "0.0.input_process."
        out[dst + ".in0.w"], out[dst + ".in0.b"] = _fold_convt(sd[ip + "0.0.0.weight"], sd[ip + "0.0.0.bias"],
                                                               sd[ip + "0.0.1.weight"], sd[ip + "0.0.1.bias"])
        out[dst + ".in1.w"], out[dst + ".in1.b"] = _fold_convt(sd[ip + "1.0.0.weight"], sd[ip + "1.0.0.bias"],
                                                               sd[ip + "1.0.1.weight"], sd[ip + "1.0.1.bias"])
        out[dst + ".in2.w"] = bf(sd[ip + "2.0.0.weight"].reshape(sd[ip + "2.0.0.weight"].shape[0], -1))
        out[dst + ".in2.b"] = sd[ip + "2.0.0.bias"]
        out[dst + ".in3a.w"] = bf(sd[ip + "3.0.0.weight"].reshape(sd[ip + "3.0.0.weight"].shape[0], -1))
        out[dst + ".in3a.b"] = sd[ip + "3.0.0.bias"]
        w3 = sd[ip + "3.0.1.weight"]
        out[dst + ".in3b.w"] = bf(w3.permute(0, 2, 3, 1).reshape(w3.shape[0], -1))
        out[dst + ".in3b.b"] = sd[ip + "3.0.1.bias"]
        sc = src + "0.0.scratch."
        for k in range(4):
            out[f"{dst}.rn{k}.w"] = _conv3_pack(sd[f"{sc}layer{k + 1}_rn.weight"])
        for r in (1, 2, 3, 4):
            q = f"{sc}refinenet{r}."
            w = sd[q + "out_conv.weight"]
            out[f"{dst}.ref{r}.out.w"], out[f"{dst}.ref{r}.out.b"] = bf(w.reshape(w.shape[0], -1)), sd[q + "out_conv.bias"]
            for u in ((1, 2) if r != 4 else (2,)):
                for cc in (1, 2):
                    w = sd[f"{q}resConfUnit{u}.conv{cc}.weight"]
                    bnp = f"{q}resConfUnit{u}.bn{cc}."
                    if bnp + "weight" in sd:
                        # use_bn (dpt_block.py:134-149): eval-mode BatchNorm folded into the bias-free conv
                        g_ = sd[bnp + "weight"].double() / torch.sqrt(sd[bnp + "running_var"].double() + 1e-5)
                        b = (sd[bnp + "bias"].double() - g_ * sd[bnp + "running_mean"].double()).float()
                        w = (w.double() * g_.view(-1, 1, 1, 1)).float()
                    else:
                        b = sd[f"{q}resConfUnit{u}.conv{cc}.bias"]
                    out[f"{dst}.ref{r}.rcu{u}.c{cc}.w"] = _conv3_pack(w)
                    out[f"{dst}.ref{r}.rcu{u}.c{cc}.b"] = b
        q = src + "0.1."
        out[dst + ".proc.c1.w"], out[dst + ".proc.c1.b"] = _conv3_pack(sd[q + "conv1.weight"]), sd[q + "conv1.bias"]
        out[dst + ".proc.c2.w"], out[dst + ".proc.c2.b"] = _conv3_pack(sd[q + "conv2.0.weight"]), sd[q + "conv2.0.bias"]
        w = sd[q + "conv2.2.weight"]
        out[dst + ".proc.out.w"], out[dst + ".proc.out.b"] = w.reshape(w.shape[0], -1).contiguous(), sd[q + "conv2.2.bias"]

    head("h1", "head1.")
    if ec.unc_out_ch:
        head("hu", "uncertainty_head.")
    if ec.refine:
        p = "classification_head."
        out["cls.fc1.w"], out["cls.fc1.b"] = bf(sd[p + "mlp.fc1.weight"]), sd[p + "mlp.fc1.bias"]
        out["cls.fc2.w"], out["cls.fc2.b"] = bf(sd[p + "mlp.fc2.weight"]), sd[p + "mlp.fc2.bias"]
        P, od = ec.cls_patch, ec.cls_out
        w = sd[p + "linear.weight"].reshape(od, P, P, -1).permute(1, 2, 0, 3).reshape(P * P * od, -1)
        out["cls.lin.w"] = bf(w)
        out["cls.lin.b"] = sd[p + "linear.bias"].reshape(od, P, P).permute(1, 2, 0).reshape(-1)
        out["cls.bias"] = sd["classification_bias"]
    if ec.unet:
        # unet_encoder.py:26-72.  3x3 convs in the implicit-GEMM layout; ConvTranspose(k = s = 2) as a GEMM with rows
        # ordered (ky, kx, out) and its bias replicated over the 4 taps; the per-pixel 1x1 convs transposed to [in][out].
        u = "unet_feature."

        def dconv(dst, src, first=True):
            if first:
                out[dst + ".c1.w"], out[dst + ".c1.b"] = _conv3_pack(sd[src + "conv.0.weight"]), sd[src + "conv.0.bias"]
            out[dst + ".c2.w"], out[dst + ".c2.b"] = _conv3_pack(sd[src + "conv.2.weight"]), sd[src + "conv.2.bias"]

        out["unet.stem.w"] = sd[u + "downs.0.conv.0.weight"].reshape(64, 27)
        out["unet.stem.b"] = sd[u + "downs.0.conv.0.bias"]
        dconv("unet.d0", u + "downs.0.", first=False)
        for l in (1, 2, 3):
            dconv(f"unet.d{l}", f"{u}downs.{l}.")
        dconv("unet.bn", u + "bottleneck.")
        for k in range(4):
            l = 3 - k
            wt = sd[f"{u}ups.{2 * k}.weight"]  # (in, out, 2, 2)
            out[f"unet.u{l}.t.w"] = bf(wt.permute(2, 3, 1, 0).reshape(4 * wt.shape[1], wt.shape[0]))
            out[f"unet.u{l}.t.b"] = sd[f"{u}ups.{2 * k}.bias"].repeat(4)
            dconv(f"unet.u{l}", f"{u}ups.{2 * k + 1}.")
        t2 = lambda w: w.reshape(w.shape[0], w.shape[1]).t()
        out["unet.final.w"], out["unet.final.b"] = t2(sd[u + "final_conv.weight"]), sd[u + "final_conv.bias"]
        if ec.unet == 1:
            out["unet.conv1.w"], out["unet.conv1.b"] = t2(sd["conv1.weight"]), sd["conv1.bias"]
        out["unet.conv2.w"], out["unet.conv2.b"] = t2(sd["conv2.weight"]), sd["conv2.bias"]
    return {k: v.contiguous() for k, v in out.items()}
