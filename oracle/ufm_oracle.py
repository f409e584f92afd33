"""CPU fp32 restatement of the UFM forward pass (TEST INFRASTRUCTURE — see ``oracle/__init__.py``).

Plain functional PyTorch on CPU tensors, driven by a config dict (:mod:`ufm_b200.configs`) and a
``state_dict`` with the reference's key names.  Each function cites the reference file:line it
restates (paths relative to ``/root/reference``; ``UC/`` = ``UniCeption/uniception/``).  It is
pinned against the reference itself by ``oracle/make_golden.py`` / ``tests/test_oracle_*.py``.

Stage dictionary returned by :func:`forward_stages` (all fp32, channels-first like the reference):
  enc{k}        encoder feature at ``encoder_kwargs.indices[k]``  (2B, C, h, w), views stacked on batch
  info_inter{k} info-sharing intermediate k, (2, B, D, h, w) (view-major)
  info_final    (2, B, D, h, w)
  dpt1 / dptu   DPTFeature output of the flow / uncertainty head, (B, F, 8h, 8w)
  head1_raw     processor output of the flow head (B, 2, H, W); headu_raw (B, 1, H, W)
  flow_reg      regression flow after the adaptor (B, 2, H, W)
  covis         sigmoid mask (B, 1, H, W)
  desc          classification descriptors (2B, 16, H, W)            [Refine]
  residual, log_softmax, flow                                        [Refine; Base: flow == flow_reg]
"""
from __future__ import annotations

import math
from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

DINOV2_ARCH = {"small": (384, 12, 6), "base": (768, 12, 12), "large": (1024, 24, 16)}
# UC/models/encoders/image_normalizations.py:17-35 (only the entries with distinct values matter)
IMAGE_NORMALIZATION = {
    "dummy": ((0.0, 0.0, 0.0), (1.0, 1.0, 1.0)),
    "croco": ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225)),
    "dust3r": ((0.5, 0.5, 0.5), (0.5, 0.5, 0.5)),
    "dinov2": ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225)),
    "identity": ((0.0, 0.0, 0.0), (1.0, 1.0, 1.0)),
    "patch_embedder": ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225)),
    "radio": ((0.0, 0.0, 0.0), (1.0, 1.0, 1.0)),
    "sea_raft": ((0.0, 0.0, 0.0), (1.0 / 255, 1.0 / 255, 1.0 / 255)),
    "unimatch": ((0.0, 0.0, 0.0), (1.0 / 255, 1.0 / 255, 1.0 / 255)),
    "roma": ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225)),
    "cosmos": ((0.0, 0.0, 0.0), (0.5, 0.5, 0.5)),
}


# ----------------------------------------------------------------------------------------------
# transformer pieces
# ----------------------------------------------------------------------------------------------
def _ln(x, sd, prefix, eps=1e-6):
    return F.layer_norm(x, (x.shape[-1],), sd[prefix + "weight"], sd[prefix + "bias"], eps)


def rope2d(t, pos, base=100.0):
    """``RoPE2D.forward`` (UC/models/libs/croco/pos_embed.py:108-155; the compiled cuRoPE2D computes the same):
    t (B, heads, N, hd), pos (B, N, 2) = (y, x) of each token.  The first hd/2 dims rotate by y, the last hd/2 by x."""
    D = t.shape[-1] // 2
    inv_freq = 1.0 / (base ** (torch.arange(0, D, 2).float() / D))

    def rope1d(v, p1):
        fr = torch.einsum("bn,j->bnj", p1.float(), inv_freq)
        fr = torch.cat((fr, fr), dim=-1)[:, None]  # (B, 1, N, D)
        v1, v2 = v[..., : D // 2], v[..., D // 2:]
        return v * fr.cos() + torch.cat((-v2, v1), dim=-1) * fr.sin()

    y, x = t.chunk(2, dim=-1)
    return torch.cat((rope1d(y, pos[:, :, 0]), rope1d(x, pos[:, :, 1])), dim=-1)


def grid_positions(B, h, w, views=1):
    """``PositionGetter`` (UC/models/utils/positional_encoding.py:8-23), repeated per view
    (global_attention_transformer.py:499-503)."""
    g = torch.cartesian_prod(torch.arange(h), torch.arange(w)).view(1, h * w, 2)
    return g.repeat(B, views, 1)


def attention(x, sd, prefix, heads, qk_norm=False, pos=None):
    """UC/models/utils/transformer_blocks.py:150-187 (and the DINOv2 SDPA wrapper, dinov2.py:128-141):
    fused qkv Linear -> (B, heads, N, hd) -> [LayerNorm over head_dim on q,k] -> [RoPE2D on q,k] ->
    softmax(q k^T / sqrt(hd)) v -> proj."""
    B, N, C = x.shape
    hd = C // heads
    qkv = F.linear(x, sd[prefix + "qkv.weight"], sd.get(prefix + "qkv.bias"))
    qkv = qkv.reshape(B, N, 3, heads, hd).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]
    if qk_norm:
        # norm_layer = partial(nn.LayerNorm, eps=1e-6) is handed down to q_norm/k_norm (transformer_blocks.py:142-143)
        q = F.layer_norm(q, (hd,), sd[prefix + "q_norm.weight"], sd[prefix + "q_norm.bias"], 1e-6)
        k = F.layer_norm(k, (hd,), sd[prefix + "k_norm.weight"], sd[prefix + "k_norm.bias"], 1e-6)
    if pos is not None:
        q, k = rope2d(q, pos), rope2d(k, pos)
    o = F.scaled_dot_product_attention(q, k, v, scale=hd ** -0.5)
    o = o.transpose(1, 2).reshape(B, N, C)
    return F.linear(o, sd[prefix + "proj.weight"], sd[prefix + "proj.bias"])


def block(x, sd, prefix, heads, qk_norm=False, pos=None):
    """Pre-LN block: ``x += ls1(attn(norm1 x)); x += ls2(mlp(norm2 x))``
    (transformer_blocks.py:386-403; hub DINOv2 ``Block.forward``; LayerScale = ``x * gamma`` :315-317)."""
    a = attention(_ln(x, sd, prefix + "norm1."), sd, prefix + "attn.", heads, qk_norm, pos)
    if prefix + "ls1.gamma" in sd:
        a = a * sd[prefix + "ls1.gamma"]
    x = x + a
    h = F.linear(_ln(x, sd, prefix + "norm2."), sd[prefix + "mlp.fc1.weight"], sd[prefix + "mlp.fc1.bias"])
    h = F.gelu(h)  # nn.GELU() default = erf form (transformer_blocks.py:65-88)
    h = F.linear(h, sd[prefix + "mlp.fc2.weight"], sd[prefix + "mlp.fc2.bias"])
    if prefix + "ls2.gamma" in sd:
        h = h * sd[prefix + "ls2.gamma"]
    return x + h


def interpolate_pos_embed(pos_embed: torch.Tensor, h0: int, w0: int, offset: float = 0.1, antialias: bool = False) -> torch.Tensor:
    """Hub DINOv2 ``interpolate_pos_encoding`` (third-party): bicubic,
    ``scale_factor=((n+0.1)/M)``, no antialias.  ``h0, w0`` = patch-grid rows, cols.  Returns (1, 1+h0*w0, D).
    Note the upstream argument naming: its "w" is the image height."""
    N = pos_embed.shape[1] - 1
    M = int(math.sqrt(N))
    D = pos_embed.shape[-1]
    if h0 * w0 == N and h0 == w0:
        return pos_embed.float()
    pe = pos_embed.float()
    patch = pe[:, 1:].reshape(1, M, M, D).permute(0, 3, 1, 2)
    patch = F.interpolate(patch, mode="bicubic", antialias=antialias,
                          scale_factor=(float(h0 + offset) / M, float(w0 + offset) / M))
    assert patch.shape[-2:] == (h0, w0)
    patch = patch.permute(0, 2, 3, 1).reshape(1, -1, D)
    return torch.cat([pe[:, :1], patch], dim=1)


def encoder(img: torch.Tensor, sd, cfg) -> List[torch.Tensor]:
    """``DINOv2IntermediateFeatureReturner.forward`` (UC/models/encoders/dinov2.py:233-266) over the hub
    ``get_intermediate_layers(x, n=indices, reshape=True, norm=True)``: patch-embed conv14/s14, +cls,
    +interpolated pos-embed, L blocks, FINAL norm on every returned layer, drop cls, -> NCHW."""
    ek = cfg["encoder_kwargs"]
    dim, depth, heads = DINOV2_ARCH[ek["size"]]
    if ek.get("keep_first_n_layers") is not None:
        depth = min(depth, ek["keep_first_n_layers"])
    patch = ek.get("patch_size", 14)
    p = "encoder.model."
    B, C, H, W = img.shape
    assert C == 3 and H % patch == 0 and W % patch == 0, "Input shape must be divisible by patch size"
    h0, w0 = H // patch, W // patch
    x = F.conv2d(img, sd[p + "patch_embed.proj.weight"], sd[p + "patch_embed.proj.bias"], stride=patch)
    x = x.flatten(2).transpose(1, 2)
    x = torch.cat([sd[p + "cls_token"].expand(B, -1, -1), x], dim=1)
    reg = bool(ek.get("with_registers"))  # hub *_reg models: 4 register tokens, antialiased interpolation, offset 0
    x = x + interpolate_pos_embed(sd[p + "pos_embed"], h0, w0, 0.0 if reg else 0.1, reg)
    R = 0
    if reg:
        rt = sd[p + "register_tokens"]
        R = rt.shape[1]
        x = torch.cat([x[:, :1], rt.expand(B, -1, -1), x[:, 1:]], dim=1)
    indices = ek["indices"]
    if isinstance(indices, int):
        take = list(range(depth - indices, depth))
    else:
        take = [depth + i if i < 0 else i for i in indices]
    outs = []
    for i in range(depth):
        x = block(x, sd, f"{p}blocks.{i}.", heads)
        if i in take:
            o = _ln(x, sd, p + "norm.") if ek.get("norm_intermediate", True) else x
            outs.append(o[:, 1 + R:].reshape(B, h0, w0, dim).permute(0, 3, 1, 2).contiguous())
    return outs


def cross_attention(q_in, kv_in, sd, prefix, heads, qk_norm=False):
    """``CrossAttention.forward`` (UC/models/utils/transformer_blocks.py:190-283): separate projq / projk / projv Linears,
    [LayerNorm over head_dim on q, k], softmax(q k^T / sqrt(hd)) v, proj."""
    B, Nq, C = q_in.shape
    hd = C // heads
    split = lambda t: t.reshape(B, -1, heads, hd).permute(0, 2, 1, 3)
    q = split(F.linear(q_in, sd[prefix + "projq.weight"], sd.get(prefix + "projq.bias")))
    k = split(F.linear(kv_in, sd[prefix + "projk.weight"], sd.get(prefix + "projk.bias")))
    v = split(F.linear(kv_in, sd[prefix + "projv.weight"], sd.get(prefix + "projv.bias")))
    if qk_norm:
        q = F.layer_norm(q, (hd,), sd[prefix + "q_norm.weight"], sd[prefix + "q_norm.bias"], 1e-6)
        k = F.layer_norm(k, (hd,), sd[prefix + "k_norm.weight"], sd[prefix + "k_norm.bias"], 1e-6)
    o = F.scaled_dot_product_attention(q, k, v, scale=hd ** -0.5).transpose(1, 2).reshape(B, Nq, C)
    return F.linear(o, sd[prefix + "proj.weight"], sd[prefix + "proj.bias"])


def cross_block(x, y, sd, prefix, heads, qk_norm=False):
    """``CrossAttentionBlock.forward`` (transformer_blocks.py:489-519): x += ls1(attn(norm1 x));
    x += ls2(cross_attn(norm2 x, norm_y y, norm_y y)); x += ls3(mlp(norm3 x))."""
    g = lambda name, t: t * sd[prefix + name] if prefix + name in sd else t
    x = x + g("ls1.gamma", attention(_ln(x, sd, prefix + "norm1."), sd, prefix + "attn.", heads, qk_norm))
    y_ = _ln(y, sd, prefix + "norm_y.") if prefix + "norm_y.weight" in sd else y
    x = x + g("ls2.gamma", cross_attention(_ln(x, sd, prefix + "norm2."), y_, sd, prefix + "cross_attn.", heads, qk_norm))
    h = F.gelu(F.linear(_ln(x, sd, prefix + "norm3."), sd[prefix + "mlp.fc1.weight"], sd[prefix + "mlp.fc1.bias"]))
    h = F.linear(h, sd[prefix + "mlp.fc2.weight"], sd[prefix + "mlp.fc2.bias"])
    return x + g("ls3.gamma", h)


def info_sharing_cross(f1: torch.Tensor, f2: torch.Tensor, sd, cfg):
    """``MultiViewCrossAttentionTransformerIFR.forward`` (UC/models/info_sharing/cross_attention_transformer.py:360-478)
    for 2 views: per-view token streams, proj_embed, then at every depth BOTH views are updated from the previous
    depth's features (view v runs its own branch with the other view's tokens as cross tokens); final norm, and the same
    norm on the intermediates at ``indices``.  Returns (final (2,B,D,h,w), [intermediate (2,B,D,h,w)])."""
    ik = cfg["info_sharing_kwargs"]
    dim, depth, heads = ik["dim"], ik["depth"], ik["num_heads"]
    p = "info_sharing."
    B, Cin, h, w = f1.shape
    xs = [f.permute(0, 2, 3, 1).reshape(B, h * w, Cin) for f in (f1, f2)]
    if p + "proj_embed.weight" in sd:
        xs = [F.linear(x, sd[p + "proj_embed.weight"], sd[p + "proj_embed.bias"]) for x in xs]
    idx = ik["indices"]
    take = list(range(depth - idx, depth)) if isinstance(idx, int) else [depth + i if i < 0 else i for i in idx]
    inter = []
    qk_norm = bool(ik.get("qk_norm"))
    to_views = lambda ts: torch.stack([t.reshape(B, h, w, dim).permute(0, 3, 1, 2) for t in ts], 0).contiguous()
    for i in range(depth):
        xs = [cross_block(xs[v], xs[1 - v], sd, f"{p}multi_view_branches.{v}.{i}.", heads, qk_norm) for v in (0, 1)]
        if i in take:
            inter.append(to_views([_ln(x, sd, p + "norm.") if ik.get("norm_intermediate", True) else x for x in xs]))
    return to_views([_ln(x, sd, p + "norm.") for x in xs]), inter


def info_sharing(f1: torch.Tensor, f2: torch.Tensor, sd, cfg):
    """``MultiViewGlobalAttentionTransformerIFR.forward`` (UC/models/info_sharing/global_attention_transformer.py:428-609)
    and ``MultiViewAlternatingAttentionTransformerIFR.forward`` (alternating_attention_transformer.py:490-728) for 2 views:
    tokens (B, 2hw, C_in) -> proj_embed -> + view_pos_table row per view -> blocks (alternating: odd layers attend
    within each view) -> final norm; intermediates at ``indices`` get the same final norm.
    Returns (final (2,B,D,h,w), [intermediate (2,B,D,h,w)])."""
    if cfg["info_sharing_str"] == "cross_attention":
        return info_sharing_cross(f1, f2, sd, cfg)
    ik = cfg["info_sharing_kwargs"]
    dim, depth, heads = ik["dim"], ik["depth"], ik["num_heads"]
    alternating = cfg["info_sharing_str"] == "alternating_attention"
    p = "info_sharing."
    B, Cin, h, w = f1.shape
    n = h * w
    x = torch.stack([f1, f2], dim=1).permute(0, 1, 3, 4, 2).reshape(B, 2 * n, Cin)
    if p + "proj_embed.weight" in sd:
        x = F.linear(x, sd[p + "proj_embed.weight"], sd[p + "proj_embed.bias"])
    tab = sd[p + "view_pos_table"]
    pe = torch.zeros(2, dim)
    pe[0] = tab[0]
    if not alternating or ik.get("use_pe_for_non_reference_views", False):
        # use_rand_idx_pe_for_non_reference_views: one draw from the global generator per forward, as the reference
        # (global_attention_transformer.py:520-524; alternating_attention_transformer.py:281-286, default True there)
        rand = ik.get("use_rand_idx_pe_for_non_reference_views", alternating)
        high = ik.get("max_num_views_for_pe", 1000) if alternating else ik.get("max_num_views")
        row = int(torch.randint(low=1, high=high, size=(1,)).item()) if rand else 1
        pe[1] = tab[row]
    x = x + pe.repeat_interleave(n, dim=0)[None]
    idx = ik["indices"]
    take = list(range(depth - idx, depth)) if isinstance(idx, int) else [depth + i if i < 0 else i for i in idx]
    inter = []
    qk_norm = bool(ik.get("qk_norm"))
    # custom_positional_encoding="rope" (global_attention_transformer.py:110-113): RoPE2D on q, k of every block
    pos = grid_positions(B, h, w, views=2) if ik.get("custom_positional_encoding") == "rope" and not alternating else None
    for i in range(depth):
        pre = f"{p}self_attention_blocks.{i}."
        if alternating and i % 2 == 1:
            x = block(x.reshape(B * 2, n, dim), sd, pre, heads, qk_norm).reshape(B, 2 * n, dim)
        else:
            x = block(x, sd, pre, heads, qk_norm, pos)
        if i in take:
            inter.append(_ln(x, sd, p + "norm.") if ik.get("norm_intermediate", True) else x)

    def to_views(t):
        return t.reshape(B, 2, h, w, dim).permute(1, 0, 4, 2, 3).contiguous()

    return to_views(_ln(x, sd, p + "norm.")), [to_views(t) for t in inter]


# ----------------------------------------------------------------------------------------------
# DPT head
# ----------------------------------------------------------------------------------------------
def _rcu(x, sd, prefix):
    """``ResidualConvUnit_custom.forward`` (UC/models/libs/croco/dpt_block.py:156-177): x + conv2(relu(conv1(relu(x))));
    with ``use_bn`` the convs have no bias and an eval-mode BatchNorm follows each (:134-149, 164-171)."""
    def bn(o, k):
        if prefix + k + ".weight" not in sd:
            return o
        return F.batch_norm(o, sd[prefix + k + ".running_mean"], sd[prefix + k + ".running_var"], sd[prefix + k + ".weight"],
                            sd[prefix + k + ".bias"], training=False, eps=1e-5)

    o = bn(F.conv2d(F.relu(x), sd[prefix + "conv1.weight"], sd.get(prefix + "conv1.bias"), padding=1), "bn1")
    o = bn(F.conv2d(F.relu(o), sd[prefix + "conv2.weight"], sd.get(prefix + "conv2.bias"), padding=1), "bn2")
    return o + x


def _fusion(sd, prefix, path, skip=None):
    """``FeatureFusionBlock_custom.forward`` (dpt_block.py:225-255), width_ratio == 1."""
    out = path
    if skip is not None:
        out = out + _rcu(skip, sd, prefix + "resConfUnit1.")
    out = _rcu(out, sd, prefix + "resConfUnit2.")
    out = F.interpolate(out, scale_factor=2, mode="bilinear", align_corners=True)
    return F.conv2d(out, sd[prefix + "out_conv.weight"], sd[prefix + "out_conv.bias"])


def dpt_feature(feats: List[torch.Tensor], sd, prefix, hooks=(0, 1, 2, 3)) -> torch.Tensor:
    """``DPTFeature.forward`` (UC/models/prediction_heads/dpt.py:180-232) with ``init`` layers (:94-178); input stage k
    reads ``list_features[hooks[k]]`` (:198-205)."""
    feats = [feats[h] for h in hooks]
    ip = prefix + "input_process."
    l0 = F.conv2d(feats[0], sd[ip + "0.0.0.weight"], sd[ip + "0.0.0.bias"])
    l0 = F.conv_transpose2d(l0, sd[ip + "0.0.1.weight"], sd[ip + "0.0.1.bias"], stride=4)
    l0 = F.conv2d(l0, sd[ip + "0.1.weight"], None, padding=1)
    l1 = F.conv2d(feats[1], sd[ip + "1.0.0.weight"], sd[ip + "1.0.0.bias"])
    l1 = F.conv_transpose2d(l1, sd[ip + "1.0.1.weight"], sd[ip + "1.0.1.bias"], stride=2)
    l1 = F.conv2d(l1, sd[ip + "1.1.weight"], None, padding=1)
    l2 = F.conv2d(feats[2], sd[ip + "2.0.0.weight"], sd[ip + "2.0.0.bias"])
    l2 = F.conv2d(l2, sd[ip + "2.1.weight"], None, padding=1)
    l3 = F.conv2d(feats[3], sd[ip + "3.0.0.weight"], sd[ip + "3.0.0.bias"])
    l3 = F.conv2d(l3, sd[ip + "3.0.1.weight"], sd[ip + "3.0.1.bias"], stride=2, padding=1)
    l3 = F.conv2d(l3, sd[ip + "3.1.weight"], None, padding=1)
    sc = prefix + "scratch."
    p4 = _fusion(sd, sc + "refinenet4.", l3)[:, :, : l2.shape[2], : l2.shape[3]]
    p3 = _fusion(sd, sc + "refinenet3.", p4, l2)
    p2 = _fusion(sd, sc + "refinenet2.", p3, l1)
    return _fusion(sd, sc + "refinenet1.", p2, l0)


def dpt_processor(x, sd, prefix, out_hw) -> torch.Tensor:
    """``DPTRegressionProcessor.forward`` (dpt.py:285-311): conv3x3 -> bilinear(align_corners=True) to image ->
    conv3x3 -> ReLU -> conv1x1."""
    x = F.conv2d(x, sd[prefix + "conv1.weight"], sd[prefix + "conv1.bias"], padding=1)
    x = F.interpolate(x, size=tuple(out_hw), mode="bilinear", align_corners=True)
    x = F.relu(F.conv2d(x, sd[prefix + "conv2.0.weight"], sd[prefix + "conv2.0.bias"], padding=1))
    return F.conv2d(x, sd[prefix + "conv2.2.weight"], sd[prefix + "conv2.2.bias"])


def flow_adaptor(x, kw, out_hw, buffers=None):
    """``FlowAdaptor.forward`` (UC/models/prediction_heads/adaptors.py:77-141).  ``buffers`` = (flow_mean, flow_std) as
    loaded from the state_dict: they are registered buffers (:70-71), so a checkpoint's values replace the kwargs."""
    H, W = out_hw
    if kw.get("output_normalized_coordinate", False):
        x = 0.5 * (x + 1) * torch.tensor([W, H], dtype=torch.float32).view(1, 2, 1, 1) + 0.5
        xs = torch.arange(W, dtype=torch.float32) + 0.5
        ys = torch.arange(H, dtype=torch.float32) + 0.5
        return x - torch.stack(torch.meshgrid(xs, ys, indexing="xy"), dim=0)
    strat = kw.get("scale_strategy", "none")
    bh, bw = kw["base_shape"]
    sx, sy = {"none": (1.0, 1.0), "scale_width": (W / bw, W / bw), "scale_height": (H / bh, H / bh),
              "scale_both": (W / bw, H / bh)}[strat]
    s = torch.tensor([sx, sy], dtype=torch.float32).view(1, 2, 1, 1)
    mean = torch.tensor(kw["flow_mean"], dtype=torch.float32).view(1, 2, 1, 1)
    std = torch.tensor(kw["flow_std"], dtype=torch.float32).view(1, 2, 1, 1)
    if buffers is not None:
        mean = buffers[0].float().view(1, 2, 1, 1) if buffers[0] is not None else mean
        std = buffers[1].float().view(1, 2, 1, 1) if buffers[1] is not None else std
    return x * (std * s) + mean * s


def covariance_adaptor(x):
    """``Covariance2DAdaptor.forward`` exp_tanh (adaptors.py:944-965) -> (cov, log_det, inv_cov)."""
    c1, c2, s = torch.split(x, 1, dim=1)
    t = s.tanh()
    de = (c1 + c2) / 2
    cov = torch.cat([c1.exp(), c2.exp(), t * torch.exp(de)], dim=1)
    log_det = c1 + c2 + torch.log(1 - t * t + 1e-8)
    inv = (1 / (1 - t * t + 1e-8)) * torch.cat([torch.exp(-c1), torch.exp(-c2), -t * torch.exp(-de)], dim=1)
    return cov, log_det, inv


def confidence_adaptor(x, kw):
    """``ConfidenceAdaptor.forward`` (adaptors.py:899-927)."""
    t, vmin, vmax = kw["confidence_type"], kw["vmin"], kw["vmax"]
    if t == "exp":
        return vmin + x.exp().clip(max=vmax - vmin)
    if t == "sigmoid":
        return torch.sigmoid(x) * (vmax - vmin) + vmin
    if t == "softmax":
        B, C, H, W = x.shape
        return F.softmax(x.reshape(B, C, -1), dim=-1).reshape(B, C, H, W) * (H * W)
    raise ValueError(t)


def run_adaptors(raw, adaptors_kwargs, out_hw, sd=None, prefix="head1.1.adaptors.") -> Dict[str, Any]:
    """``AdaptorMap.forward`` (UC/models/prediction_heads/base.py:186-210): split channels in dict order."""
    out, c = {}, 0
    for name, a in adaptors_kwargs.items():
        cls = a["class"]
        if cls == "FlowAdaptor":
            buf = None if sd is None else (sd.get(f"{prefix}{name}.flow_mean"), sd.get(f"{prefix}{name}.flow_std"))
            out[name] = flow_adaptor(raw[:, c:c + 2], a["kwargs"], out_hw, buf)
            c += 2
        elif cls == "MaskAdaptor":
            out[name] = {"logits": raw[:, c:c + 1], "mask": torch.sigmoid(raw[:, c:c + 1])}
            c += 1
        elif cls == "Covariance2DAdaptor":
            out[name] = covariance_adaptor(raw[:, c:c + 3])
            c += 3
        elif cls == "ConfidenceAdaptor":
            out[name] = confidence_adaptor(raw[:, c:c + 1], a["kwargs"])
            c += 1
        else:
            raise NotImplementedError(cls)
    return out


# ----------------------------------------------------------------------------------------------
# Refine: classification head + local refinement
# ----------------------------------------------------------------------------------------------
def mlp_feature(x, sd, prefix, patch):
    """``MLPFeature.forward`` (UC/models/prediction_heads/mlp_feature.py:77-101)."""
    t = x.permute(0, 2, 3, 1)
    t = F.linear(t, sd[prefix + "mlp.fc1.weight"], sd[prefix + "mlp.fc1.bias"])
    t = F.linear(F.gelu(t), sd[prefix + "mlp.fc2.weight"], sd[prefix + "mlp.fc2.bias"])
    t = t.permute(0, 3, 1, 2).contiguous()
    t = F.conv2d(t, sd[prefix + "linear.weight"], sd[prefix + "linear.bias"])
    return F.pixel_shuffle(t, patch)


def _double_conv(x, sd, p):
    """``DoubleConv.forward`` (uniflowmatch/models/unet_encoder.py:10-23)."""
    x = F.relu(F.conv2d(x, sd[p + "conv.0.weight"], sd[p + "conv.0.bias"], padding=1))
    return F.relu(F.conv2d(x, sd[p + "conv.2.weight"], sd[p + "conv.2.bias"], padding=1))


def unet(img, sd, prefix="unet_feature.", n_levels=4, return_last=False):
    """``UNet.forward`` (uniflowmatch/models/unet_encoder.py:50-72): 4 x [DoubleConv, maxpool 2], bottleneck, 4 x
    [ConvTranspose 2x2/s2, nearest resize to the skip if the sizes differ, cat(skip, x), DoubleConv], 1x1 conv."""
    skips = []
    x = img
    for k in range(n_levels):
        x = _double_conv(x, sd, f"{prefix}downs.{k}.")
        skips.append(x)
        x = F.max_pool2d(x, kernel_size=2, stride=2)
    x = _double_conv(x, sd, prefix + "bottleneck.")
    for k in range(n_levels):
        x = F.conv_transpose2d(x, sd[f"{prefix}ups.{2 * k}.weight"], sd[f"{prefix}ups.{2 * k}.bias"], stride=2)
        skip = skips[n_levels - 1 - k]
        if x.shape != skip.shape:
            x = F.interpolate(x, size=skip.shape[2:])  # default mode: nearest
        x = _double_conv(torch.cat((skip, x), dim=1), sd, f"{prefix}ups.{2 * k + 1}.")
    out = F.conv2d(x, sd[prefix + "final_conv.weight"], sd[prefix + "final_conv.bias"])
    return (out, x) if return_last else out


def combine_unet(cls_feat, unet_feat, sd, method):
    """ufm.py:967-982: ``conv``: conv2(relu(conv1(cat(cls, unet)))); ``modulate``: conv2(cls * tanh(unet))."""
    if method == "conv":
        t = F.relu(F.conv2d(torch.cat([cls_feat, unet_feat], dim=1), sd["conv1.weight"], sd["conv1.bias"]))
        return F.conv2d(t, sd["conv2.weight"], sd["conv2.bias"])
    if method == "modulate":
        return F.conv2d(cls_feat * torch.tanh(unet_feat), sd["conv2.weight"], sd["conv2.bias"])
    raise ValueError(method)


def _cubic_weights(t: torch.Tensor, A: float = -0.75):
    """PyTorch bicubic convolution coefficients for fractional offset t (taps at -1,0,1,2)."""
    def c1(x):
        return ((A + 2) * x - (A + 3)) * x * x + 1

    def c2(x):
        return ((A * x - 5 * A) * x + 8 * A) * x - 4 * A

    return torch.stack([c2(t + 1), c1(t), c1(1 - t), c2(2 - t)], dim=-1)


def local_refinement(flow, feat1, feat2, bias, temperature, P=5, chunk_rows=16):
    """``classification_refinement`` (uniflowmatch/models/ufm.py:1012-1039) =
    ``obtain_neighborhood_features`` (:1112-1178) + ``compute_refinement_attention`` (:1041-1095),
    restated without ``grid_sample``: for pixel (x,y) sample feat2 at p=(x+fx+dj, y+fy+di) with bicubic
    (A=-0.75, zero padding per tap, the normalise/denormalise of align_corners=False cancels), score_k = q.k/T + bias_k
    with k = 5a+b, di=a-R, dj=b-R; residual = sum softmax_k * (dj, di).
    Returns (residual (B,2,H,W), log_softmax (B,H,W,P,P))."""
    B, C, H, W = feat2.shape
    R = (P - 1) // 2
    res = torch.empty(B, 2, H, W)
    lsm = torch.empty(B, H, W, P, P)
    xs = torch.arange(W, dtype=torch.float32).view(1, 1, W)
    f2flat = feat2.permute(0, 2, 3, 1).reshape(B * H * W, C)
    boff = (torch.arange(B) * (H * W)).view(B, 1, 1)
    dj = torch.arange(-R, R + 1, dtype=torch.float32)
    for y0 in range(0, H, chunk_rows):
        y1 = min(H, y0 + chunk_rows)
        ys = torch.arange(y0, y1, dtype=torch.float32).view(1, -1, 1)
        # same float ops as the reference: ((p + 0.5)/S*2-1) then grid_sample's unnormalise ((g+1)*S-1)/2
        px = xs + flow[:, 0, y0:y1]
        py = ys + flow[:, 1, y0:y1]
        scores = torch.zeros(B, y1 - y0, W, P, P)
        q = feat1[:, :, y0:y1].permute(0, 2, 3, 1)  # B,h,W,C
        for a in range(P):
            for b in range(P):
                gx = ((px + dj[b]) + 0.5) / W * 2 - 1
                gy = ((py + dj[a]) + 0.5) / H * 2 - 1
                ix = ((gx + 1) * W - 1) / 2
                iy = ((gy + 1) * H - 1) / 2
                x0 = torch.floor(ix)
                y0f = torch.floor(iy)
                wx = _cubic_weights(ix - x0)
                wy = _cubic_weights(iy - y0f)
                k = torch.zeros(B, y1 - y0, W, C)
                for ty in range(4):
                    yy = (y0f + (ty - 1)).long()
                    vy = (yy >= 0) & (yy < H)
                    yyc = yy.clamp(0, H - 1)
                    for tx in range(4):
                        xx = (x0 + (tx - 1)).long()
                        v = vy & (xx >= 0) & (xx < W)
                        xxc = xx.clamp(0, W - 1)
                        val = f2flat[(boff + yyc * W + xxc).reshape(-1)].view(B, y1 - y0, W, C)
                        wgt = (wy[..., ty] * wx[..., tx] * v).unsqueeze(-1)
                        k = k + val * wgt
                scores[..., a, b] = (q * k).sum(-1)
        scores = scores / temperature + bias.view(1, 1, 1, P, P)
        flat = scores.reshape(B, y1 - y0, W, P * P)
        att = torch.softmax(flat, dim=-1).reshape(B, y1 - y0, W, P, P)
        lsm[:, y0:y1] = torch.log_softmax(flat, dim=-1).reshape(B, y1 - y0, W, P, P)
        res[:, 0, y0:y1] = (att * dj.view(1, 1, 1, 1, P)).sum((-1, -2))
        res[:, 1, y0:y1] = (att * dj.view(1, 1, 1, P, 1)).sum((-1, -2))
    return res, lsm


# ----------------------------------------------------------------------------------------------
# model forward
# ----------------------------------------------------------------------------------------------
def forward_stages(img1: torch.Tensor, img2: torch.Tensor, sd, cfg, stop_after: Optional[str] = None) -> Dict[str, torch.Tensor]:
    """``UniFlowMatchConfidence.forward`` (ufm.py:562-662) / ``UniFlowMatchClassificationRefinement.forward``
    (:843-1009) on normalised fp32 (B,3,H,W) images; returns every stage boundary."""
    if img1.shape[-2:] != img2.shape[-2:]:
        raise NotImplementedError("Unequal Image sizes are not supported now")  # ufm.py:316-317
    st: Dict[str, torch.Tensor] = {}
    B, _, H, W = img1.shape
    feats = encoder(torch.cat([img1, img2], dim=0), sd, cfg)  # ufm.py:291-319
    for k, f in enumerate(feats):
        st[f"enc{k}"] = f
    if stop_after == "enc":
        return st
    f1, f2 = feats[-1][:B], feats[-1][B:]
    final, inter = info_sharing(f1, f2, sd, cfg)
    st["info_final"] = final
    for k, t in enumerate(inter):
        st[f"info_inter{k}"] = t
    if stop_after == "info":
        return st
    head_in = [f1, inter[0][0], inter[1][0], final[0]]  # ufm.py:602-608, view 1 only
    hooks = lambda k: tuple((cfg.get(k) or {}).get("dpt_feature", {}).get("hooks", (0, 1, 2, 3)))
    st["dpt1"] = dpt_feature(head_in, sd, "head1.0.0.", hooks("feature_head_kwargs"))
    st["head1_raw"] = dpt_processor(st["dpt1"], sd, "head1.0.1.", (H, W))
    a1 = run_adaptors(st["head1_raw"], cfg["adaptors_kwargs"], (H, W), sd)
    st["flow_reg"] = a1["flow"]
    if cfg.get("uncertainty_head_kwargs"):
        st["dptu"] = dpt_feature(head_in, sd, "uncertainty_head.0.0.", hooks("uncertainty_head_kwargs"))
        st["headu_raw"] = dpt_processor(st["dptu"], sd, "uncertainty_head.0.1.", (H, W))
        au = run_adaptors(st["headu_raw"], cfg["uncertainty_adaptors_kwargs"], (H, W))
        if "non_occluded_mask" in au:
            st["covis"] = au["non_occluded_mask"]["mask"]
            st["covis_logits"] = au["non_occluded_mask"]["logits"]
        if "flow_cov" in au:
            st["flow_cov"], st["flow_cov_log_det"], st["flow_cov_inv"] = au["flow_cov"]
        if "keypoint_confidence" in au:
            st["keypoint_confidence"] = au["keypoint_confidence"]  # (B,1,H,W); ufm.py:654 squeezes dim 1
    st["flow"] = st["flow_reg"]
    if "classification_head_kwargs" in cfg:
        ck = cfg["classification_head_kwargs"]
        cf = torch.cat([torch.cat([feats[0][:B], final[0]], dim=1),
                        torch.cat([feats[0][B:], final[1]], dim=1)], dim=0)  # ufm.py:954-963
        st["desc"] = mlp_feature(cf, sd, "classification_head.", ck["patch_size"])
        if cfg.get("use_unet_feature"):
            st["desc_cls"] = st["desc"]
            st["unet_feat"], st["unet_x0"] = unet(torch.cat([img1, img2], dim=0), sd, return_last=True)  # ufm.py:916-918
            st["desc"] = combine_unet(st["desc_cls"], st["unet_feat"], sd, cfg.get("feature_combine_method", "conv"))
        res, lsm = local_refinement(st["flow_reg"], st["desc"][:B], st["desc"][B:], sd["classification_bias"],
                                    float(cfg.get("temperature", 4)), int(cfg.get("refinement_range", 5)))
        st["residual"], st["log_softmax"] = res, lsm
        st["flow"] = st["flow_reg"] + res  # ufm.py:991
    return st


# ----------------------------------------------------------------------------------------------
# API shell: pre/post (base.py:137-334, flow_resizing.py)
# ----------------------------------------------------------------------------------------------
def canonicalize_inputs(src: torch.Tensor, tgt: torch.Tensor, required_norm: str, data_norm_type: Optional[str]):
    """``predict_correspondences_batched`` front half (uniflowmatch/models/base.py:160-231) -> 2x (B,3,H,W) fp32."""
    assert isinstance(src, torch.Tensor) and isinstance(tgt, torch.Tensor), "source_image and target_image must be torch.Tensors"
    assert src.dim() in [3, 4], "source_image must have dimensions 3 or 4"
    assert tgt.dim() in [3, 4], "target_image must have dimensions 3 or 4"
    if src.dim() == 3:
        src, tgt = src.unsqueeze(0), tgt.unsqueeze(0)
    if src.shape[1] == 3 and tgt.shape[1] == 3:
        pass
    elif src.shape[-1] == 3 and tgt.shape[-1] == 3:
        src, tgt = src.permute(0, 3, 1, 2), tgt.permute(0, 3, 1, 2)
    else:
        raise ValueError("source_image and target_image must have 3 channels in either BCHW or BHWC format")
    mean, std = (torch.tensor(v).view(1, 3, 1, 1) for v in IMAGE_NORMALIZATION[required_norm])
    if src.dtype == torch.float32:
        assert data_norm_type is not None, "data_norm_type must be provided for float32 images"
        assert data_norm_type in IMAGE_NORMALIZATION, f"data_norm_type must be one of {list(IMAGE_NORMALIZATION)}"
        if data_norm_type != required_norm:
            pm, ps = (torch.tensor(v).view(1, 3, 1, 1) for v in IMAGE_NORMALIZATION[data_norm_type])
            src = src * (ps / std) + (pm - mean) / std
            tgt = tgt * (ps / std) + (pm - mean) / std
    elif src.dtype == torch.uint8:
        src = (src.float() / 255.0 - mean) / std
        tgt = (tgt.float() / 255.0 - mean) / std
    else:
        raise ValueError("source_image and target_image must be of type torch.float32 or torch.uint8")
    return src, tgt


def select_resolution(resolutions, H0, W0, H1, W1):
    """``AutomaticShapeSelection._closest_aspect_strategy`` (flow_resizing.py:667-694); resolutions are (W, H)."""
    cands = [(r[1], r[0]) for r in resolutions]
    if not cands:
        raise ValueError("No valid shape found for the given resolution.")
    return min(cands, key=lambda s: abs(s[0] / s[1] - H0 / W0) + abs(s[0] / s[1] - H1 / W1))


def unmap_flow(flow, rep0, src0, src1, shape0):
    """``unmap_predicted_flow`` (flow_resizing.py:749-877).  Regions are [top, bottom, left, right]."""
    B = flow.shape[0]
    roi = flow[..., rep0[0]:rep0[1], rep0[2]:rep0[3]]
    sc = torch.stack(torch.meshgrid(torch.arange(0, roi.shape[3]) + 0.5, torch.arange(0, roi.shape[2]) + 0.5,
                                    indexing="xy"), dim=-1).permute(2, 0, 1).unsqueeze(0)
    svs = [src0[1] - src0[0], src0[3] - src0[2]]
    tvs = [src1[1] - src1[0], src1[3] - src1[2]]
    scv = F.interpolate(sc.float(), size=svs, mode="bilinear", align_corners=False)
    tcv = F.interpolate(roi.float(), size=svs, mode="nearest") + scv
    scv = scv * torch.tensor([svs[1] / (rep0[3] - rep0[2]), svs[0] / (rep0[1] - rep0[0])]).view(1, 2, 1, 1)
    tcv = tcv * torch.tensor([tvs[1] / (rep0[3] - rep0[2]), tvs[0] / (rep0[1] - rep0[0])]).view(1, 2, 1, 1)
    scv = scv + torch.tensor([src0[2], src0[0]]).view(1, 2, 1, 1)
    tcv = tcv + torch.tensor([src1[2], src1[0]]).view(1, 2, 1, 1)
    out = torch.zeros(B, 2, shape0[0], shape0[1])
    out[..., src0[0]:src0[1], src0[2]:src0[3]] = tcv - scv
    return out


def unmap_channels(ch, rep0, src0, shape0):
    """``unmap_predicted_channels`` (flow_resizing.py:955-1010): crop, nearest resize, paste."""
    roi = ch[..., rep0[0]:rep0[1], rep0[2]:rep0[3]]
    r = F.interpolate(roi, size=[src0[1] - src0[0], src0[3] - src0[2]], mode="nearest")
    out = torch.zeros(ch.shape[0], ch.shape[1], shape0[0], shape0[1], dtype=ch.dtype)
    out[..., src0[0]:src0[1], src0[2]:src0[3]] = r
    return out


def predict_correspondences_batched(src, tgt, sd, cfg, data_norm_type: Optional[str] = None) -> Dict[str, torch.Tensor]:
    """``predict_correspondences_batched`` + ``_predict_correspondences_batched`` (base.py:137-334) on CPU fp32.
    Returns {"flow": (B,2,H0,W0), "covisibility": (B,H0,W0) [if present], "flow_covariance" [if present]}."""
    src, tgt = canonicalize_inputs(src, tgt, cfg["encoder_kwargs"]["data_norm_type"], data_norm_type)
    H0, W0 = src.shape[2:]
    H1, W1 = tgt.shape[2:]
    res = cfg.get("inference_resolution") or [(560, 420)]
    if isinstance(res[0], int):
        res = [res]
    th, tw = select_resolution(res, H0, W0, H1, W1)
    # ResizeToFixedManipulation.__call__ (flow_resizing.py:276-354)
    s0 = F.interpolate(src, size=(th, tw), mode="bilinear", align_corners=False, antialias=True)
    s1 = F.interpolate(tgt, size=(th, tw), mode="bilinear", align_corners=False, antialias=True)
    rep0 = (torch.tensor([th / H0, th / H0, tw / W0, tw / W0]) * torch.tensor([0, H0, 0, W0])).to(torch.int64).tolist()
    src0, src1 = [0, H0, 0, W0], [0, H1, 0, W1]
    st = forward_stages(s0, s1, sd, cfg)
    out = {"flow": unmap_flow(st["flow"], rep0, src0, src1, (H0, W0))}
    if "flow_cov" in st:
        cov = unmap_channels(st["flow_cov"], rep0, src0, (H0, W0))
        wr, hr = W0 / tw, H0 / th
        out["flow_covariance"] = cov * torch.tensor([wr ** 2, hr ** 2, wr * hr]).view(1, 3, 1, 1)
    if "covis" in st:
        out["covisibility"] = unmap_channels(st["covis"], rep0, src0, (H0, W0)).squeeze(1)
    return out
