"""Run the UNMODIFIED reference classes from /root/reference (build container) or from the staged byte-for-byte copy
of its import closure under ``oracle/_ref`` (GPU box; made by ``oracle/build_ref.py``).

Test infrastructure: used by ``oracle/make_golden.py``, ``bench.py``'s reference legs and by tests that are skipped
when neither location exists.  The single substitution is
``torch.hub.load`` (reference ``UniCeption/uniception/models/encoders/dinov2.py:85-91``), which has
no network here; it is replaced by :class:`HubDinoVisionTransformer`, a restatement of the published
``facebookresearch/dinov2`` ``DinoVisionTransformer`` (``dinov2/models/vision_transformer.py``,
``dinov2/layers/{block,attention,mlp,layer_scale,patch_embed}.py``; unpinned upstream — SURVEY.md
Appendix B).  The reference wrapper then runs as shipped: it deletes ``mask_token``
(``dinov2.py:93-95``), truncates ``blocks`` (``:98-99``), swaps the attention forward for SDPA
(``:116-144``) and calls ``get_intermediate_layers`` (``:259-261``).
"""
from __future__ import annotations

import math
import os
import sys
from typing import Any, Dict

import torch
import torch.nn as nn
import torch.nn.functional as F

_STAGED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _find_root() -> str:
    """$UFM_REFERENCE_ROOT, else /root/reference (build container), else the byte-for-byte import closure that
    ``oracle/build_ref.py`` staged under ``oracle/_ref`` (git-ignored; it travels to the GPU box with gpurun)."""
    env = os.environ.get("UFM_REFERENCE_ROOT")
    if env:
        return env
    for cand in ("/root/reference", _STAGED):
        if os.path.isdir(os.path.join(cand, "uniflowmatch")):
            return cand
    return "/root/reference"


REFERENCE_ROOT = _find_root()


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "uniflowmatch"))


class _Attn(nn.Module):
    def __init__(self, dim, heads):
        super().__init__()
        self.num_heads = heads
        self.qkv = nn.Linear(dim, 3 * dim)
        self.attn_drop = nn.Dropout(0.0)
        self.proj = nn.Linear(dim, dim)
        self.proj_drop = nn.Dropout(0.0)

    def forward(self, x, attn_bias=None):  # replaced by the reference's SDPA wrapper
        B, N, C = x.shape
        qkv = self.qkv(x).reshape(B, N, 3, self.num_heads, C // self.num_heads).permute(2, 0, 3, 1, 4)
        a = (qkv[0] * (C // self.num_heads) ** -0.5) @ qkv[1].transpose(-2, -1)
        x = (a.softmax(-1) @ qkv[2]).transpose(1, 2).reshape(B, N, C)
        return self.proj_drop(self.proj(x))


class _LayerScale(nn.Module):
    def __init__(self, dim, init_values=1.0):
        super().__init__()
        self.gamma = nn.Parameter(init_values * torch.ones(dim))

    def forward(self, x):
        return x * self.gamma


class _Mlp(nn.Module):
    def __init__(self, dim, hidden):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.act = nn.GELU()
        self.fc2 = nn.Linear(hidden, dim)

    def forward(self, x):
        return self.fc2(self.act(self.fc1(x)))


class _Block(nn.Module):
    def __init__(self, dim, heads):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, eps=1e-6)
        self.attn = _Attn(dim, heads)
        self.ls1 = _LayerScale(dim)
        self.norm2 = nn.LayerNorm(dim, eps=1e-6)
        self.mlp = _Mlp(dim, 4 * dim)
        self.ls2 = _LayerScale(dim)

    def forward(self, x):
        x = x + self.ls1(self.attn(self.norm1(x)))
        x = x + self.ls2(self.mlp(self.norm2(x)))
        return x


class _PatchEmbed(nn.Module):
    def __init__(self, dim, patch):
        super().__init__()
        self.proj = nn.Conv2d(3, dim, kernel_size=patch, stride=patch)

    def forward(self, x):
        return self.proj(x).flatten(2).transpose(1, 2)


class HubDinoVisionTransformer(nn.Module):
    """``dinov2_vit{s,b,l}14``: img_size 518, patch 14, LayerScale init 1.0, interpolate_offset 0.1,
    interpolate_antialias False, no register tokens."""

    def __init__(self, dim, depth, heads, patch=14, img=518, num_register_tokens=0):
        super().__init__()
        self.patch_size = patch
        self.embed_dim = dim
        # the *_reg hub entries: 4 register tokens, interpolate_antialias=True, interpolate_offset=0.0 (hub/backbones.py)
        self.num_register_tokens = num_register_tokens
        self.interpolate_offset = 0.0 if num_register_tokens else 0.1
        self.interpolate_antialias = bool(num_register_tokens)
        if num_register_tokens:
            self.register_tokens = nn.Parameter(torch.zeros(1, num_register_tokens, dim))
        self.patch_embed = _PatchEmbed(dim, patch)
        n = (img // patch) ** 2
        self.cls_token = nn.Parameter(torch.zeros(1, 1, dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, n + 1, dim))
        self.mask_token = nn.Parameter(torch.zeros(1, dim))
        self.blocks = nn.ModuleList([_Block(dim, heads) for _ in range(depth)])
        self.norm = nn.LayerNorm(dim, eps=1e-6)

    def interpolate_pos_encoding(self, x, w, h):
        previous_dtype = x.dtype
        npatch = x.shape[1] - 1
        N = self.pos_embed.shape[1] - 1
        if npatch == N and w == h:
            return self.pos_embed
        pos_embed = self.pos_embed.float()
        class_pos_embed = pos_embed[:, 0]
        patch_pos_embed = pos_embed[:, 1:]
        dim = x.shape[-1]
        w0 = w // self.patch_size
        h0 = h // self.patch_size
        M = int(math.sqrt(N))
        sx = float(w0 + self.interpolate_offset) / M
        sy = float(h0 + self.interpolate_offset) / M
        patch_pos_embed = F.interpolate(
            patch_pos_embed.reshape(1, M, M, dim).permute(0, 3, 1, 2),
            mode="bicubic", antialias=self.interpolate_antialias, scale_factor=(sx, sy))
        assert (w0, h0) == patch_pos_embed.shape[-2:]
        patch_pos_embed = patch_pos_embed.permute(0, 2, 3, 1).view(1, -1, dim)
        return torch.cat((class_pos_embed.unsqueeze(0), patch_pos_embed), dim=1).to(previous_dtype)

    def prepare_tokens_with_masks(self, x, masks=None):
        B, nc, w, h = x.shape
        x = self.patch_embed(x)
        x = torch.cat((self.cls_token.expand(x.shape[0], -1, -1), x), dim=1)
        x = x + self.interpolate_pos_encoding(x, w, h)
        if self.num_register_tokens:  # registers sit between cls and the patch tokens and carry no positional embedding
            x = torch.cat((x[:, :1], self.register_tokens.expand(x.shape[0], -1, -1), x[:, 1:]), dim=1)
        return x

    def forward_features(self, x, masks=None):
        x = self.prepare_tokens_with_masks(x)
        for blk in self.blocks:
            x = blk(x)
        x_norm = self.norm(x)
        R = self.num_register_tokens
        return {"x_norm_clstoken": x_norm[:, 0], "x_norm_regtokens": x_norm[:, 1:R + 1], "x_norm_patchtokens": x_norm[:, R + 1:],
                "x_prenorm": x}

    def get_intermediate_layers(self, x, n=1, reshape=False, return_class_token=False, norm=True):
        xin = x
        x = self.prepare_tokens_with_masks(x)
        total = len(self.blocks)
        take = range(total - n, total) if isinstance(n, int) else n
        outs = []
        for i, blk in enumerate(self.blocks):
            x = blk(x)
            if i in take:
                outs.append(x)
        assert len(outs) == len(take)
        if norm:
            outs = [self.norm(o) for o in outs]
        outs = [o[:, 1 + self.num_register_tokens:] for o in outs]
        if reshape:
            B, _, w, h = xin.shape
            outs = [o.reshape(B, w // self.patch_size, h // self.patch_size, -1).permute(0, 3, 1, 2).contiguous()
                    for o in outs]
        return tuple(outs)


_HUB = {"dinov2_vits14": (384, 12, 6), "dinov2_vitb14": (768, 12, 12), "dinov2_vitl14": (1024, 24, 16)}


def _hub_load(repo, name, **kw):
    assert repo == "facebookresearch/dinov2", repo
    if name.endswith("_reg"):
        return HubDinoVisionTransformer(*_HUB[name[:-4]], num_register_tokens=4)
    return HubDinoVisionTransformer(*_HUB[name])


def import_reference():
    """Put the reference on ``sys.path`` and substitute ``torch.hub.load``; return its ``ufm`` module."""
    assert reference_available(), f"{REFERENCE_ROOT} is not mounted"
    for p in (os.path.join(REFERENCE_ROOT, "UniCeption"), REFERENCE_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    torch.hub.load = _hub_load
    import uniflowmatch.models.ufm as ufm  # noqa: E402

    return ufm


def build_reference(cfg: Dict[str, Any], state_dict=None):
    """Construct the reference model for ``cfg`` (Refine if it has a classification head)."""
    import contextlib
    import io

    ufm = import_reference()
    cls = ufm.UniFlowMatchClassificationRefinement if "classification_head_kwargs" in cfg else ufm.UniFlowMatchConfidence
    with contextlib.redirect_stdout(io.StringIO()):
        model = cls(**cfg).eval()
    if state_dict is not None:
        model.load_state_dict(state_dict, strict=True)
    return model
