"""Drop-in model classes: same names, constructor kwargs, ``state_dict`` keys and
``predict_correspondences_batched`` contract as the reference (uniflowmatch/models/base.py:85-334,
uniflowmatch/models/ufm.py:120-1009), with the forward pass executed by the sm_100a engine.
"""
from __future__ import annotations

import json
import os

from dataclasses import dataclass
from typing import Any, Dict, Optional

import torch
import torch.nn as nn

from . import _lib, prepost
from .engine import Engine
from .packing import inference_hw
from .weights import make_state_dict, model_spec


# ---- output dataclasses (base.py:11-73) ------------------------------------------------------------------------------
@dataclass
class UFMFlowFieldOutput:
    flow_output: torch.Tensor
    flow_covariance: Optional[torch.Tensor] = None
    flow_covariance_inv: Optional[torch.Tensor] = None
    flow_covariance_log_det: Optional[torch.Tensor] = None


@dataclass
class UFMMaskFieldOutput:
    mask: torch.Tensor
    logits: Optional[torch.Tensor]


@dataclass
class UFMClassificationRefinementOutput:
    regression_flow_output: torch.Tensor
    residual: torch.Tensor
    log_softmax: torch.Tensor
    feature_map_0: torch.Tensor
    feature_map_1: torch.Tensor


@dataclass
class UFMOutputInterface:
    flow: Optional[UFMFlowFieldOutput] = None
    classification_refinement: Optional[UFMClassificationRefinementOutput] = None
    covisibility: Optional[UFMMaskFieldOutput] = None


def modify_state_dict(original_state_dict, mappings):
    """Rename / drop checkpoint keys: ``{substring: replacement or None}``, first matching rule wins, ``None`` drops the
    key (semantics of ``modify_state_dict``, uniflowmatch/models/ufm.py:85-117)."""
    out = {}
    for k, v in original_state_dict.items():
        new_key, skip = None, False
        for old, new in mappings.items():
            if old in k:
                if new is None:
                    skip = True
                else:
                    new_key = k.replace(old, new)
                break
        if not skip:
            out[new_key if new_key is not None else k] = v
    return out


def _aliases(key: str):
    """Other reference keys that name the same DPT tensor (dpt.py:107-172: ``scratch.layer{K}_rn``, ``scratch.layer_rn[K-1]``
    and ``input_process[K-1][1]`` are one module)."""
    import re

    m = re.match(r"(.*\.)scratch\.layer(\d)_rn\.(.*)", key)
    if m:
        k = int(m.group(2)) - 1
        return [f"{m.group(1)}scratch.layer_rn.{k}.{m.group(3)}", f"{m.group(1)}input_process.{k}.1.{m.group(3)}"]
    m = re.match(r"(.*\.)scratch\.layer_rn\.(\d)\.(.*)", key)
    if m:
        k = int(m.group(2))
        return [f"{m.group(1)}scratch.layer{k + 1}_rn.{m.group(3)}", f"{m.group(1)}input_process.{k}.1.{m.group(3)}"]
    m = re.match(r"(.*\.)input_process\.(\d)\.1\.(.*)", key)
    if m:
        k = int(m.group(2))
        return [f"{m.group(1)}scratch.layer{k + 1}_rn.{m.group(3)}", f"{m.group(1)}scratch.layer_rn.{k}.{m.group(3)}"]
    return []


class _EncoderShim:
    def __init__(self, data_norm_type):
        self.data_norm_type = data_norm_type


class UniFlowMatchModelsBase(nn.Module):
    """Holds the reference-named tensors as buffers (inference only) and owns one engine per CUDA device."""

    _refine = False

    def __init__(self, max_pairs_per_call: int = 16, seed: int = 0, **cfg):
        super().__init__()
        assert cfg.pop("pretrained_checkpoint_path", None) is None, "Pretrained weights are not supported for now"  # ufm.py:558
        backbone_ckpt = cfg.pop("pretrained_backbone_checkpoint_path", None)
        if cfg.get("info_sharing_and_head_structure", "dual+single") != "dual+single":
            raise AssertionError("Only dual+single is supported now")  # ufm.py:180
        if self._refine and "classification_head_kwargs" not in cfg:
            raise ValueError("UniFlowMatchClassificationRefinement needs classification_head_kwargs")
        if not self._refine:
            for k in ("classification_head_kwargs", "temperature", "refinement_range", "use_unet_feature",
                      "classification_head_type", "feature_combine_method"):
                cfg.pop(k, None)
        self.config = cfg
        self.max_pairs_per_call = max_pairs_per_call
        self.inference_resolution = cfg.get("inference_resolution") or [560, 420]
        self.encoder = _EncoderShim(cfg["encoder_kwargs"].get("data_norm_type", "dinov2"))
        self._names = {}
        for key, t in make_state_dict(cfg, seed=seed).items():
            safe = key.replace(".", "__")
            self._names[safe] = key
            self.register_buffer(safe, t)
        self._engines: Dict[Any, Engine] = {}
        self._last_unc: Dict[str, torch.Tensor] = {}
        if backbone_ckpt is not None:
            self._load_backbone_checkpoint(backbone_ckpt)

    # checkpoint ingestion ---------------------------------------------------------------------------
    def _load_backbone_checkpoint(self, path: str):
        """``UniFlowMatch.__init__`` (ufm.py:197-217): a training checkpoint ({"state_dict": {"model.<key>": ...}}, with the
        feature-matching projection and the DINOv2 mask token dropped, strict) or an exported one ({"model": ...}, no
        missing keys allowed).  Only the backbone keys (encoder / info_sharing / head1) are taken."""
        ckpt = torch.load(path, map_location="cpu", weights_only=False)
        mine = set(self._names.values())
        backbone = {k for k in mine if k.split(".")[0] in ("encoder", "info_sharing", "head1")}
        if "state_dict" in ckpt:
            sd = {k[6:]: v for k, v in ckpt["state_dict"].items() if k.startswith("model.")}
            sd = modify_state_dict(sd, {"feature_matching_proj": None, "encoder.model.mask_token": None})
            missing, unexpected = sorted(backbone - set(sd)), sorted(set(sd) - backbone)
            if missing or unexpected:
                raise RuntimeError(f"backbone checkpoint: missing {missing[:5]}, unexpected {unexpected[:5]}")
        else:
            sd = ckpt["model"]
            missing = sorted(backbone - set(sd))
            assert len(missing) == 0, f"Missing keys: {missing}"
        self.load_state_dict({**self.state_dict(), **{k: v for k, v in sd.items() if k in backbone}})

    @classmethod
    def from_pretrained_ckpt(cls, pretrained_model_name_or_path, strict=True, **kw):
        """``UniFlowMatch.from_pretrained_ckpt`` (ufm.py:219-240): {"model_args": ctor kwargs, "model": state_dict}."""
        if os.path.isfile(pretrained_model_name_or_path):
            ckpt = torch.load(pretrained_model_name_or_path, map_location="cpu", weights_only=False)
            model = cls(**ckpt["model_args"], **kw)
            model.load_state_dict(ckpt["model"], strict=strict)
            return model
        raise ValueError(f"Pretrained model {pretrained_model_name_or_path} not found.")

    @classmethod
    def from_pretrained(cls, pretrained_model_name_or_path, strict: bool = True, **kw):
        """``PyTorchModelHubMixin.from_pretrained`` for the layout the reference models are published in
        ("infinity1096/UFM-Base", ufm.py:1254-1256): a directory holding ``config.json`` (constructor kwargs) and
        ``model.safetensors`` (reference key names).  A hub id is resolved through the local Hugging Face cache only:
        this build has no network access."""
        path = str(pretrained_model_name_or_path)
        if not os.path.isdir(path):
            try:
                from huggingface_hub import snapshot_download

                path = snapshot_download(path, local_files_only=True)
            except Exception as e:
                raise ValueError(f"{pretrained_model_name_or_path!r} is neither a directory with config.json + "
                                 f"model.safetensors nor a model in the local Hugging Face cache ({e})")
        with open(os.path.join(path, "config.json")) as f:
            cfg = json.load(f)
        from safetensors.torch import load_file

        sd = load_file(os.path.join(path, "model.safetensors"))
        # the published checkpoints may carry the DINOv2 mask token the wrapper deletes (dinov2.py:93-95)
        sd = modify_state_dict(sd, {"encoder.model.mask_token": None})
        model = cls(**{**cfg, **kw})
        # safetensors stores one copy of tensors that alias each other (scratch.layerK_rn == scratch.layer_rn.K-1 ==
        # input_process.K-1.1): restore the aliases before the strict check
        for key in model._names.values():
            if key not in sd:
                for other in _aliases(key):
                    if other in sd:
                        sd[key] = sd[other]
                        break
        model.load_state_dict(sd, strict=strict)
        return model

    def save_pretrained(self, save_directory: str):
        """config.json + model.safetensors in the layout ``from_pretrained`` (and the reference's hub mixin) reads."""
        os.makedirs(save_directory, exist_ok=True)
        from safetensors.torch import save_file

        with open(os.path.join(save_directory, "config.json"), "w") as f:
            json.dump(self.config, f, indent=1)
        save_file({k: v.detach().cpu().contiguous().clone() for k, v in self.state_dict().items()},
                  os.path.join(save_directory, "model.safetensors"))

    # reference key names in/out -------------------------------------------------------------------------------------
    def state_dict(self, *a, **k):
        return {self._names[n]: t for n, t in self.named_buffers(remove_duplicate=False) if n in self._names}

    def load_state_dict(self, sd, strict: bool = True):
        want = set(self._names.values())
        missing = sorted(want - set(sd))
        unexpected = sorted(set(sd) - want)
        if strict and (missing or unexpected):
            raise RuntimeError(f"load_state_dict: missing {missing[:5]}, unexpected {unexpected[:5]}")
        for safe, key in self._names.items():
            if key in sd:
                getattr(self, safe).copy_(sd[key])
        for e in self._engines.values():
            e.close()
        self._engines = {}
        return torch.nn.modules.module._IncompatibleKeys(missing, unexpected)

    def _resolutions(self):
        """Trained resolutions as (H, W) (base.py:86-98 accepts one (W, H) or a list of them)."""
        res = self.inference_resolution
        res = [res] if isinstance(res[0], int) else res
        return [(int(r[1]), int(r[0])) for r in res]

    def _engine(self, device, hw=None) -> Engine:
        """One engine (weight replica + workspace) per device and trained resolution, created on first use."""
        device = torch.device(device)
        hw = tuple(hw) if hw is not None else self._resolutions()[0]
        assert hw in self._resolutions(), f"{hw} is not one of the model's inference resolutions {self._resolutions()}"
        key = (device.type, device.index if device.index is not None else torch.cuda.current_device(), hw)
        if key not in self._engines:
            cfg = dict(self.config, inference_resolution=[hw[1], hw[0]])
            self._engines[key] = Engine(cfg, self.state_dict(), device, self.max_pairs_per_call)
        return self._engines[key]

    def _device(self):
        return next(self.buffers()).device

    # forward (ufm.py:562 / 843) ---------------------------------------------------------------------------------------
    def _run(self, img0: torch.Tensor, img1: torch.Tensor, layout: int, keep_aux: bool = False):
        """img0/img1 at the inference resolution, uint8 or normalised fp32, on CUDA."""
        if img0.shape != img1.shape:
            raise NotImplementedError("Unequal Image sizes are not supported now")  # ufm.py:316-317
        hw = tuple(img0.shape[1:3] if layout == _lib.UFM_NHWC else img0.shape[2:4])
        eng = self._engine(img0.device, hw)
        eng.set_keep_aux(keep_aux)
        eng.draw_view_index()
        flows, covs, aux, unc = [], [], [], []
        B = img0.shape[0]
        for s in range(0, B, self.max_pairs_per_call):
            a = img0[s:s + self.max_pairs_per_call].contiguous()
            b = img1[s:s + self.max_pairs_per_call].contiguous()
            f, c = eng.forward(a, b, layout)
            flows.append(f)
            covs.append(c)
            if keep_aux and self._refine:
                aux.append({k: eng.stage(k) for k in ("flow_reg", "residual", "log_softmax", "desc")})
            # uncertainty-head adaptors beyond the mask (ufm.py:644-660) are read back from their stage buffers
            names = (["flow_cov", "flow_cov_inv", "flow_cov_log_det"] if eng.has_cov else []) + \
                    (["keypoint_confidence"] if eng.has_conf else [])
            if names:
                unc.append({k: eng.stage(k) for k in names})
        flow = flows[0] if len(flows) == 1 else torch.cat(flows)
        cov = None if covs[0] is None else (covs[0] if len(covs) == 1 else torch.cat(covs))
        self._last_unc = {k: torch.cat([u[k] for u in unc]) for k in unc[0]} if unc else {}
        return flow, cov, aux

    def _run_host(self, img0: torch.Tensor, img1: torch.Tensor, layout: int):
        """Host uint8 / normalised fp32 images at the inference resolution -> pinned host (flow, covisibility)."""
        if img0.shape != img1.shape:
            raise NotImplementedError("Unequal Image sizes are not supported now")  # ufm.py:316-317
        hw = tuple(img0.shape[1:3] if layout == _lib.UFM_NHWC else img0.shape[2:4])
        eng = self._engine(self._device(), hw)
        eng.set_keep_aux(False)
        eng.draw_view_index()
        B = img0.shape[0]
        flow = torch.empty(B, 2, eng.H, eng.W, dtype=torch.float32, pin_memory=True)
        cov = torch.empty(B, eng.H, eng.W, dtype=torch.float32, pin_memory=True) if eng.has_mask else None
        unc = []
        for s in range(0, B, self.max_pairs_per_call):
            e = min(B, s + self.max_pairs_per_call)
            eng.forward_host(img0[s:e], img1[s:e], layout, flow[s:e], cov[s:e] if cov is not None else None)
            names = (["flow_cov", "flow_cov_inv", "flow_cov_log_det"] if eng.has_cov else []) + \
                    (["keypoint_confidence"] if eng.has_conf else [])
            if names:
                unc.append({k: eng.stage(k) for k in names})
        self._last_unc = {k: torch.cat([u[k] for u in unc]) for k in unc[0]} if unc else {}
        return flow, cov

    def forward(self, view1, view2) -> UFMOutputInterface:
        img1, img2 = view1["img"], view2["img"]
        assert view1.get("data_norm_type", self.encoder.data_norm_type) == self.encoder.data_norm_type
        assert img1.dim() == 4 and img1.shape[1] == 3, "Input must be of shape (B, C, H, W)"
        flow, cov, aux = self._run(img1.float().contiguous(), img2.float().contiguous(), _lib.UFM_NCHW,
                                   keep_aux=self._refine)
        res = UFMOutputInterface(flow=UFMFlowFieldOutput(flow_output=flow))
        if cov is not None:
            res.covisibility = UFMMaskFieldOutput(mask=cov.unsqueeze(1), logits=None)
        u = self._last_unc
        if "flow_cov" in u:  # ufm.py:649-652
            res.flow.flow_covariance, res.flow.flow_covariance_inv = u["flow_cov"], u["flow_cov_inv"]
            res.flow.flow_covariance_log_det = u["flow_cov_log_det"]
        if "keypoint_confidence" in u:  # ufm.py:654 (set dynamically there as well)
            res.keypoint_confidence = u["keypoint_confidence"]
        if self._refine and aux:
            cat = lambda k: torch.cat([a[k] for a in aux])
            desc = cat("desc")  # (B, 2, C/8, H, W, 8) bf16: the engine's channel-group planar descriptor layout
            fmap = lambda v: desc[:, v].permute(0, 1, 4, 2, 3).reshape(desc.shape[0], -1, desc.shape[3], desc.shape[4]).float()
            P = int(self.config.get("refinement_range", 5))
            res.classification_refinement = UFMClassificationRefinementOutput(
                regression_flow_output=flow, residual=cat("residual"),
                log_softmax=cat("log_softmax").reshape(flow.shape[0], flow.shape[2], flow.shape[3], P, P),
                feature_map_0=fmap(0), feature_map_1=fmap(1))
        return res

    # API shell (base.py:137-334) --------------------------------------------------------------------------------------
    def predict_correspondences_batched(self, source_image: torch.Tensor, target_image: torch.Tensor,
                                        data_norm_type: Optional[str] = None) -> UFMOutputInterface:
        assert isinstance(source_image, torch.Tensor) and isinstance(target_image, torch.Tensor), \
            "source_image and target_image must be torch.Tensors"
        assert source_image.dim() in [3, 4], "source_image must have dimensions 3 or 4"
        assert target_image.dim() in [3, 4], "target_image must have dimensions 3 or 4"
        if source_image.dim() == 3:
            source_image, target_image = source_image.unsqueeze(0), target_image.unsqueeze(0)
        if source_image.shape[1] == 3 and target_image.shape[1] == 3:
            layout = _lib.UFM_NCHW
        elif source_image.shape[-1] == 3 and target_image.shape[-1] == 3:
            layout = _lib.UFM_NHWC
        else:
            raise ValueError("source_image and target_image must have 3 channels in either BCHW or BHWC format")
        required = self.encoder.data_norm_type
        if source_image.dtype == torch.float32:
            assert data_norm_type is not None, "data_norm_type must be provided for float32 images"
            assert data_norm_type in prepost.IMAGE_NORMALIZATION, \
                f"data_norm_type must be one of {list(prepost.IMAGE_NORMALIZATION.keys())}"
        elif source_image.dtype != torch.uint8:
            raise ValueError("source_image and target_image must be of type torch.float32 or torch.uint8")
        if source_image.shape[0] == 0 or target_image.shape[0] == 0:
            # the reference fails on an empty batch too (a reshape of 0 elements with an inferred dimension inside the
            # encoder wrapper); stated here instead of surfacing as an index error further down
            raise RuntimeError("predict_correspondences_batched: empty batch (the reference raises on B = 0 as well)")
        in_device = source_image.device
        dev = self._device()
        if dev.type != "cuda":
            raise RuntimeError("ufm_b200 models run on CUDA only: call .to('cuda') (there is no CPU fallback)")
        hw0 = tuple(source_image.shape[1:3] if layout == _lib.UFM_NHWC else source_image.shape[2:4])
        hw1 = tuple(target_image.shape[1:3] if layout == _lib.UFM_NHWC else target_image.shape[2:4])
        # AutomaticShapeSelection (flow_resizing.py:667-694): the trained resolution of closest aspect, first entry on ties
        res_wh = [(w, h) for h, w in self._resolutions()]
        pick = prepost.closest_aspect(res_wh, hw0[0], hw0[1], hw1[0], hw1[1])
        native = hw0 == hw1 == tuple(pick)  # resize and un-map are then the identity
        renorm = source_image.dtype == torch.float32 and data_norm_type != required
        if native and not renorm and in_device.type == "cpu" and target_image.device.type == "cpu" \
                and source_image.dtype == target_image.dtype:
            # host tensors at the native size: one ufm_forward_host call per chunk -- H2D, forward and the D2H copies
            # (overlapped with the tail of the pass) inside the C ABI, results land in pinned host tensors
            flow, cov = self._run_host(source_image.contiguous(), target_image.contiguous(), layout)
            out = UFMOutputInterface(flow=UFMFlowFieldOutput(flow_output=flow))
            fcov = self._last_unc.get("flow_cov")
            if fcov is not None:
                out.flow.flow_covariance = fcov.to(in_device)
            if cov is not None:
                out.covisibility = UFMMaskFieldOutput(mask=cov, logits=None)
            return out
        src = source_image.to(dev, non_blocking=True)
        tgt = target_image.to(dev, non_blocking=True)
        if native and not renorm:
            # the engine reads the images as they are (u8 normalisation is fused into the patch gather);
            # the reference's un-map at native size is the identity up to (flow + c) - c rounding
            flow, cov, _ = self._run(src.contiguous(), tgt.contiguous(), layout)
            fcov = self._last_unc.get("flow_cov")  # base.py:295-319 at native size: ratios are 1
        else:
            # non-native size or foreign normalisation: normalise + anti-alias resize, forward, un-map -- two CUDA
            # kernels around the engine (csrc/prepost.cu), no torch ops on the data path
            th, tw = pick
            nhwc = layout == _lib.UFM_NHWC
            s0 = prepost.prep_resize(src, nhwc, th, tw, required, data_norm_type)
            s1 = prepost.prep_resize(tgt, nhwc, th, tw, required, data_norm_type)
            f, c, _ = self._run(s0, s1, _lib.UFM_NCHW)
            rep0 = prepost.scaled_region(th, tw, hw0[0], hw0[1])
            src0, src1 = [0, hw0[0], 0, hw0[1]], [0, hw1[0], 0, hw1[1]]
            flow, cov = prepost.unmap(f, c, rep0, src0, src1, hw0)
            fcov = self._last_unc.get("flow_cov")
            if fcov is not None:
                # unmap_predicted_channels + covariance rescaling (base.py:295-319): the three planes as 3B images
                wr, hr = hw0[1] / tw, hw0[0] / th
                _, fc = prepost.unmap(None, fcov.reshape(-1, th, tw), rep0, src0, src1, hw0, cov_mul=(wr ** 2, hr ** 2, wr * hr))
                fcov = fc.reshape(-1, 3, hw0[0], hw0[1])
        out = UFMOutputInterface(flow=UFMFlowFieldOutput(flow_output=flow.to(in_device)))
        if fcov is not None:
            out.flow.flow_covariance = fcov.to(in_device)
        if cov is not None:
            out.covisibility = UFMMaskFieldOutput(mask=cov.to(in_device), logits=None)
        return out


class UniFlowMatchConfidence(UniFlowMatchModelsBase):
    """Engine-backed ``UniFlowMatchConfidence`` (ufm.py:474-708): flow head + uncertainty (covisibility) head."""
    _refine = False


class UniFlowMatchClassificationRefinement(UniFlowMatchModelsBase):
    """Engine-backed ``UniFlowMatchClassificationRefinement`` (ufm.py:710-1178)."""
    _refine = True
