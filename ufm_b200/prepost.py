"""Pre/post-processing shell of ``predict_correspondences_batched`` for inputs that are NOT at the inference
resolution (uniflowmatch/utils/flow_resizing.py:276-354, 667-744, 749-877, 955-1010).  Host-side torch ops on the
device the data lives on; at the native resolution the engine consumes the raw images directly and none of this runs.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch
import torch.nn.functional as F

# UC/models/encoders/image_normalizations.py:17-35
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


def norm_tensors(name: str, device):
    mean, std = IMAGE_NORMALIZATION[name]
    return (torch.tensor(mean, device=device).view(1, 3, 1, 1), torch.tensor(std, device=device).view(1, 3, 1, 1))


def closest_aspect(resolutions: Sequence[Tuple[int, int]], H0, W0, H1, W1) -> Tuple[int, int]:
    """``AutomaticShapeSelection._closest_aspect_strategy``; ``resolutions`` are (W, H); returns (H, W)."""
    cands = [(r[1], r[0]) for r in resolutions]
    if not cands:
        raise ValueError("No valid shape found for the given resolution.")
    return min(cands, key=lambda s: abs(s[0] / s[1] - H0 / W0) + abs(s[0] / s[1] - H1 / W1))


def resize_to_fixed(img: torch.Tensor, th: int, tw: int) -> torch.Tensor:
    """``ResizeToFixedManipulation.__call__`` image part: bilinear, align_corners=False, antialias=True (BCHW fp32)."""
    return F.interpolate(img, size=(th, tw), mode="bilinear", align_corners=False, antialias=True)


def scaled_region(th, tw, H0, W0) -> List[int]:
    mult = torch.tensor([th / H0, th / H0, tw / W0, tw / W0])
    return (mult * torch.tensor([0, H0, 0, W0])).to(torch.int64).tolist()


def unmap_flow(flow, rep0, src0, src1, shape0):
    """``unmap_predicted_flow``: crop, nearest-resample the flow onto the source grid, rescale end points."""
    dev = flow.device
    B = flow.shape[0]
    roi = flow[..., rep0[0]:rep0[1], rep0[2]:rep0[3]]
    sc = torch.stack(torch.meshgrid(torch.arange(0, roi.shape[3]) + 0.5, torch.arange(0, roi.shape[2]) + 0.5,
                                    indexing="xy"), dim=-1).permute(2, 0, 1).unsqueeze(0).to(dev)
    svs = [src0[1] - src0[0], src0[3] - src0[2]]
    tvs = [src1[1] - src1[0], src1[3] - src1[2]]
    scv = F.interpolate(sc.float(), size=svs, mode="bilinear", align_corners=False)
    tcv = F.interpolate(roi.float(), size=svs, mode="nearest") + scv
    scv = scv * torch.tensor([svs[1] / (rep0[3] - rep0[2]), svs[0] / (rep0[1] - rep0[0])]).view(1, 2, 1, 1).to(dev)
    tcv = tcv * torch.tensor([tvs[1] / (rep0[3] - rep0[2]), tvs[0] / (rep0[1] - rep0[0])]).view(1, 2, 1, 1).to(dev)
    scv = scv + torch.tensor([src0[2], src0[0]]).view(1, 2, 1, 1).to(dev)
    tcv = tcv + torch.tensor([src1[2], src1[0]]).view(1, 2, 1, 1).to(dev)
    out = torch.zeros(B, 2, shape0[0], shape0[1], dtype=flow.dtype, device=dev)
    out[..., src0[0]:src0[1], src0[2]:src0[3]] = tcv - scv
    return out


def unmap_channels(ch, rep0, src0, shape0):
    """``unmap_predicted_channels``: crop -> nearest resize -> paste."""
    roi = ch[..., rep0[0]:rep0[1], rep0[2]:rep0[3]]
    r = F.interpolate(roi, size=[src0[1] - src0[0], src0[3] - src0[2]], mode="nearest")
    out = torch.zeros(ch.shape[0], ch.shape[1], shape0[0], shape0[1], dtype=ch.dtype, device=ch.device)
    out[..., src0[0]:src0[1], src0[2]:src0[3]] = r
    return out
