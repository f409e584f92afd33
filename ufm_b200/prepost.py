"""Pre/post-processing shell of ``predict_correspondences_batched`` for inputs that are NOT at the inference
resolution (uniflowmatch/utils/flow_resizing.py:276-354, 667-744, 749-877, 955-1010).  At the native resolution the
engine consumes the raw images directly and none of this runs.

The product path is ``prep_resize`` / ``unmap`` below: two CUDA kernels (csrc/prepost.cu) behind the C ABI
(``ufm_op_prep_resize``, ``ufm_op_unmap``).  The torch restatements of the same reference code that the GPU tests check
the kernels against live with the rest of the test infrastructure in ``oracle/ufm_oracle.py`` (``unmap_flow``,
``unmap_channels``), not in this package.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import ctypes as C

import numpy as np
import torch

from . import _lib

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


def scaled_region(th, tw, H0, W0) -> List[int]:
    mult = torch.tensor([th / H0, th / H0, tw / W0, tw / W0])
    return (mult * torch.tensor([0, H0, 0, W0])).to(torch.int64).tolist()


# ---------------------------------------------------------------------------------------------------------------------
# kernel-backed product path
# ---------------------------------------------------------------------------------------------------------------------
def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


def prep_resize(img: torch.Tensor, nhwc: bool, th: int, tw: int, required: str, given: str = None) -> torch.Tensor:
    """Normalise (base.py:215-229) and anti-alias resize (flow_resizing.py:276-354) in one kernel.

    img: CUDA uint8 or fp32, [B,H0,W0,3] (nhwc) or [B,3,H0,W0]; fp32 images are already normalised with ``given`` and are
    re-normalised to ``required`` when the two differ.  Returns normalised fp32 [B,3,th,tw]."""
    assert img.is_cuda and img.dtype in (torch.uint8, torch.float32)
    img = img.contiguous()
    B = img.shape[0]
    H0, W0 = (img.shape[1], img.shape[2]) if nhwc else (img.shape[2], img.shape[3])
    mean, std = IMAGE_NORMALIZATION[required]
    mul, add = (1.0, 1.0, 1.0), (0.0, 0.0, 0.0)
    if img.dtype == torch.float32 and given is not None and given != required:
        pm, ps = IMAGE_NORMALIZATION[given]
        # same fp32 arithmetic as the torch expression  x * (ps / std) + (pm - mean) / std
        mul = [float(np.float32(a) / np.float32(b)) for a, b in zip(ps, std)]
        add = [float((np.float32(a) - np.float32(m)) / np.float32(b)) for a, m, b in zip(pm, mean, std)]
    out = torch.empty(B, 3, th, tw, dtype=torch.float32, device=img.device)
    lib = _lib.load()
    with torch.cuda.device(img.device):
        rc = lib.ufm_op_prep_resize(img.data_ptr(), int(img.dtype == torch.uint8), int(nhwc), B, H0, W0, th, tw, _f3(mean),
                                    _f3(std), _f3(mul), _f3(add), out.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream)
    if rc != 0:
        raise RuntimeError(f"ufm_op_prep_resize failed for {H0}x{W0} -> {th}x{tw} (down-scale factor too large?)")
    return out


def unmap(flow, cov, rep0, src0, src1, shape0, cov_mul=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """``unmap_predicted_flow`` + ``unmap_predicted_channels`` (flow_resizing.py:749-877, 955-1010) in one kernel.

    flow [B,2,H,W] or None, cov [B,H,W] or None (fp32, CUDA, inference resolution) -> flow [B,2,H0,W0], cov [B,H0,W0];
    plane b of cov is multiplied by cov_mul[b % len(cov_mul)] (the covariance rescaling of base.py:311-315)."""
    ref = flow if flow is not None else cov
    if flow is not None:
        flow = flow.contiguous()
    B, H, W = ref.shape[0], ref.shape[-2], ref.shape[-1]
    H0, W0 = int(shape0[0]), int(shape0[1])
    svs = [src0[1] - src0[0], src0[3] - src0[2]]
    tvs = [src1[1] - src1[0], src1[3] - src1[2]]
    rh, rw = rep0[1] - rep0[0], rep0[3] - rep0[2]
    i4 = lambda v: (C.c_int * 4)(*[int(x) for x in v])
    f2 = lambda a, b: (C.c_float * 2)(float(a), float(b))
    out = torch.empty(B, 2, H0, W0, dtype=torch.float32, device=ref.device) if flow is not None else None
    cout = torch.empty(B, H0, W0, dtype=torch.float32, device=ref.device) if cov is not None else None
    cm = (C.c_float * len(cov_mul))(*[float(np.float32(v)) for v in cov_mul]) if cov_mul else None
    if cov is not None:
        cov = cov.contiguous()
    lib = _lib.load()
    with torch.cuda.device(ref.device):
        rc = lib.ufm_op_unmap(None if flow is None else flow.data_ptr(), None if cov is None else cov.data_ptr(), B, H, W, H0,
                              W0, i4(rep0), i4(src0), i4(src1), f2(svs[1] / rw, svs[0] / rh), f2(tvs[1] / rw, tvs[0] / rh),
                              None if out is None else out.data_ptr(), None if cout is None else cout.data_ptr(), cm,
                              len(cov_mul) if cov_mul else 0, torch.cuda.current_stream().cuda_stream)
    if rc != 0:
        raise RuntimeError("ufm_op_unmap: invalid regions")
    return out, cout
