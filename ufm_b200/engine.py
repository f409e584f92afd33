"""Python handle on one ``ufm_engine`` (one per GPU): weights in, device pointers in/out.  PyTorch is used only
for device memory and streams; every kernel on the path is in ``libufm_b200.so``."""
from __future__ import annotations

import ctypes as C
from typing import Any, Dict, Optional, Tuple

import torch

from . import _lib
from .packing import engine_config, pack_weights, view_index_range

_DT = {torch.float32: _lib.UFM_F32, torch.bfloat16: _lib.UFM_BF16, torch.uint8: _lib.UFM_U8}
_TD = {_lib.UFM_F32: torch.float32, _lib.UFM_BF16: torch.bfloat16}


class UfmEngineError(RuntimeError):
    pass


class Engine:
    def __init__(self, cfg: Dict[str, Any], state_dict: Dict[str, torch.Tensor], device: torch.device, max_pairs: int = 16):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise UfmEngineError("ufm_b200 needs a B200 (sm_100) GPU: there is no CPU fallback")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise UfmEngineError(f"engine device must be CUDA, got {self.device}")
        self.index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.cfg = cfg
        self.ec = engine_config(cfg, state_dict)
        self.H, self.W = self.ec.img_h, self.ec.img_w
        kinds = [self.ec.unc_kind[i] for i in range(self.ec.unc_n_adaptors)]
        self.has_mask = self.ec.unc_out_ch > 0 and (not kinds or _lib.ADAPT_MASK in kinds)
        self.has_cov = _lib.ADAPT_COV_EXP_TANH in kinds
        self.has_conf = any(k >= _lib.ADAPT_CONF_EXP for k in kinds)
        self.max_pairs = max_pairs
        h = C.c_void_p()
        rc = self.lib.ufm_create(C.byref(self.ec), self.index, C.byref(h))
        if rc != 0:
            raise UfmEngineError(f"ufm_create failed with status {rc} (no sm_100 device?)")
        self.h = h
        for name, t in pack_weights(cfg, state_dict).items():
            shape = (C.c_int64 * t.dim())(*t.shape)
            self._ck(self.lib.ufm_set_weight(self.h, name.encode(), t.data_ptr(), _DT[t.dtype], shape, t.dim()))
        self._ck(self.lib.ufm_prepare(self.h, max_pairs))
        self.view_mode, self.view_high = view_index_range(cfg)
        if self.view_mode == "none" and cfg["info_sharing_str"] != "cross_attention":
            n_rows = state_dict["info_sharing.view_pos_table"].shape[0]
            self._ck(self.lib.ufm_set_view_index(self.h, n_rows))  # the appended zero row

    def draw_view_index(self):
        """use_rand_idx_pe_for_non_reference_views: one draw from torch's global CPU generator per model forward (the model
        classes call this once per batch, before the per-chunk engine calls), exactly how the reference draws it
        (global_attention_transformer.py:521-522), so a seeded run reproduces the reference."""
        if self.view_mode == "rand":
            row = int(torch.randint(low=1, high=self.view_high, size=(1,)).item())
            self._ck(self.lib.ufm_set_view_index(self.h, row))

    def _ck(self, rc):
        if rc != 0:
            raise UfmEngineError(self.lib.ufm_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.ufm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_keep_aux(self, keep: bool):
        self._ck(self.lib.ufm_set_keep_aux(self.h, int(keep)))

    def forward(self, src: torch.Tensor, tgt: torch.Tensor, layout: int) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        """src/tgt: CUDA tensors [B,H,W,3] (NHWC) or [B,3,H,W] (NCHW), uint8 or normalised float32, B <= max_pairs.
        Returns (flow [B,2,H,W] fp32, covisibility [B,H,W] fp32 or None), asynchronous on the current stream."""
        assert src.is_cuda and tgt.is_cuda and src.is_contiguous() and tgt.is_contiguous()
        assert src.dtype == tgt.dtype and src.shape == tgt.shape
        B = src.shape[0]
        flow = torch.empty(B, 2, self.H, self.W, dtype=torch.float32, device=src.device)
        cov = torch.empty(B, self.H, self.W, dtype=torch.float32, device=src.device) if self.has_mask else None
        stream = torch.cuda.current_stream(src.device).cuda_stream
        self._ck(self.lib.ufm_forward(self.h, src.data_ptr(), tgt.data_ptr(), B, _DT[src.dtype], layout,
                                      flow.data_ptr(), cov.data_ptr() if cov is not None else None, stream))
        return flow, cov

    def forward_host(self, src: torch.Tensor, tgt: torch.Tensor, layout: int, flow: torch.Tensor,
                     cov: Optional[torch.Tensor]):
        """Host tensors in (ideally pinned), host tensors out; H2D + forward + D2H + sync inside the C call."""
        assert not src.is_cuda and not flow.is_cuda
        self._ck(self.lib.ufm_forward_host(self.h, src.data_ptr(), tgt.data_ptr(), src.shape[0], _DT[src.dtype], layout,
                                           flow.data_ptr(), cov.data_ptr() if cov is not None else None))

    def set_profile(self, on: bool):
        self._ck(self.lib.ufm_set_profile(self.h, int(on)))

    def profile_read(self) -> Dict[str, Dict[str, float]]:
        """Per kernel family since the last read: summed device ms, algorithmic FLOPs / HBM bytes, launches."""
        n = len(_lib.PROFILE_CATS)
        ms, fl, by, la = (C.c_double * n)(), (C.c_double * n)(), (C.c_double * n)(), (C.c_int64 * n)()
        self._ck(self.lib.ufm_profile_read(self.h, ms, fl, by, la))
        return {name: {"ms": ms[i], "flops": fl[i], "bytes": by[i], "launches": int(la[i])}
                for i, name in enumerate(_lib.PROFILE_CATS)}

    def launch_count(self) -> int:
        return self.lib.ufm_last_launch_count(self.h)

    def stage(self, name: str) -> torch.Tensor:
        """Copy of an internal stage buffer of the last forward, in the engine's own layout."""
        dt, nd = C.c_int(), C.c_int()
        shape = (C.c_int64 * 8)()
        self._ck(self.lib.ufm_stage_desc(self.h, name.encode(), C.byref(dt), shape, C.byref(nd)))
        t = torch.empty(tuple(shape[i] for i in range(nd.value)), dtype=_TD[dt.value], device=self.device)
        self._ck(self.lib.ufm_stage_copy(self.h, name.encode(), t.data_ptr(), t.numel() * t.element_size()))
        return t
