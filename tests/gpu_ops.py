"""Thin torch wrappers over the kernel-level C-ABI entry points (test helper, GPU only)."""
import ctypes as C

import torch

from ufm_b200 import _lib


def _p(t):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def gemm(a, wt, *, X=None, Y=1, Nimg=1, C_=None, a_strides=None, tw=128, th=1, taps=1, cin_pad=None, bias=None, scale=None,
         res1=None, res2=None, r2=(0, 0, 0), out_f32=None, out_bf16=None, out_strides=None, out_off=0, flags=0,
         tail=None, out2=None):
    """a: bf16 tensor; logical extents (C_, X, Y, Nimg) with element strides a_strides=(sx, sy, sn)."""
    lib = _lib.load()
    n_out, ktot = wt.shape
    if cin_pad is None:
        cin_pad = ktot // taps
    err = C.create_string_buffer(1024)
    tw2, tb2, ts, tsh, tout = None, None, None, None, 0
    keep = []
    if tail is not None:
        w2, b2, sc, sh = tail
        tout = w2.shape[0]
        tw2 = w2.data_ptr()
        arr = lambda v: (C.c_float * 4)(*(list(v) + [0.0] * (4 - len(v))))
        tb2, ts, tsh = arr(b2), arr(sc), arr(sh)
        keep += [tb2, ts, tsh]
        tb2, ts, tsh = C.cast(tb2, C.c_void_p), C.cast(ts, C.c_void_p), C.cast(tsh, C.c_void_p)
    sx, sy, sn = a_strides
    osx, osy, osn = out_strides
    rc = lib.ufm_op_gemm(_p(a), C_, X, Y, Nimg, sx, sy, sn, tw, th, _p(wt), n_out, taps, cin_pad, _p(bias), _p(scale),
                         _p(res1), _p(res2), r2[0], r2[1], r2[2], _p(out_f32), _p(out_bf16), osn, osy, osx, out_off,
                         flags, tw2, tb2, ts, tsh, tout, _p(out2), _stream(), err, 1024)
    if rc != 0:
        raise RuntimeError(err.value.decode())


def linear(a2d, wt, **kw):
    """a2d [M, K] bf16 row-major, wt [N, K] bf16."""
    M, K = a2d.shape
    N = wt.shape[0]
    return gemm(a2d, wt, X=M, C_=K, a_strides=(K, M * K, M * K), out_strides=(N, 0, 0), **kw)


def conv3x3(x_nhwc, w_packed, *, tw=-1, th=0, **kw):
    n, H, W, Cc = x_nhwc.shape
    N = w_packed.shape[0]
    return gemm(x_nhwc, w_packed, X=W, Y=H, Nimg=n, C_=Cc, a_strides=(Cc, W * Cc, H * W * Cc), tw=tw, th=th, taps=9,
                out_strides=(N, W * N, H * W * N), **kw)


def attention(qkv, n_seq, seq_len, heads, scale):
    lib = _lib.load()
    out = torch.empty(n_seq * seq_len, heads * 64, dtype=torch.bfloat16, device=qkv.device)
    err = C.create_string_buffer(1024)
    rc = lib.ufm_op_attention(qkv.data_ptr(), n_seq, seq_len, heads, scale, out.data_ptr(), _stream(), err, 1024)
    if rc != 0:
        raise RuntimeError(err.value.decode())
    return out
