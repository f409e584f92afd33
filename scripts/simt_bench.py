#!/usr/bin/env python3
"""Times the bandwidth-bound kernels of the UFM-Base step standalone, at the step's own shapes (B pairs at 560x420):
LayerNorm of the encoder / info-sharing residual streams and the align-corners resizes of the DPT head.  Prints the
achieved GB/s of each (algorithmic bytes = input + output of one launch) next to the measured copy bandwidth.

    python scripts/simt_bench.py [--pairs 16] [--iters 20] [--only ln|resize]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ufm_b200 import _lib  # noqa: E402


def timeit(fn, iters, flush):
    st = torch.cuda.current_stream()
    for _ in range(3):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        flush.add_(1)  # evict: 512 MB > L2
        a.record(st)
        fn()
        b.record(st)
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=16)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    lib = _lib.load()
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    flush = torch.zeros(128 << 20, dtype=torch.float32, device=dev)
    peak = None
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        try:
            peak = json.load(open(pk))
        except Exception:
            peak = None
    print("peaks:", peak)
    B = args.pairs
    if args.only in ("", "ln"):
        for D, S in ((1024, 1201), (768, 1200)):
            nseq = 2 * B
            x = torch.randn(nseq, S, D, device=dev)
            w = torch.randn(D, device=dev)
            b = torch.randn(D, device=dev)
            ob = torch.empty(nseq * S, D, dtype=torch.bfloat16, device=dev)

            def f():
                assert lib.ufm_op_layernorm(x.data_ptr(), nseq, S, 0, S, D, w.data_ptr(), b.data_ptr(), 1e-6, 0, ob.data_ptr(),
                                            D, 0, None, 0, 0, st) == 0

            ms = timeit(f, args.iters, flush)
            gb = nseq * S * D * 6 / 1e9
            print(f"layernorm D={D} rows={nseq * S}: {ms * 1e3:8.1f} us  {gb / ms * 1e3:8.1f} GB/s")
    if args.only in ("", "resize"):
        # DPT fusion-block x2 upsamples (fp32 in; bf16 or fp32 out) and the head's 8x-token -> image resize (bf16 -> bf16)
        for (h, w_, C, oh, ow, kind) in ((30, 40, 256, 60, 80, "f32"), (60, 80, 256, 120, 160, "f32"), (120, 160, 256, 240, 320, "f32"),
                                          (120, 160, 256, 240, 320, "f2b"), (240, 320, 128, 420, 560, "b2b")):
            n = B
            if kind == "b2b":
                xin = torch.randn(n, h, w_, C, device=dev).bfloat16()
                out = torch.empty(n, oh, ow, C, dtype=torch.bfloat16, device=dev)
                args_ = (None, xin.data_ptr(), n, h, w_, C, oh, ow, oh, ow, None, out.data_ptr(), st)
                gb = n * C * 2 * (h * w_ + oh * ow) / 1e9
            else:
                xin = torch.randn(n, h, w_, C, device=dev)
                if kind == "f32":
                    out = torch.empty(n, oh, ow, C, device=dev)
                    args_ = (xin.data_ptr(), None, n, h, w_, C, oh, ow, oh, ow, out.data_ptr(), None, st)
                    gb = n * C * 4 * (h * w_ + oh * ow) / 1e9
                else:
                    out = torch.empty(n, oh, ow, C, dtype=torch.bfloat16, device=dev)
                    args_ = (xin.data_ptr(), None, n, h, w_, C, oh, ow, oh, ow, None, out.data_ptr(), st)
                    gb = n * C * (4 * h * w_ + 2 * oh * ow) / 1e9

            def f():
                assert lib.ufm_op_resize_ac(*args_) == 0

            ms = timeit(f, args.iters, flush)
            print(f"resize {kind} {h}x{w_}x{C} -> {oh}x{ow}: {ms * 1e3:8.1f} us  {gb / ms * 1e3:8.1f} GB/s")


if __name__ == "__main__":
    main()
