#!/usr/bin/env python
"""Micro-benchmark of the local-window refinement kernel: B = 16 pairs at 560x420, 16-channel bf16
descriptors, with an INJECTED flow f = smooth field U(-40,40) px + N(0,1) so that the gather locality is not the trivially
best case random-init weights give (|flow| < 0.07 px).  CUDA-event timing; reports ms and GB/s of algorithmic bytes
(F1, F2 bf16 + flow in + flow out = 80 B / pixel).

    python scripts/refine_bench.py [--pairs 16] [--reps 20]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=16)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    import torch
    import torch.nn.functional as F

    import __graft_entry__ as g

    g.build()
    from ufm_b200 import _lib

    lib = _lib.load()
    B, H, W = a.pairs, 420, 560
    gen = torch.Generator(device="cuda").manual_seed(0)
    desc = torch.randn(B, 2, 2, H, W, 8, device="cuda", generator=gen).bfloat16()
    bias = torch.zeros(25, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for name, amp, noise in [("natural (|flow| < 0.1 px)", 0.1, 0.0), ("injected smooth +-40 px + N(0,1)", 40.0, 1.0)]:
        base = F.interpolate((torch.rand(B, 2, 4, 5, device="cuda", generator=gen) * 2 - 1) * amp, size=(H, W),
                             mode="bilinear", align_corners=True)
        flow = (base + noise * torch.randn(B, 2, H, W, device="cuda", generator=gen)).contiguous()
        out = torch.empty_like(flow)
        run = lambda: lib.ufm_op_refine(desc.data_ptr(), B, H, W, flow.data_ptr(), bias.data_ptr(), 0.25, out.data_ptr(),
                                        None, None, st)
        for _ in range(3):
            assert run() == 0
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.reps
        by = B * H * W * (2 * 16 * 2 + 2 * 4 + 2 * 4)
        print(f"refine {name:36s} B={B}: {ms:7.3f} ms  {by / ms / 1e6:7.1f} GB/s algorithmic ({by / 1e6:.0f} MB)")


if __name__ == "__main__":
    main()
