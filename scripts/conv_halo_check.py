#!/usr/bin/env python
"""3x3 implicit-GEMM conv: legacy 9-box A operand vs halo-resident A operand (ufm_debug_flag 3), parity and timing.

    python scripts/conv_halo_check.py [modes, e.g. 0,1,3]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch
import torch.nn.functional as F


def main():
    import __graft_entry__ as g

    g.build()
    import gpu_ops as ops
    from ufm_b200 import _lib
    from ufm_b200.packing import _conv3_pack

    lib = _lib.load()
    if os.environ.get("GEMM_PAIR"):
        lib.ufm_debug_flag(1, int(os.environ["GEMM_PAIR"]))  # bit 0: CTA pairs for BN = 256, bit 1: also for BN = 128
    modes = [int(m) for m in (sys.argv[1] if len(sys.argv) > 1 else "0,1,3").split(",")]
    gen = torch.Generator(device="cpu").manual_seed(0)
    only = os.environ.get("CONV_ONLY", "")  # substring filter on the timing shapes; also skips the parity part
    # parity on small shapes (edges, ragged tiles, channel remainder)
    for (n, H, W, Ci, Co) in [] if only else [(2, 30, 40, 96, 64), (1, 15, 20, 64, 64), (2, 24, 32, 256, 256), (1, 42, 56, 32, 128), (1, 21, 28, 128, 64)]:
        x = torch.randn(n, H, W, Ci, generator=gen).cuda().bfloat16()
        w = (torch.randn(Co, Ci, 3, 3, generator=gen) * (9 * Ci) ** -0.5).bfloat16().float()
        bias = torch.randn(Co, generator=gen).cuda()
        ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.cuda(), bias, padding=1).permute(0, 2, 3, 1)
        wp = _conv3_pack(w).cuda()
        for m in modes:
            lib.ufm_debug_flag(3, m)
            of = torch.zeros(n, H, W, Co, device="cuda")
            ops.conv3x3(x, wp, bias=bias, out_f32=of)
            torch.cuda.synchronize()
            r = ((of - ref).norm() / ref.norm()).item()
            print(f"parity mode {m} shape {(n, H, W, Ci, Co)}: rel-L2 {r:.3e} {'OK' if r < 1e-5 else 'MISMATCH'}", flush=True)
    # timing on the engine's shapes (B = 16 heads: 16 images; UNet: 32 images)
    shapes = [("rcu@120x160", 16, 120, 160, 256, 256), ("rcu@240x320", 16, 240, 320, 256, 256),
              ("proc.c1@240x320", 16, 240, 320, 256, 128), ("proc.c2@420x560", 16, 420, 560, 128, 128),
              ("unet 64->64@420x560", 32, 420, 560, 64, 64), ("unet 128->64@420x560", 32, 420, 560, 128, 64),
              ("unet 256->128@210x280", 32, 210, 280, 256, 128), ("unet 512->256@105x140", 32, 105, 140, 512, 256)]
    for name, n, H, W, Ci, Co in shapes:
        if only and only not in name:
            continue
        x = torch.randn(n, H, W, Ci, device="cuda").bfloat16()
        wp = _conv3_pack(torch.randn(Co, Ci, 3, 3) * (9 * Ci) ** -0.5).cuda()
        ob = torch.zeros(n, H, W, Co, dtype=torch.bfloat16, device="cuda")
        line = f"{name:24s}"
        for m in modes:
            if m == 3:
                continue
            lib.ufm_debug_flag(3, m)
            for _ in range(3):
                ops.conv3x3(x, wp, out_bf16=ob, flags=_lib.GEMM_RELU_BF16)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                ops.conv3x3(x, wp, out_bf16=ob, flags=_lib.GEMM_RELU_BF16)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            line += f"  mode {m}: {ms:7.3f} ms {2.0 * n * H * W * 9 * Ci * Co / ms / 1e9:7.1f} TFLOP/s"
        print(line, flush=True)
    lib.ufm_debug_flag(3, 1)


if __name__ == "__main__":
    main()
