#!/usr/bin/env python
"""Micro-benchmark of the tcgen05 GEMM / implicit-GEMM conv kernel on the UFM-Base shapes (B = 16 pairs), through the
kernel-level C-ABI entry point.  CUDA-event timing, inputs larger than L2 for the big shapes; prints ms and TFLOP/s per
shape and epilogue variant.

    python scripts/gemm_bench.py [--variants 0,1] [--reps 20]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants", default="0:1,1:1", help="comma list of epilogue:pair[:split[:rows2[:warp_shift[:gelu_packed]]]]  (split: last-wave column "
                    "slices, default 1; rows2: 512 x 256 pair tiles, ufm_debug_flag 6, default 0; warp_shift: ufm_debug_flag 7; "
                    "gelu_packed: bit 0 = ufm_debug_flag 9, bit 1 = ufm_debug_flag 10 (L2 prefetch of in-place residual slabs))")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--only", default="")
    ap.add_argument("--no-conv", action="store_true", help="row-matrix GEMM shapes only")
    a = ap.parse_args()
    import torch

    import __graft_entry__ as g

    g.build()
    import gpu_ops as ops
    from ufm_b200 import _lib

    lib = _lib.load()
    dev = "cuda"
    Me, Mi = 32 * 1201, 16 * 2400
    GELU = _lib.GEMM_GELU
    # name, M, K, N, epilogue kind
    lin = [("enc.qkv", Me, 1024, 3072, "bf16"), ("enc.proj", Me, 1024, 1024, "res"), ("enc.fc1", Me, 1024, 4096, "gelu"),
           ("enc.fc2", Me, 4096, 1024, "res"), ("info.qkv", Mi, 768, 2304, "bf16"), ("info.proj", Mi, 768, 768, "res"),
           ("info.fc1", Mi, 768, 3072, "gelu"), ("info.fc2", Mi, 3072, 768, "res")]
    # name, n, H, W, Cin, Cout, kind
    conv = [("rcu@120x160", 16, 120, 160, 256, 256, "convres"), ("rcu@240x320", 4, 240, 320, 256, 256, "convres"),
            ("rcu1.c2@120x160", 16, 120, 160, 256, 256, "convres2"), ("rcu2.c2@120x160", 16, 120, 160, 256, 256, "convresbf"),
            ("rcu2.c2@60x80", 32, 60, 80, 256, 256, "convresbf"),
            ("proc.c1@240x320", 4, 240, 320, 256, 128, "convbf16"), ("proc.c2@420x560", 4, 420, 560, 128, 128, "convbf16")]
    variants = [tuple(int(x) for x in v.split(":")) for v in a.variants.split(",")]
    variants = [v if len(v) > 2 else v + (1,) for v in variants]
    variants = [v if len(v) > 3 else v + (0,) for v in variants]
    variants = [v if len(v) > 4 else v + (0,) for v in variants]
    variants = [v if len(v) > 5 else v + (0,) for v in variants]
    gen = torch.Generator(device=dev).manual_seed(0)

    def timeit(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / a.reps

    print(f"{'shape':18s} {'M':>7s} {'K':>5s} {'N':>5s} " + "   ".join(f"e{v[0]}p{v[1]}s{v[2]}r{v[3]}w{v[4]}g{v[5]}: ms TFLOP/s" for v in variants))
    for name, M, K, N, kind in lin:
        if a.only and a.only not in name:
            continue
        A = (torch.randn(M, K, device=dev, generator=gen) * 0.5).bfloat16()
        Wt = (torch.randn(N, K, device=dev, generator=gen) * 0.03).bfloat16()
        bias = torch.randn(N, device=dev, generator=gen)
        x = torch.randn(M, N, device=dev, generator=gen) if kind == "res" else None
        ob = torch.empty(M, N, dtype=torch.bfloat16, device=dev) if kind != "res" else None
        res = []
        for v in variants:
            lib.ufm_debug_flag(0, v[0])
            lib.ufm_debug_flag(1, v[1])
            lib.ufm_debug_flag(5, v[2])
            lib.ufm_debug_flag(6, v[3])
            lib.ufm_debug_flag(7, v[4])
            lib.ufm_debug_flag(9, v[5] & 1)
            lib.ufm_debug_flag(10, (v[5] >> 1) & 1)
            if kind == "res":
                fn = lambda: ops.linear(A, Wt, bias=bias, res1=x, out_f32=x)
            elif kind == "gelu":
                fn = lambda: ops.linear(A, Wt, bias=bias, out_bf16=ob, flags=GELU)
            else:
                fn = lambda: ops.linear(A, Wt, bias=bias, out_bf16=ob)
            ms = timeit(fn)
            res.append(f"{ms:7.3f} {2.0 * M * K * N / ms / 1e9:7.1f}")
        print(f"{name:18s} {M:7d} {K:5d} {N:5d} " + "   ".join(res))
        del A, Wt, x, ob
    for name, n, H, W, Ci, Co, kind in ([] if a.no_conv else conv):
        if a.only and a.only not in name:
            continue
        X = (torch.randn(n, H, W, Ci, device=dev, generator=gen) * 0.5).bfloat16()
        Wp = (torch.randn(Co, 9 * Ci, device=dev, generator=gen) * 0.02).bfloat16()
        bias = torch.randn(Co, device=dev, generator=gen)
        r = torch.randn(n, H, W, Co, device=dev, generator=gen)
        r2 = torch.randn(n, H, W, Co, device=dev, generator=gen)
        of = torch.empty(n, H, W, Co, device=dev)
        ob = torch.empty(n, H, W, Co, dtype=torch.bfloat16, device=dev)
        res = []
        for v in variants:
            lib.ufm_debug_flag(0, v[0])
            lib.ufm_debug_flag(1, v[1])
            lib.ufm_debug_flag(7, v[4])
            if kind == "convres2":  # fusion block: path + skip + conv, fp32 sum + ReLU'd bf16 copy
                fn = lambda: ops.conv3x3(X, Wp, bias=bias, res1=r, res2=r2, r2=(H * W * Co, W * Co, Co), out_f32=of,
                                         out_bf16=ob, flags=_lib.GEMM_RELU_BF16)
            elif kind == "convresbf":  # x + conv, bf16 only
                fn = lambda: ops.conv3x3(X, Wp, bias=bias, res1=r, out_bf16=ob)
            elif kind == "convres":
                fn = lambda: ops.conv3x3(X, Wp, bias=bias, res1=r, out_f32=of, out_bf16=ob, flags=_lib.GEMM_RELU_BF16)
            else:
                fn = lambda: ops.conv3x3(X, Wp, bias=bias, out_bf16=ob)
            ms = timeit(fn)
            res.append(f"{ms:7.3f} {2.0 * n * H * W * 9 * Ci * Co / ms / 1e9:7.1f}")
        print(f"{name:18s} {n * H * W:7d} {9 * Ci:5d} {Co:5d} " + "   ".join(res))
    lib.ufm_debug_flag(0, 1)
    lib.ufm_debug_flag(1, -1)
    lib.ufm_debug_flag(5, 1)
    lib.ufm_debug_flag(6, -1)
    lib.ufm_debug_flag(7, 0)
    lib.ufm_debug_flag(9, -1)
    lib.ufm_debug_flag(10, -1)


if __name__ == "__main__":
    main()
