#!/usr/bin/env python
"""One launch of cuBLAS (torch.matmul, bf16) and one of gemm_kernel per UFM encoder / info-sharing GEMM shape, for an
`ncu --set full` capture that puts the library's tile / cluster choice (kernel name), L2 -> SM traffic and tensor-pipe
share next to the repo's kernel on the same shapes (calibration only; nothing here is on the product path).

    python scripts/gemm_probe.py [--reps 1] [--shapes enc.fc1,enc.fc2]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=1)
    ap.add_argument("--shapes", default="enc.qkv,enc.proj,enc.fc1,enc.fc2,info.fc2")
    a = ap.parse_args()
    import torch

    import __graft_entry__ as g

    g.build()
    import gpu_ops as own
    from ufm_b200 import _lib

    Me, Mi = 32 * 1201, 16 * 2400
    lin = {"enc.qkv": (Me, 1024, 3072, "bf16"), "enc.proj": (Me, 1024, 1024, "res"), "enc.fc1": (Me, 1024, 4096, "gelu"),
           "enc.fc2": (Me, 4096, 1024, "res"), "info.qkv": (Mi, 768, 2304, "bf16"), "info.proj": (Mi, 768, 768, "res"),
           "info.fc1": (Mi, 768, 3072, "gelu"), "info.fc2": (Mi, 3072, 768, "res")}
    gen = torch.Generator(device="cuda").manual_seed(0)
    for name in a.shapes.split(","):
        M, K, N, kind = lin[name]
        A = (torch.randn(M, K, device="cuda", generator=gen) * 0.5).bfloat16()
        Wt = (torch.randn(N, K, device="cuda", generator=gen) * 0.03).bfloat16()
        bias = torch.randn(N, device="cuda", generator=gen)
        x = torch.randn(M, N, device="cuda", generator=gen) if kind == "res" else None
        ob = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        for _ in range(a.reps):
            torch.matmul(A, Wt.t(), out=ob)
            if kind == "res":
                own.linear(A, Wt, bias=bias, res1=x, out_f32=x)
            elif kind == "gelu":
                own.linear(A, Wt, bias=bias, out_bf16=ob, flags=_lib.GEMM_GELU)
            else:
                own.linear(A, Wt, bias=bias, out_bf16=ob)
        torch.cuda.synchronize()
        print(f"{name}: M {M} K {K} N {N} done", flush=True)


if __name__ == "__main__":
    main()
