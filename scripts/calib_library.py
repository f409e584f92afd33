#!/usr/bin/env python
"""Library calibration on the same box: what cuBLAS (`torch.matmul`, bf16) and the SDPA
backends (cuDNN, flash, memory-efficient) reach on the exact UFM shapes, next to the repo's own kernels in the same
process.  CUDA-event timing, 3 warm-ups, `--reps` timed calls; operands exceed L2 for the large shapes and each timed
call gets a fresh L2 state from the previous call's output writes.  Library kernels are only ever run here and in
`bench.py --impl torch_gpu`, never on the product path.

    python scripts/calib_library.py [--reps 20] [--skip-own]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--skip-own", action="store_true")
    a = ap.parse_args()
    import torch
    import torch.nn.functional as F
    from torch.nn.attention import SDPBackend, sdpa_kernel

    dev = "cuda"
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

    own = None
    if not a.skip_own:
        import __graft_entry__ as g

        g.build()
        import gpu_ops as own  # noqa: F811
        from ufm_b200 import _lib

        GELU = _lib.GEMM_GELU

    Me, Mi = 32 * 1201, 16 * 2400
    lin = [("enc.qkv", Me, 1024, 3072, "bf16"), ("enc.proj", Me, 1024, 1024, "res"), ("enc.fc1", Me, 1024, 4096, "gelu"),
           ("enc.fc2", Me, 4096, 1024, "res"), ("info.qkv", Mi, 768, 2304, "bf16"), ("info.proj", Mi, 768, 768, "res"),
           ("info.fc1", Mi, 768, 3072, "gelu"), ("info.fc2", Mi, 3072, 768, "res")]
    print("# GEMM: cuBLAS = torch.matmul(A, W^T) bf16 -> bf16 (no bias / activation / residual); cuBLAS+epi = F.linear + the "
          "layer's epilogue in eager torch (bias; GELU; fp32 residual add in place); own = gemm_kernel with the fused epilogue")
    print(f"{'shape':10s} {'M':>6s} {'K':>5s} {'N':>5s} | {'cuBLAS ms':>9s} {'TF/s':>7s} | {'cuBLAS+epi ms':>13s} {'TF/s':>7s} | {'own ms':>7s} {'TF/s':>7s} | own/cuBLAS own/cuBLAS+epi")
    for name, M, K, N, kind in lin:
        A = (torch.randn(M, K, device=dev, generator=gen) * 0.5).bfloat16()
        Wt = (torch.randn(N, K, device=dev, generator=gen) * 0.03).bfloat16()
        bias = torch.randn(N, device=dev, generator=gen)
        biasb = bias.bfloat16()
        x = torch.randn(M, N, device=dev, generator=gen) if kind == "res" else None
        ob = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        fl = 2.0 * M * K * N
        t_mm = timeit(lambda: torch.matmul(A, Wt.t(), out=ob))
        if kind == "res":
            fepi = lambda: x.add_(F.linear(A, Wt, biasb))
        elif kind == "gelu":
            fepi = lambda: F.gelu(F.linear(A, Wt, biasb))
        else:
            fepi = lambda: F.linear(A, Wt, biasb)
        t_epi = timeit(fepi)
        t_own = float("nan")
        if own is not None:
            if kind == "res":
                fn = lambda: own.linear(A, Wt, bias=bias, res1=x, out_f32=x)
            elif kind == "gelu":
                fn = lambda: own.linear(A, Wt, bias=bias, out_bf16=ob, flags=GELU)
            else:
                fn = lambda: own.linear(A, Wt, bias=bias, out_bf16=ob)
            t_own = timeit(fn)
        print(f"{name:10s} {M:6d} {K:5d} {N:5d} | {t_mm:9.3f} {fl / t_mm / 1e9:7.1f} | {t_epi:13.3f} {fl / t_epi / 1e9:7.1f} | "
              f"{t_own:7.3f} {fl / t_own / 1e9:7.1f} | {t_mm / t_own:6.3f} {t_epi / t_own:6.3f}", flush=True)
        del A, Wt, x, ob

    print("\n# attention: F.scaled_dot_product_attention on (seqs, heads, tokens, 64) bf16, per backend, from CONTIGUOUS head-major "
          "q/k/v (the permute copies the reference makes are not timed); own = attn_kernel reading the fused qkv activation")
    shapes = [("enc", 32, 16, 1201), ("info", 16, 12, 2400), ("hires.enc", 8, 16, 4801), ("hires.info", 4, 12, 9600)]
    backends = [("cudnn", SDPBackend.CUDNN_ATTENTION), ("flash", SDPBackend.FLASH_ATTENTION), ("mem_eff", SDPBackend.EFFICIENT_ATTENTION)]
    print(f"{'shape':10s} {'seqs':>4s} {'heads':>5s} {'tokens':>6s} | " + " | ".join(f"{b[0]:>7s} ms {'TF/s':>7s}" for b in backends)
          + f" | {'own ms':>7s} {'TF/s':>7s} | own/best-lib")
    for name, n_seq, heads, S in shapes:
        D = heads * 64
        qkv = torch.randn(n_seq * S, 3 * D, device=dev, generator=gen).bfloat16()
        t = qkv.view(n_seq, S, 3, heads, 64).permute(2, 0, 3, 1, 4).contiguous()
        q, k, v = t[0], t[1], t[2]
        fl = 4.0 * n_seq * heads * S * S * 64
        cells, best = [], float("inf")
        for bname, be in backends:
            try:
                with sdpa_kernel(be):
                    ms = timeit(lambda: F.scaled_dot_product_attention(q, k, v, scale=0.125))
                best = min(best, ms)
                cells.append(f"{ms:10.3f} {fl / ms / 1e9:7.1f}")
            except Exception as e:  # noqa: BLE001
                cells.append(f"{'n/a':>10s} {type(e).__name__[:7]:>7s}")
        t_own = float("nan")
        if own is not None:
            t_own = timeit(lambda: own.attention(qkv, n_seq, S, heads, 0.125))
        print(f"{name:10s} {n_seq:4d} {heads:5d} {S:6d} | " + " | ".join(cells) + f" | {t_own:7.3f} {fl / t_own / 1e9:7.1f} | {best / t_own:6.3f}",
              flush=True)
        del qkv, t, q, k, v


if __name__ == "__main__":
    main()
