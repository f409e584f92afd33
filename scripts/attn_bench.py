#!/usr/bin/env python
"""Micro-benchmark of the attention kernel on the UFM shapes (B = 16 pairs): encoder 32 x 16 heads x 1201 tokens, info sharing
16 x 12 heads x 2400 tokens, hi-res encoder 8 x 16 x 4801; CUDA-event timing; also checks the result against torch SDPA
(fp32 math on the same bf16 inputs).

    python scripts/attn_bench.py [variant,variant,...]     (variants = ufm_debug_flag(2, v) tuning switches)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.nn.functional as F

    import __graft_entry__ as g

    g.build()
    import gpu_ops as ops
    from ufm_b200 import _lib

    lib = _lib.load()
    variants = [int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "4").split(",")]
    for name, n_seq, S, heads in [("enc", 32, 1201, 16), ("info", 16, 2400, 12), ("hires.enc", 8, 4801, 16)]:
        D = heads * 64
        gen = torch.Generator(device="cuda").manual_seed(0)
        qkv = torch.randn(n_seq * S, 3 * D, device="cuda", generator=gen).bfloat16()
        t = qkv[: 2 * S].float().view(2, S, 3, heads, 64).permute(2, 0, 3, 1, 4)
        ref = F.scaled_dot_product_attention(t[0], t[1], t[2], scale=0.125).transpose(1, 2).reshape(2 * S, D)
        for var in variants:
            lib.ufm_debug_flag(2, var)
            for _ in range(3):
                out = ops.attention(qkv, n_seq, S, heads, 0.125)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 20
            e0.record()
            for _ in range(reps):
                out = ops.attention(qkv, n_seq, S, heads, 0.125)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            fl = 4.0 * n_seq * heads * S * S * 64
            err = ((out[: 2 * S].float() - ref).norm() / ref.norm()).item()
            print(f"v{var:<3d} {name:10s} seqs {n_seq:3d} tokens {S:5d} heads {heads:2d}: {ms:7.3f} ms  "
                  f"{fl / ms / 1e9:7.1f} TFLOP/s  rel-L2 {err:.2e}")


if __name__ == "__main__":
    main()
