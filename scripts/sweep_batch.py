#!/usr/bin/env python
"""SURVEY.md 8(d) config 5, single-GPU leg: pairs/s of UFM-Base 560x420 against the batch size B (device tensors in,
CUDA-event timing, engine calls of at most --chunk pairs).  Multi-GPU rows come from `bench.py --gpus N` (weak scaling,
contiguous shards of ceil(B/G) pairs per rank: a B-pair job on G GPUs runs at the B/G row of this table per GPU).

    python scripts/sweep_batch.py [--workload base16] [--chunk 32] [--batches 1,2,4,8,16,32,64,128]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="base16")
    ap.add_argument("--chunk", type=int, default=32)
    ap.add_argument("--batches", default="1,2,4,8,16,32,64,128")
    ap.add_argument("--host", action="store_true", help="pinned host tensors in, host tensors out (ufm_forward_host)")
    a = ap.parse_args()
    import torch

    import __graft_entry__ as g
    import bench

    g.build()
    from oracle.golden_util import make_inputs
    from ufm_b200 import UniFlowMatchClassificationRefinement, UniFlowMatchConfidence
    from ufm_b200.weights import make_state_dict

    cfg, _, desc, _ = bench._make_cfg(a.workload)
    refine = "classification_head_kwargs" in cfg
    cls = UniFlowMatchClassificationRefinement if refine else UniFlowMatchConfidence
    model = cls(max_pairs_per_call=a.chunk, **cfg)
    model.load_state_dict(make_state_dict(cfg, seed=0))
    model = model.to("cuda:0")
    W, H = cfg["inference_resolution"]
    print(f"# {desc}; engine calls of <= {a.chunk} pairs; {'pinned host' if a.host else 'device'} tensors in/out")
    print("B,ms_per_call,pairs_per_s")
    for B in [int(b) for b in a.batches.split(",")]:
        src, tgt = make_inputs(B, H, W, seed=1)
        dsrc, dtgt = (src.pin_memory(), tgt.pin_memory()) if a.host else (src.cuda(), tgt.cuda())
        reps = max(3, min(20, 256 // B))
        for _ in range(2):
            model.predict_correspondences_batched(dsrc, dtgt)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            model.predict_correspondences_batched(dsrc, dtgt)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        if a.host:
            import time

            t0 = time.perf_counter()
            for _ in range(reps):
                model.predict_correspondences_batched(dsrc, dtgt)
            ms = (time.perf_counter() - t0) * 1e3 / reps  # the host path synchronises inside the call
        print(f"{B},{ms:.2f},{B / ms * 1e3:.1f}", flush=True)


if __name__ == "__main__":
    main()
