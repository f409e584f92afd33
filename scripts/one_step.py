#!/usr/bin/env python
"""One forward pass of a bench workload, for ncu: `warm` untimed passes, then one pass.

    python scripts/one_step.py [--workload base16] [--pairs N] [--warm 2]

Prints the launch count of one pass so that `ncu -s <warm*launches> -c <launches>` isolates the last pass.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="base16")
    ap.add_argument("--pairs", type=int, default=0)
    ap.add_argument("--warm", type=int, default=2)
    a = ap.parse_args()
    import torch

    import __graft_entry__ as g
    import bench

    g.build()
    from oracle.golden_util import make_inputs
    from ufm_b200 import UniFlowMatchClassificationRefinement, UniFlowMatchConfidence, _lib
    from ufm_b200.weights import make_state_dict

    cfg, pairs, desc, name = bench._make_cfg(a.workload)
    pairs = a.pairs or pairs
    refine = "classification_head_kwargs" in cfg
    cls = UniFlowMatchClassificationRefinement if refine else UniFlowMatchConfidence
    model = cls(max_pairs_per_call=pairs, **cfg)
    model.load_state_dict(make_state_dict(cfg, seed=0))
    model = model.to("cuda:0")
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(pairs, H, W, seed=1)
    dsrc, dtgt = src.cuda(), tgt.cuda()
    eng = model._engine(torch.device("cuda:0"))
    for _ in range(a.warm + 1):
        eng.forward(dsrc, dtgt, _lib.UFM_NHWC)
    torch.cuda.synchronize()
    print(f"one_step: workload {a.workload} pairs {pairs} launches_per_pass {eng.launch_count()}")


if __name__ == "__main__":
    main()
