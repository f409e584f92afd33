#!/usr/bin/env python
"""BASELINE.json configs[4] / SURVEY.md 8(d) config 5: pairs/s of UFM-Base 560x420 for a TOTAL batch of B pairs sharded
over G GPUs through the repo's own multi-GPU API (ufm_b200.dp.predict_sharded: contiguous shards of ceil(B / G) pairs, one
engine replica per rank, no collective on the data path; the optional NCCL all_gather of the finished outputs on / off).

    torchrun --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port P scripts/sweep_dp.py [--batches 1,2,...]
    python scripts/sweep_dp.py            (G = 1)

Every rank holds the full batch in PINNED HOST memory; a step = H2D of the rank's shard + forward (+ all_gather); timed
with CUDA events between barriers, max over ranks; rank 0 prints CSV rows
    G,B,gather,ms_per_step,pairs_per_s,per_gpu_efficiency
where efficiency = pairs/s / (G x the G = 1, gather-off rate at the same PER-GPU batch if measured in this run, else blank).
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="1,2,4,8,16,32,64,128,256")
    ap.add_argument("--chunk", type=int, default=32, help="max pairs per engine call")
    ap.add_argument("--workload", default="base16")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist

    import __graft_entry__ as g
    import bench

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    out_fd = os.dup(1)
    os.dup2(2, 1)  # NCCL banners go to stderr; stdout carries the CSV only
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    g.build()
    from oracle.golden_util import make_inputs
    from ufm_b200 import UniFlowMatchConfidence
    from ufm_b200.dp import predict_sharded, shard_sizes
    from ufm_b200.weights import make_state_dict

    cfg, _, desc, _ = bench._make_cfg(a.workload)
    model = UniFlowMatchConfidence(max_pairs_per_call=a.chunk, **cfg)
    model.load_state_dict(make_state_dict(cfg, seed=0))
    model = model.to(dev)
    W, H = cfg["inference_resolution"]

    def predict(s, t):
        o = model.predict_correspondences_batched(s.to(dev, non_blocking=True), t.to(dev, non_blocking=True))
        return o.flow.flow_output, o.covisibility.mask

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def emit(line):
        if rank == 0:
            os.write(out_fd, (line + "\n").encode())

    emit(f"# {desc}; total batch B sharded over G = {world} GPU(s) by dp.predict_sharded, engine calls of <= {a.chunk} pairs, "
         f"pinned host uint8 in, device fp32 out (gather: NCCL all_gather_into_tensor of flow + covisibility to every rank)")
    emit("G,B,gather,pairs_per_gpu_max,ms_per_step,pairs_per_s")
    for B in [int(b) for b in a.batches.split(",")]:
        src, tgt = make_inputs(B, H, W, seed=1)
        src, tgt = src.pin_memory(), tgt.pin_memory()
        reps = max(3, min(20, 256 // max(1, B // world)))
        for gather in ((False, True) if world > 1 else (False,)):
            for _ in range(2):
                predict_sharded(predict, src, tgt, gather=gather)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                out = predict_sharded(predict, src, tgt, gather=gather)
            e1.record()
            barrier()
            ms = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            if gather and rank == 0:
                assert out[0].shape[0] == B
            emit(f"{world},{B},{int(gather)},{max(shard_sizes(B, world))},{ms.item():.3f},{B / ms.item() * 1e3:.1f}")
        del src, tgt
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
