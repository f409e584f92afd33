"""Data-parallel host logic (ufm_b200/dp.py) on CPU: contiguous batch shards over world_size 2 with the gloo backend, ragged
and empty shards, gather order.  The engine itself has no CPU path, so the per-rank `predict` is a deterministic stand-in
(a per-pair function of the inputs); what is tested is the sharding / gathering around it."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ufm_b200.dp import predict_sharded, shard_range, shard_sizes


def test_shard_ranges_partition_the_batch():
    for B in [0, 1, 2, 5, 16, 17, 255, 256]:
        for G in [1, 2, 4, 8]:
            r = [shard_range(B, G, k) for k in range(G)]
            assert r[0][0] == 0 and r[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))            # contiguous, in rank order
            assert max(hi - lo for lo, hi in r) == (B + G - 1) // G       # ceil(B / G) per rank, never padded
            assert sum(shard_sizes(B, G)) == B
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _fake_predict(src, tgt):
    """Stand-in for model.predict_correspondences_batched: a per-pair function of the pixels, so a wrong shard or a wrong
    gather order changes the result."""
    b, H, W, _ = src.shape
    s = src.float().mean(dim=(1, 2, 3)) - tgt.float().mean(dim=(1, 2, 3))
    flow = s.view(b, 1, 1, 1).expand(b, 2, H, W).contiguous()
    cov = torch.sigmoid(s).view(b, 1, 1).expand(b, H, W).contiguous()
    return flow, cov


def _worker(rank, world, port, B, out_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(7)
        src = torch.randint(0, 256, (B, 6, 8, 3), dtype=torch.uint8, generator=g)
        tgt = torch.randint(0, 256, (B, 6, 8, 3), dtype=torch.uint8, generator=g)
        calls = []

        def predict(s, t):
            calls.append(s.shape[0])
            return _fake_predict(s, t)

        flow, cov = predict_sharded(predict, src, tgt, gather=True)
        ref_flow, ref_cov = _fake_predict(src, tgt)
        lo, hi = shard_range(B, world, rank)
        ok = torch.equal(flow, ref_flow) and torch.equal(cov, ref_cov) and calls == ([hi - lo] if hi > lo else [])
        lf, lc = predict_sharded(_fake_predict, src, tgt, gather=False)
        if hi > lo:
            ok = ok and torch.equal(lf, ref_flow[lo:hi]) and torch.equal(lc, ref_cov[lo:hi])
        else:
            ok = ok and lf is None
        out_q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("B", [5, 1, 4])
def test_predict_sharded_world2_gloo(B):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, B, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True), (1, True)]
