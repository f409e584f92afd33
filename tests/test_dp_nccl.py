"""dp.predict_sharded with the REAL engine on two GPUs (NCCL): the gathered full-batch result must be bit-identical to
the single-GPU result of the same batch (pairs are independent: sharding may not change a pair's output).  Skipped on a
box with fewer than two GPUs (the CPU-side logic is covered by tests/test_dp_gloo.py with world size 2)."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, B, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import __graft_entry__ as g

        g.build()
        from oracle.golden_util import make_inputs
        from ufm_b200 import UniFlowMatchConfidence, get_config
        from ufm_b200.dp import predict_sharded
        from ufm_b200.weights import make_state_dict

        cfg = get_config("tiny_base")
        m = UniFlowMatchConfidence(max_pairs_per_call=4, **cfg).to(dev)
        m.load_state_dict(make_state_dict(cfg, seed=0, out_gain=8.0))
        W, H = cfg["inference_resolution"]
        src, tgt = make_inputs(B, H, W, seed=17)

        def predict(s, t):
            o = m.predict_correspondences_batched(s.to(dev), t.to(dev))
            return o.flow.flow_output, o.covisibility.mask

        flow, cov = predict_sharded(predict, src, tgt, gather=True)  # host tensors in: rank 1's shard is empty for B = 1
        one_f, one_c = predict(src, tgt)  # the whole batch on this GPU alone
        torch.cuda.synchronize()
        q.put((rank, bool(torch.equal(flow, one_f) and torch.equal(cov, one_c) and flow.shape[0] == B)))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("B", [6, 5, 1])
def test_predict_sharded_two_gpus_bit_identical(B):
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, B, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res == [(0, True), (1, True)]
