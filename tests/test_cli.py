"""Headless CLI: pair-list parsing and batching on the CPU; an end-to-end run on the GPU."""
import os

import numpy as np
import pytest
import torch

from ufm_b200 import cli


def test_pair_list_and_grouping(tmp_path):
    text = "a.png b.png\n# comment\n\n/abs/c.png, d.png  # trailing\n"
    pairs = cli.parse_pair_list(text, base="/data")
    assert pairs == [("/data/a.png", "/data/b.png"), ("/abs/c.png", "/data/d.png")]
    with pytest.raises(ValueError):
        cli.parse_pair_list("only_one.png\n")
    s1, s2 = ((10, 20), (10, 20)), ((12, 20), (10, 20))
    assert cli.group_by_shape([s1, s2, s1, s1, s2], batch=2) == [[0, 2], [3], [1, 4]]
    args = cli.make_parser().parse_args(["batch", "list.txt", "-o", "out", "--model", "refine", "--batch", "4"])
    assert (args.command, args.pairs, args.output, args.model, args.batch) == ("batch", "list.txt", "out", "refine", 4)
    args = cli.make_parser().parse_args(["infer", "a.png", "b.png"])
    assert (args.source, args.target, args.model) == ("a.png", "b.png", "base")


def test_flow_colouring():
    f = np.zeros((2, 8, 8), np.float32)
    f[0, :, 4:] = 3.0
    img = cli.flow_to_color(f)
    assert img.shape == (8, 8, 3) and img.dtype == np.uint8
    assert img[:, :4].max() == 0 and img[:, 4:].max() == 255


@pytest.mark.gpu
def test_cli_batch_matches_api(tmp_path):
    import cv2

    import __graft_entry__ as g

    g.build()
    from ufm_b200 import UniFlowMatchConfidence, get_config

    rng = np.random.default_rng(0)
    # native twice, one foreign size (resize + un-map path), one pair whose target has another size than its source (as the
    # reference's example pairs bike_0 / bike_1 do)
    sizes = [(168, 224), (168, 224), (120, 200), (216, 216)]
    tsizes = [(168, 224), (168, 224), (120, 200), (121, 216)]
    lines = []
    for i, (h, w) in enumerate(sizes):
        for side, (hh, ww) in (("s", (h, w)), ("t", tsizes[i])):
            cv2.imwrite(str(tmp_path / f"{i}{side}.png"), rng.integers(0, 256, (hh, ww, 3), dtype=np.uint8))
        lines.append(f"{i}s.png {i}t.png")
    (tmp_path / "pairs.txt").write_text("\n".join(lines) + "\n")
    out = tmp_path / "out"
    assert cli.main(["batch", str(tmp_path / "pairs.txt"), "-o", str(out), "--config", "tiny_base", "--batch", "2"]) == 0
    model = UniFlowMatchConfidence(**get_config("tiny_base")).to("cuda")
    for i, (h, w) in enumerate(sizes):
        flow = np.load(out / f"{i:06d}_flow.npy")
        assert flow.shape == (2, h, w)
        s = torch.from_numpy(cv2.cvtColor(cv2.imread(str(tmp_path / f"{i}s.png")), cv2.COLOR_BGR2RGB))
        t = torch.from_numpy(cv2.cvtColor(cv2.imread(str(tmp_path / f"{i}t.png")), cv2.COLOR_BGR2RGB))
        ref = model.predict_correspondences_batched(s.cuda(), t.cuda())
        assert np.allclose(flow, ref.flow.flow_output[0].cpu().numpy(), atol=1e-5)
        cov = cv2.imread(str(out / f"{i:06d}_covisibility.png"), cv2.IMREAD_GRAYSCALE)
        assert np.array_equal(cov, (ref.covisibility.mask[0].cpu().numpy() * 255).astype(np.uint8))
        if i >= 2:
            # ... and against the CPU oracle's restatement of the reference shell on the decoded images (foreign sizes)
            from oracle import ufm_oracle as O
            from oracle.golden_util import rel_l2

            with torch.no_grad():
                sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
                orc = O.predict_correspondences_batched(s.unsqueeze(0), t.unsqueeze(0), sd, get_config("tiny_base"))
            assert rel_l2(torch.from_numpy(flow)[None], orc["flow"]) < 1.6e-2
    assert cli.main(["infer", str(tmp_path / "0s.png"), str(tmp_path / "0t.png"), "-o", str(out / "one"), "--config", "tiny_base"]) == 0
    assert os.path.exists(out / "one" / "flow.npy") and os.path.exists(out / "one" / "covisibility_mask.png")
    assert os.path.exists(out / "one" / "flow_visualization.png")
