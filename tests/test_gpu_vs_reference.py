"""Engine vs the UNMODIFIED reference run on the same B200 (SURVEY.md H5 iv), at the full architecture sizes.

Three results per case, on the same seeded weights and inputs:
  T  the reference on the CPU in fp32 (autocast is a no-op without CUDA: the "truth" every golden fixture records),
  R  the reference on the GPU under its own recipe -- bf16 autocast for encoder + info sharing (ref base.py:272-274),
     fp32 with TF32 matmuls / cuDNN convs for the heads, the classification MLP and the refinement (ufm.py:635, :921,
     transformer_blocks.py:20),
  E  the engine (bf16 operands everywhere, fp32 accumulation / LayerNorm / residuals).
|R - T| is the error the reference's own GPU recipe has; the engine is held to a small multiple of it, and |E - R| is
reported.  Skipped when no reference tree is present (oracle/_ref is staged by oracle/build_ref.py and travels to the
GPU box; /root/reference exists only in the build container).
"""
import json
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ref_harness as R  # noqa: E402
from oracle.golden_util import make_inputs, max_abs, rel_l2  # noqa: E402
from ufm_b200 import UniFlowMatchClassificationRefinement, UniFlowMatchConfidence, get_config  # noqa: E402
from ufm_b200.weights import make_state_dict  # noqa: E402

needs_ref = pytest.mark.skipif(not R.reference_available(), reason="no reference tree (oracle/_ref not staged)")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")


@pytest.fixture(scope="module", autouse=True)
def _built():
    import __graft_entry__ as g

    g.build()
    assert torch.cuda.is_available(), "GPU tests need a B200"
    torch.set_num_threads(len(os.sched_getaffinity(0)))


def _three_way(cfg_name, res, gain, batch, seed=7, cpu_truth=True):
    from ufm_b200 import configs

    cfg = get_config(cfg_name) if res is None else getattr(configs, cfg_name)(res)
    sd = make_state_dict(cfg, seed=0, out_gain=gain)
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(batch, H, W, seed=seed)
    refine = "classification_head_kwargs" in cfg
    ref = R.build_reference(cfg, sd)
    out = {}
    with torch.no_grad():
        if cpu_truth:
            t = ref.predict_correspondences_batched(src, tgt)
            out["T"] = (t.flow.flow_output, t.covisibility.mask)
        ref = ref.to("cuda")
        r = ref.predict_correspondences_batched(src.cuda(), tgt.cuda())
        out["R"] = (r.flow.flow_output.cpu(), r.covisibility.mask.cpu())
    del ref
    torch.cuda.empty_cache()
    cls = UniFlowMatchClassificationRefinement if refine else UniFlowMatchConfidence
    m = cls(max_pairs_per_call=batch, **cfg).to("cuda")
    m.load_state_dict(sd)
    e = m.predict_correspondences_batched(src.cuda(), tgt.cuda())
    torch.cuda.synchronize()
    out["E"] = (e.flow.flow_output.cpu(), e.covisibility.mask.cpu())
    del m
    torch.cuda.empty_cache()
    return out, refine


def _rows(tag, out):
    rows = {}
    pairs = [("E", "R")] + ([("R", "T"), ("E", "T")] if "T" in out else [])
    for a, b in pairs:
        rows[f"{a}-{b}"] = {"flow_rel_l2": rel_l2(out[a][0], out[b][0]), "flow_max_abs_px": max_abs(out[a][0], out[b][0]),
                            "covis_max_abs": max_abs(out[a][1], out[b][1])}
    rows["flow_max_px"] = out["R"][0].abs().max().item()
    print(f"\n[{tag}] max |flow| {rows['flow_max_px']:.3f} px")
    for k, v in rows.items():
        if isinstance(v, dict):
            print(f"  {k}: flow rel-L2 {v['flow_rel_l2']:.3e}  max-abs {v['flow_max_abs_px']:.3e} px  covis max-abs {v['covis_max_abs']:.3e}")
    if os.path.isdir(OUT):
        path = os.path.join(OUT, "precision_table.json")
        d = json.load(open(path)) if os.path.exists(path) else {}
        d[tag] = rows
        json.dump(d, open(path, "w"), indent=1)
    return rows


@needs_ref
@pytest.mark.parametrize("cfg_name", ["ufm_base_assumed", "ufm_refine_assumed"])
def test_engine_vs_reference_gpu_amplified(cfg_name):
    """560x420, full architecture, amplified head gain (flow up to tens of px, logits up to +-80) so errors are visible."""
    out, refine = _three_way(cfg_name, None, 8.0, 2)
    rows = _rows(f"{cfg_name} gain8", out)
    # the engine may not be further from the fp32 truth than 2x the reference's own GPU recipe (+ a floor)
    assert rows["E-T"]["flow_rel_l2"] <= max(2.0 * rows["R-T"]["flow_rel_l2"], 1e-2 if refine else 6e-3), rows
    assert rows["E-R"]["flow_rel_l2"] <= (2e-2 if refine else 1e-2), rows


@needs_ref
@pytest.mark.parametrize("cfg_name", ["ufm_base_assumed", "ufm_refine_assumed"])
def test_north_star_vs_reference_gpu_natural_scale(cfg_name):
    """BASELINE.json tolerances at the output scale a random-init model has: flow <= 1e-2 px max-abs and covisibility
    <= 1e-3 against the reference forward -- here both against its fp32 CPU run and against its GPU recipe."""
    out, refine = _three_way(cfg_name, None, 0.01, 1, seed=21)
    rows = _rows(f"{cfg_name} natural", out)
    # UFM-Base meets the absolute north-star bar.  On UFM-Refine the reference's OWN GPU recipe is 4.8e-2 px away from its
    # fp32 run (the 25-way softmax expectation of the local refinement amplifies the bf16 encoder noise into +-2 px
    # residual flips), so the bar there is "no worse than 1.5x what the reference does to itself".
    tol = max(1e-2, 1.5 * rows["R-T"]["flow_max_abs_px"]) if refine else 1e-2
    for k in ("E-T", "E-R"):
        assert rows[k]["flow_max_abs_px"] <= tol, (k, tol, rows)
        assert rows[k]["covis_max_abs"] <= 1e-3, (k, rows)


@needs_ref
def test_hires_full_architecture_vs_reference_gpu():
    """BASELINE.json configs[3]: UFM-Base at 1120x840 with the FULL architecture (4801-token encoder attention at D=1024,
    9600-token global attention at d=768, 840x1120 convs), one pair, reference on the GPU so it takes seconds."""
    out, _ = _three_way("ufm_base_assumed", (1120, 840), 8.0, 1, cpu_truth=False)
    rows = _rows("ufm_base_assumed 1120x840 gain8", out)
    assert out["E"][0].shape == (1, 2, 840, 1120)
    assert rows["E-R"]["flow_rel_l2"] <= 1e-2, rows
