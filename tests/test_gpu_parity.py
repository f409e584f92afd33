"""End-to-end parity on a B200, through the drop-in classes and the C ABI, against (a) the CPU fp32 oracle run on the
same seeded inputs and weights and (b) the committed golden fixtures produced by the unmodified reference.

Tolerances (BASELINE.json: "flow within 1e-2 px max-abs, or a stated relative tolerance for bf16; covisibility within
1e-3"): the engine follows the reference's GPU recipe (bf16 operands, fp32 accumulation / LayerNorm / residuals) and
additionally keeps the DPT conv operands in bf16, so against the fp32 oracle the stated tolerances are
  stage features   rel-L2 <= 2e-2     (reference fp32<->bf16 autocast itself: 6.7e-3 / 6.4e-3)
  final flow       rel-L2 <= 1.6e-2 (measured <= 1.35e-2 over all cases)
  Refine flow      rel-L2 <= 3e-2 (measured <= 2.1e-2; the 25-way softmax expectation, residual in [-2, 2] px, amplifies
                   descriptor noise -- the unmodified reference under its OWN GPU recipe is 1.4e-2 / 4.1 px max-abs away
                   from its fp32 run on ufm_refine_assumed, tests/test_gpu_vs_reference.py; the refinement kernel itself is
                   checked to 5e-4 px by feeding the oracle the engine's own descriptors)
  covisibility     logits rel-L2 <= 2.5e-2 and |mask - ref| <= 0.25 * max|logit - ref logit| (sigmoid is 1/4-Lipschitz): the
                   fixtures amplify the random-init outputs (out_gain=8, logits up to +-80) so that errors are visible.
At the natural random-init output scale (test_north_star_absolute_tolerances) the absolute north-star tolerances hold:
flow <= 1e-2 px max-abs and covisibility <= 1e-3 max-abs.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ufm_oracle as O  # noqa: E402
from oracle.golden_util import (GOLDEN_CASES, SHAPE_CASES, SHAPE_GOLDEN, case_config, make_inputs, make_shape_inputs,  # noqa: E402
                                max_abs, rel_l2, subsample)
from ufm_b200 import UniFlowMatchClassificationRefinement, UniFlowMatchConfidence, _lib  # noqa: E402
from ufm_b200.weights import make_state_dict  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
FEAT_TOL, FLOW_TOL, REFINE_FLOW_TOL = 2e-2, 1.6e-2, 3e-2
LOGIT_TOL = 2.5e-2  # raw covisibility logits (one output channel, zero-mean): measured 0.8e-2 ... 2.0e-2 over the seventeen cases


@pytest.fixture(scope="module", autouse=True)
def _built():
    import __graft_entry__ as g

    g.build()
    assert torch.cuda.is_available(), "GPU tests need a B200"
    torch.set_num_threads(len(os.sched_getaffinity(0)))


def _model(name):
    case = GOLDEN_CASES[name]
    cfg = case_config(name)
    sd = make_state_dict(cfg, seed=0, out_gain=case["out_gain"])
    cls = UniFlowMatchClassificationRefinement if "classification_head_kwargs" in cfg else UniFlowMatchConfidence
    m = cls(**cfg).to("cuda")
    m.load_state_dict(sd)
    return m, cfg, sd


def _engine_stages(model, B, h, w):
    """Engine stage buffers converted to the oracle's layouts."""
    eng = next(iter(model._engines.values()))
    out = {}

    def tok(name, key, sl=None):
        try:
            t = eng.stage(name).float()
        except Exception:
            return
        if sl is not None:
            t = t[..., :sl]
        Bn, V, n, D = t.shape
        out[key] = t.permute(1, 0, 3, 2).reshape(V, Bn, D, h, w)

    tok("enc_last", "enc_last")
    tok("info_inter0", "info_inter0")
    tok("info_inter1", "info_inter1")
    tok("info_final", "info_final")
    for k in ("dpt1", "dptu"):
        try:
            out[k] = eng.stage(k).float().permute(0, 3, 1, 2)
        except Exception:
            pass
    for k in ("head1_raw", "headu_raw", "flow_reg", "residual"):
        try:
            out[k] = eng.stage(k)
        except Exception:
            pass
    try:
        d = eng.stage("desc")
        # engine layout (B, view, C/8, H, W, 8) bf16 -> (view*B, C, H, W) fp32
        out["desc"] = d.permute(1, 0, 2, 5, 3, 4).reshape(-1, d.shape[2] * d.shape[5], d.shape[3], d.shape[4]).float()
        out["enc_first"] = eng.stage("enc_first").float()
    except Exception:
        pass
    try:
        out["log_softmax"] = eng.stage("log_softmax")
    except Exception:
        pass
    try:
        # UNet branch: (B, view, H, W, 64) bf16 / (B, view, 16, H, W) fp32 / planar descriptors -> (view*B, C, H, W)
        x0 = eng.stage("unet_x0").float()
        out["unet_x0"] = x0.permute(1, 0, 4, 2, 3).reshape(-1, x0.shape[4], x0.shape[2], x0.shape[3])
        uf = eng.stage("unet_feat")
        out["unet_feat"] = uf.permute(1, 0, 2, 3, 4).reshape(-1, *uf.shape[2:])
        d = eng.stage("desc_cls")
        out["desc_cls"] = d.permute(1, 0, 2, 5, 3, 4).reshape(-1, d.shape[2] * d.shape[5], d.shape[3], d.shape[4]).float()
    except Exception:
        pass
    return out


def _report(tag, rows):
    print(f"\n[{tag}]")
    for k, r, m, s in rows:
        print(f"  {k:14s} rel-L2 {r:.3e}  max-abs {m:.3e}  (ref max {s:.3e})")


@pytest.mark.parametrize("name", ["tiny_base", "tiny_refine", "tiny_alt_qknorm", "tiny_cross", "tiny_unet", "tiny_unet_mod", "tiny_rope", "tiny_coord", "tiny_cross_raw", "tiny_bn", "tiny_nonorm", "tiny_randidx", "tiny_hooks", "tiny_reg"])
def test_tiny_stagewise_vs_oracle(name):
    model, cfg, sd = _model(name)
    W, H = cfg["inference_resolution"]
    B = 3
    src, tgt = make_inputs(B, H, W, seed=5)
    refine = "classification_head_kwargs" in cfg
    nt = cfg["encoder_kwargs"]["data_norm_type"]
    s0, s1 = O.canonicalize_inputs(src, tgt, nt, None)
    torch.manual_seed(77)
    with torch.no_grad():
        ref = O.forward_stages(s0, s1, sd, cfg)
    torch.manual_seed(77)
    res = model.forward({"img": s0.cuda(), "data_norm_type": nt}, {"img": s1.cuda(), "data_norm_type": nt})
    torch.cuda.synchronize()
    st = _engine_stages(model, B, H // 14, W // 14)
    enc_last = torch.stack([ref[f"enc{len(cfg['encoder_kwargs']['indices']) - 1}"][:B],
                            ref[f"enc{len(cfg['encoder_kwargs']['indices']) - 1}"][B:]])
    pairs = [("enc_last", st["enc_last"], enc_last), ("info_inter0", st["info_inter0"], ref["info_inter0"]),
             ("info_inter1", st["info_inter1"], ref["info_inter1"]), ("info_final", st["info_final"], ref["info_final"]),
             ("dpt1", st["dpt1"], ref["dpt1"]), ("head1_raw", st["head1_raw"], ref["head1_raw"]),
             ("dptu", st["dptu"], ref["dptu"]), ("headu_raw", st["headu_raw"], ref["headu_raw"])]
    if refine:
        pairs.append(("desc", st["desc"], ref["desc"]))
    if cfg.get("use_unet_feature"):
        pairs += [("unet_x0", st["unet_x0"], ref["unet_x0"]), ("unet_feat", st["unet_feat"], ref["unet_feat"]),
                  ("desc_cls", st["desc_cls"], ref["desc_cls"])]
    rows = [(k, rel_l2(a.cpu(), b), max_abs(a.cpu(), b), b.abs().max().item()) for k, a, b in pairs]
    flow = res.flow.flow_output.cpu()
    rows.append(("flow", rel_l2(flow, ref["flow"]), max_abs(flow, ref["flow"]), ref["flow"].abs().max().item()))
    cov = res.covisibility.mask.cpu()
    rows.append(("covis", rel_l2(cov, ref["covis"]), max_abs(cov, ref["covis"]), 1.0))
    _report(name, rows)
    logit_err = max_abs(st["headu_raw"].cpu(), ref["headu_raw"])
    for k, r, m, s in rows:
        if k == "covis":
            assert m <= 0.25 * logit_err + 1e-5, (k, m, logit_err)
        elif k == "flow":
            assert r <= (REFINE_FLOW_TOL if refine else FLOW_TOL), (k, r)
        elif k == "headu_raw":
            assert r <= LOGIT_TOL, (k, r)
        elif k == "head1_raw":
            assert r <= FLOW_TOL, (k, r)
        else:
            assert r <= FEAT_TOL, (k, r)
    if refine:
        # the refinement itself, isolated from upstream bf16 noise: feed the oracle the ENGINE's descriptors and flow
        d = st["desc"].cpu()
        r_ref, l_ref = O.local_refinement(st["flow_reg"].cpu(), d[:B], d[B:], sd["classification_bias"],
                                          float(cfg["temperature"]), 5)
        assert max_abs(st["residual"].cpu(), r_ref) < 5e-4
        assert max_abs(st["log_softmax"].cpu().view(B, H, W, 5, 5), l_ref) < 5e-3
        assert max_abs(flow, st["flow_reg"].cpu() + r_ref) < 1e-3


@pytest.mark.parametrize("name", list(GOLDEN_CASES))
def test_api_vs_reference_golden(name):
    """predict_correspondences_batched on the golden inputs vs outputs recorded from the unmodified reference."""
    model, cfg, sd = _model(name)
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(GOLDEN_CASES[name]["batch"], H, W)
    torch.manual_seed(GOLDEN_CASES[name].get("rng_seed", 0))  # the draw of use_rand_idx_pe_for_non_reference_views
    out = model.predict_correspondences_batched(src.cuda(), tgt.cuda())
    torch.cuda.synchronize()
    flow, cov = out.flow.flow_output.cpu(), out.covisibility.mask.cpu()
    rf = torch.from_numpy(g["sub:api_flow"])
    rc = torch.from_numpy(g["sub:api_covis"])
    r, m = rel_l2(subsample(flow), rf), max_abs(subsample(flow), rf)
    mc = max_abs(subsample(cov), rc)
    eng = next(iter(model._engines.values()))
    lg, rl = subsample(eng.stage("headu_raw").cpu()), torch.from_numpy(g["sub:headu_raw"])
    rlg, mlg = rel_l2(lg, rl), max_abs(lg, rl)
    print(f"\n[{name}] flow rel-L2 {r:.3e} max-abs {m:.3e} (ref max {rf.abs().max():.2f} px); covis max-abs {mc:.3e}; "
          f"logits rel-L2 {rlg:.3e} max-abs {mlg:.3e}")
    refine = "classification_head_kwargs" in cfg
    assert r <= (REFINE_FLOW_TOL if refine else FLOW_TOL)
    if not refine:
        assert m <= FLOW_TOL * rf.abs().max().item() * 2.5
    assert rlg <= LOGIT_TOL
    assert mc <= 0.25 * mlg + 1e-5
    assert out.covisibility.logits is None and flow.shape == (src.shape[0], 2, H, W)


@pytest.mark.parametrize("host", [False, True])
def test_api_foreign_and_different_sizes_vs_reference_golden(host):
    """Source and target images of foreign and DIFFERENT sizes (the shapes of the reference's example pairs at 1/5 scale;
    base.py:253-334): prep_resize_kernel in, unmap_kernel out, against outputs recorded from the unmodified reference
    (tests/golden/tiny_shapes.npz).  The covisibility of a head with gain 8 is compared through the mean: single pixels near
    a logit zero crossing flip by bf16 noise."""
    model, cfg, sd = _model("tiny_base")
    g = np.load(os.path.join(GOLDEN, SHAPE_GOLDEN + ".npz"))
    for i, (h0, w0, h1, w1) in enumerate(SHAPE_CASES):
        src, tgt = make_shape_inputs(i)
        out = model.predict_correspondences_batched(src if host else src.cuda(), tgt if host else tgt.cuda())
        flow, cov = out.flow.flow_output, out.covisibility.mask
        assert flow.shape == (1, 2, h0, w0) and cov.shape == (1, h0, w0)
        assert flow.device.type == ("cpu" if host else "cuda")
        rf, rc = torch.from_numpy(g[f"sub:flow{i}"]), torch.from_numpy(g[f"sub:covis{i}"])
        r = rel_l2(subsample(flow.cpu()), rf)
        mc = (subsample(cov.cpu()) - rc).abs().mean().item()
        print(f"\n[shapes {h0}x{w0} / {h1}x{w1}] flow rel-L2 {r:.3e} (ref max {rf.abs().max():.1f} px); covis mean-abs {mc:.3e}")
        assert r <= FLOW_TOL
        assert mc <= 1e-2


def test_api_contract_shapes_layouts_errors():
    model, cfg, sd = _model("tiny_base")
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(2, H, W, seed=9)
    a = model.predict_correspondences_batched(src.cuda(), tgt.cuda()).flow.flow_output
    # BCHW uint8, unbatched HWC, CPU tensors in -> CPU tensors out
    b = model.predict_correspondences_batched(src.permute(0, 3, 1, 2).contiguous().cuda(),
                                              tgt.permute(0, 3, 1, 2).contiguous().cuda()).flow.flow_output
    assert torch.equal(a, b)
    c = model.predict_correspondences_batched(src[0], tgt[0])
    assert c.flow.flow_output.device.type == "cpu" and c.flow.flow_output.shape == (1, 2, H, W)
    assert torch.allclose(c.flow.flow_output, a[:1].cpu(), atol=1e-5)
    # float32 input with the encoder's normalisation == uint8 path up to bf16 rounding of the patch operand
    s0, s1 = O.canonicalize_inputs(src, tgt, "dinov2", None)
    d = model.predict_correspondences_batched(s0.cuda(), s1.cuda(), data_norm_type="dinov2").flow.flow_output
    assert rel_l2(d.cpu(), a.cpu()) < 1e-5
    with pytest.raises(AssertionError):
        model.predict_correspondences_batched(s0.cuda(), s1.cuda())
    with pytest.raises(ValueError):
        model.predict_correspondences_batched(torch.zeros(1, 5, 8, 4).cuda(), torch.zeros(1, 5, 8, 4).cuda())
    with pytest.raises(ValueError):
        model.predict_correspondences_batched(src.cuda().half(), tgt.cuda().half())
    with pytest.raises(RuntimeError):  # B = 0: the reference raises (reshape of 0 elements), so does the engine shell
        model.predict_correspondences_batched(src[:0].cuda(), tgt[:0].cuda())
    # non-native input size: resized in, un-mapped out, compared with the oracle's shell
    g = torch.Generator().manual_seed(3)
    s2 = torch.randint(0, 256, (300, 400, 3), dtype=torch.uint8, generator=g)
    t2 = torch.randint(0, 256, (300, 400, 3), dtype=torch.uint8, generator=g)
    o = model.predict_correspondences_batched(s2.cuda(), t2.cuda())
    with torch.no_grad():
        ref = O.predict_correspondences_batched(s2, t2, sd, cfg)
    assert o.flow.flow_output.shape == (1, 2, 300, 400) and o.covisibility.mask.shape == (1, 300, 400)
    assert rel_l2(o.flow.flow_output.cpu(), ref["flow"]) < FLOW_TOL
    assert (o.covisibility.mask.cpu() - ref["covisibility"]).abs().mean().item() < 1e-2
    # up-scaled input, BCHW fp32 normalised for another model family (re-normalisation path), batch of 2
    s3 = torch.randint(0, 256, (2, 3, 40, 52), dtype=torch.uint8, generator=g)
    t3 = torch.randint(0, 256, (2, 3, 40, 52), dtype=torch.uint8, generator=g)
    f3 = lambda u: (u.float() / 255.0 - 0.5) / 0.5
    o3 = model.predict_correspondences_batched(f3(s3).cuda(), f3(t3).cuda(), data_norm_type="dust3r")
    with torch.no_grad():
        ref3 = O.predict_correspondences_batched(f3(s3), f3(t3), sd, cfg, data_norm_type="dust3r")
    assert o3.flow.flow_output.shape == (2, 2, 40, 52)
    assert rel_l2(o3.flow.flow_output.cpu(), ref3["flow"]) < FLOW_TOL


def test_batch_chunking_and_determinism():
    model, cfg, sd = _model("tiny_base")
    model.max_pairs_per_call = 2
    model._engines = {}
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(5, H, W, seed=11)
    a = model.predict_correspondences_batched(src.cuda(), tgt.cuda()).flow.flow_output
    b = model.predict_correspondences_batched(src.cuda(), tgt.cuda()).flow.flow_output
    assert torch.equal(a, b)
    one = model.predict_correspondences_batched(src[4:5].cuda(), tgt[4:5].cuda()).flow.flow_output
    assert torch.equal(one, a[4:5])  # pairs are independent: batching must not change a pair's result
    # host tensors (pageable and pinned) take the ufm_forward_host path: same bits, results on the host
    for hs, ht in ((src, tgt), (src.pin_memory(), tgt.pin_memory())):
        o = model.predict_correspondences_batched(hs, ht)
        assert o.flow.flow_output.device.type == "cpu" and o.covisibility.mask.device.type == "cpu"
        assert torch.equal(o.flow.flow_output, a.cpu())


@pytest.mark.parametrize("layout", ["bhwc", "bchw"])
def test_patch_im2col_u8_kernel_matches_generic(layout, monkeypatch):
    """uint8 images take the table-driven one-warp-per-patch gather (simt.cu patch_im2col_u8_kernel); UFM_PATCH_GENERIC=1
    routes them through the generic kernel.  Same normalisation arithmetic, so the outputs of a pass are bit-identical."""
    model, cfg, sd = _model("tiny_base")
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(3, H, W, seed=5)
    if layout == "bchw":
        src, tgt = src.permute(0, 3, 1, 2).contiguous(), tgt.permute(0, 3, 1, 2).contiguous()
    assert src.dtype == torch.uint8
    outs = []
    for generic in (False, True):
        if generic:
            monkeypatch.setenv("UFM_PATCH_GENERIC", "1")
        else:
            monkeypatch.delenv("UFM_PATCH_GENERIC", raising=False)
        o = model.predict_correspondences_batched(src.cuda(), tgt.cuda())
        outs.append((o.flow.flow_output.clone(), o.covisibility.mask.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize("name", ["tiny_base", "tiny_refine", "tiny_unet", "tiny_cov"])
def test_branches_bit_identical(name):
    """Passes of <= 4 pairs run the uncertainty head, the classification head (+ UNet branch) and the input chains of each
    DPT head on side streams (ufm_debug_flag 8, engine.cu `branches`): the same kernels on the same operands, so device
    and host paths (plain pass, graph capture, graph replay) must give the bits of the one-stream pass."""
    lib = _lib.load()
    cfg = case_config(name)
    sd = make_state_dict(cfg, seed=0, out_gain=1.0)
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(3, H, W, seed=13)
    cls = UniFlowMatchClassificationRefinement if "classification_head_kwargs" in cfg else UniFlowMatchConfidence
    res = {}
    try:
        for mode in (0, 1):
            assert lib.ufm_debug_flag(8, mode) == 0
            m = cls(max_pairs_per_call=4, **cfg).to("cuda")
            m.load_state_dict(sd)
            outs = []
            for pairs in (1, 3):
                for rep in range(3):  # plain pass + capture, then two replays
                    o = m.predict_correspondences_batched(src[:pairs].cuda(), tgt[:pairs].cuda())
                    outs.append((o.flow.flow_output.clone(), o.covisibility.mask.clone()))
                oh = m.predict_correspondences_batched(src[:pairs], tgt[:pairs])  # host path
                outs.append((oh.flow.flow_output.cuda(), oh.covisibility.mask.cuda()))
            torch.cuda.synchronize()
            res[mode] = outs
            del m
    finally:
        lib.ufm_debug_flag(8, 1)
    for (f0, c0), (f1, c1) in zip(res[0], res[1]):
        assert torch.isfinite(f1).all()
        assert torch.equal(f0, f1) and torch.equal(c0, c1)


@pytest.mark.parametrize("name", ["tiny_base", "ufm_base_assumed"])
def test_north_star_absolute_tolerances(name):
    """BASELINE.json: flow within 1e-2 px max-abs and covisibility within 1e-3 of the reference forward, at the output
    scale a random-init model has (flow spans +-0.07 px) -- out_gain=0.01 brings the seeded weights there."""
    cfg = case_config(name)
    sd = make_state_dict(cfg, seed=0, out_gain=0.01)
    m = UniFlowMatchConfidence(**cfg).to("cuda")
    m.load_state_dict(sd)
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(1, H, W, seed=21)
    out = m.predict_correspondences_batched(src.cuda(), tgt.cuda())
    torch.cuda.synchronize()
    with torch.no_grad():
        ref = O.predict_correspondences_batched(src, tgt, sd, cfg)
    mf = max_abs(out.flow.flow_output.cpu(), ref["flow"])
    mc = max_abs(out.covisibility.mask.cpu(), ref["covisibility"])
    print(f"\n[{name} natural scale] flow max-abs {mf:.3e} px (ref max {ref['flow'].abs().max():.3f}); covis max-abs {mc:.3e}")
    assert mf <= 1e-2 and mc <= 1e-3


@pytest.mark.parametrize("conf_type", ["sigmoid", "exp", "softmax"])
def test_uncertainty_adaptors(conf_type):
    """Covariance2DAdaptor + ConfidenceAdaptor + MaskAdaptor on one uncertainty head (adaptors.py:899-985,
    consumed at ufm.py:644-660, covariance rescaled at base.py:295-319)."""
    from ufm_b200.configs import tiny

    cfg = tiny(False, unc_adaptors="cov_conf_mask", conf_type=conf_type)
    sd = make_state_dict(cfg, seed=0, out_gain=1.0)
    m = UniFlowMatchConfidence(**cfg).to("cuda")
    m.load_state_dict(sd)
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(3, H, W, seed=5)
    m.max_pairs_per_call = 2  # two chunks: the per-chunk stage read-back must concatenate
    s0, s1 = O.canonicalize_inputs(src, tgt, "dinov2", None)
    res = m.forward({"img": s0.cuda(), "data_norm_type": "dinov2", "symmetrized": False},
                    {"img": s1.cuda(), "data_norm_type": "dinov2", "symmetrized": False})
    torch.cuda.synchronize()
    assert res.flow.flow_covariance.shape == (3, 3, H, W) and res.keypoint_confidence.shape == (3, H, W)
    assert res.covisibility.mask.shape == (3, 1, H, W)
    # adaptor arithmetic in isolation: the oracle's AdaptorMap on the ENGINE's raw head output (last chunk)
    eng = next(iter(m._engines.values()))
    raw = eng.stage("headu_raw").cpu()
    au = O.run_adaptors(raw, cfg["uncertainty_adaptors_kwargs"], (H, W))
    cov, log_det, inv = au["flow_cov"]
    n = raw.shape[0]
    # element-wise error relative to |ref| + a floor (the off-diagonal terms cross zero)
    rel = lambda a, b: ((a - b).abs() / (b.abs() + 1e-2 * b.abs().mean())).max().item()
    assert rel(res.flow.flow_covariance[-n:].cpu(), cov) < 1e-4
    # 1 / (1 - tanh^2 + 1e-8) is ill-conditioned near |tanh| = 1: an error of a few ulp of tanhf (device tanhf vs the
    # host's) is amplified by 2 |tanh| / (1 - tanh^2).  The bound is therefore per element: well-conditioned pixels must
    # agree to 2e-3, the others to 8 ulp of tanh times their own amplification.
    t = raw[:, 2:3].double().tanh()
    amp = (2 * t.abs() / (1 - t * t + 1e-8)).float()
    bound = 2e-3 + 8 * 6e-8 * amp
    got_inv = res.flow.flow_covariance_inv[-n:].cpu()
    err_inv = (got_inv - inv).abs() / (inv.abs() + 1e-2 * inv.abs().mean())
    assert (err_inv <= bound).all(), (err_inv / bound).max().item()
    err_ld = (res.flow.flow_covariance_log_det[-n:].cpu() - log_det).abs()
    assert (err_ld <= bound).all(), (err_ld / bound).max().item()  # log of the same 1 - tanh^2 term
    assert rel(res.keypoint_confidence[-n:].cpu(), au["keypoint_confidence"].squeeze(1)) < 2e-4
    assert max_abs(res.covisibility.mask[-n:].cpu(), au["non_occluded_mask"]["mask"]) < 1e-6
    # end to end against the fp32 oracle, native and non-native size (covariance is nearest-resampled and rescaled)
    with torch.no_grad():
        ref = O.predict_correspondences_batched(src, tgt, sd, cfg)
    out = m.predict_correspondences_batched(src.cuda(), tgt.cuda())
    assert rel_l2(out.flow.flow_covariance.cpu(), ref["flow_covariance"]) < 5e-2
    g = torch.Generator().manual_seed(4)
    s2 = torch.randint(0, 256, (2, 120, 150, 3), dtype=torch.uint8, generator=g)
    t2 = torch.randint(0, 256, (2, 120, 150, 3), dtype=torch.uint8, generator=g)
    with torch.no_grad():
        ref2 = O.predict_correspondences_batched(s2, t2, sd, cfg)
    out2 = m.predict_correspondences_batched(s2.cuda(), t2.cuda())
    assert out2.flow.flow_covariance.shape == (2, 3, 120, 150)
    assert rel_l2(out2.flow.flow_covariance.cpu(), ref2["flow_covariance"]) < 5e-2
    assert rel_l2(out2.flow.flow_output.cpu(), ref2["flow"]) < FLOW_TOL


def test_unet_branch_full_resolution():
    """SURVEY.md 8f N3 at 560x420 (105 -> 104 -> nearest-resized 105 rows on the third level): the UNet branch and the
    combine against the oracle, isolated from upstream bf16 noise by feeding the oracle the engine's own classification
    features."""
    from ufm_b200 import get_config

    cfg = get_config("ufm_refine_unet")
    sd = make_state_dict(cfg, seed=0, out_gain=8.0)
    model = UniFlowMatchClassificationRefinement(**cfg).to("cuda")
    model.load_state_dict(sd)
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(1, H, W, seed=3)
    s0, s1 = O.canonicalize_inputs(src, tgt, "dinov2", None)
    model.forward({"img": s0.cuda(), "data_norm_type": "dinov2"}, {"img": s1.cuda(), "data_norm_type": "dinov2"})
    torch.cuda.synchronize()
    st = _engine_stages(model, 1, H // 14, W // 14)
    with torch.no_grad():
        uf, x0 = O.unet(torch.cat([s0, s1], dim=0), sd, return_last=True)
        desc_ref = O.combine_unet(st["desc_cls"].cpu(), uf, sd, "conv")
    rows = [(k, rel_l2(a.cpu(), b), max_abs(a.cpu(), b), b.abs().max().item())
            for k, a, b in (("unet_x0", st["unet_x0"], x0), ("unet_feat", st["unet_feat"], uf), ("desc", st["desc"], desc_ref))]
    _report("ufm_refine_unet", rows)
    for k, r, m, s in rows:
        assert r <= FEAT_TOL, (k, r)


def test_tiny_hires_vs_oracle():
    """BASELINE.json configs[3] geometry (1120x840: 4801-token encoder attention, 9600-token global attention, convs up
    to 840x1120) on the slim architecture, so the CPU oracle finishes in seconds: every kernel at its largest tiling."""
    from ufm_b200.configs import tiny

    cfg = tiny(False, inference_resolution=(1120, 840))
    sd = make_state_dict(cfg, seed=0, out_gain=8.0)
    model = UniFlowMatchConfidence(max_pairs_per_call=1, **cfg).to("cuda")
    model.load_state_dict(sd)
    src, tgt = make_inputs(1, 840, 1120, seed=9)
    s0, s1 = O.canonicalize_inputs(src, tgt, "dinov2", None)
    with torch.no_grad():
        ref = O.forward_stages(s0, s1, sd, cfg)
    out = model.predict_correspondences_batched(src.cuda(), tgt.cuda())
    torch.cuda.synchronize()
    eng = next(iter(model._engines.values()))
    rows = []
    for k in ("head1_raw", "headu_raw"):
        a = eng.stage(k).cpu()
        rows.append((k, rel_l2(a, ref[k]), max_abs(a, ref[k]), ref[k].abs().max().item()))
    flow = out.flow.flow_output.cpu()
    rows.append(("flow", rel_l2(flow, ref["flow"]), max_abs(flow, ref["flow"]), ref["flow"].abs().max().item()))
    _report("tiny@1120x840", rows)
    for k, r, m, s in rows:
        assert r <= FLOW_TOL, (k, r)
    cov = out.covisibility.mask.cpu()
    assert max_abs(cov, ref["covis"].squeeze(1)) <= 0.25 * rows[1][2] + 1e-5


def test_multiple_inference_resolutions():
    """base.py:86-98: a list of trained resolutions; AutomaticShapeSelection picks the closest aspect (first on ties).
    One engine per resolution behind the same model object; checked against the oracle's API shell."""
    from ufm_b200.configs import tiny

    cfg = tiny(False)
    cfg["inference_resolution"] = [[224, 168], [168, 224]]  # (W, H): landscape and portrait
    sd = make_state_dict(cfg, seed=0, out_gain=8.0)
    model = UniFlowMatchConfidence(**cfg).to("cuda")
    model.load_state_dict(sd)
    for (H0, W0) in [(168, 224), (224, 168), (300, 200), (120, 200)]:
        src, tgt = make_inputs(2, H0, W0, seed=H0)
        with torch.no_grad():
            ref = O.predict_correspondences_batched(src, tgt, sd, cfg)
        out = model.predict_correspondences_batched(src.cuda(), tgt.cuda())
        flow = out.flow.flow_output.cpu()
        assert flow.shape == (2, 2, H0, W0)
        r = rel_l2(flow, ref["flow"])
        print(f"\n[multi-res {H0}x{W0}] flow rel-L2 {r:.3e}")
        assert r <= FLOW_TOL, (H0, W0, r)
        assert (out.covisibility.mask.cpu() - ref["covisibility"]).abs().mean().item() < 1e-2
    assert len(model._engines) == 2


def test_ingested_checkpoint_matches_oracle(tmp_path):
    """SURVEY.md 8f N4 on the GPU: a checkpoint written by the UNMODIFIED reference class through its hub mixin
    (config.json + model.safetensors, ufm.py:120) -- or by save_pretrained when no reference tree is staged -- is read by
    from_pretrained and must produce the oracle's output for those weights.  The checkpoint carries FlowAdaptor buffers
    that differ from the constructor kwargs (adaptors.py:70-71): the buffers must win."""
    from oracle import ref_harness as R

    cfg = case_config("tiny_refine")
    sd = make_state_dict(cfg, seed=4, out_gain=8.0)
    sd["head1.1.adaptors.flow.flow_mean"] = torch.tensor([1.5, -0.75]).view(1, 2, 1, 1)
    sd["head1.1.adaptors.flow.flow_std"] = torch.tensor([1.25, 0.5]).view(1, 2, 1, 1)
    path = str(tmp_path / "ckpt")
    if R.reference_available():
        R.build_reference(cfg, sd).save_pretrained(path)
    else:
        m0 = UniFlowMatchClassificationRefinement(**cfg)
        m0.load_state_dict(sd)
        m0.save_pretrained(path)
    model = UniFlowMatchClassificationRefinement.from_pretrained(path).to("cuda")
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(2, H, W, seed=13)
    out = model.predict_correspondences_batched(src.cuda(), tgt.cuda())
    torch.cuda.synchronize()
    with torch.no_grad():
        ref = O.predict_correspondences_batched(src, tgt, sd, cfg)
        plain = O.predict_correspondences_batched(src, tgt, {k: v for k, v in sd.items() if "adaptors.flow" not in k}, cfg)
    r = rel_l2(out.flow.flow_output.cpu(), ref["flow"])
    print(f"\n[ingested checkpoint] flow rel-L2 {r:.3e} (kwargs-only affine would be {rel_l2(plain['flow'], ref['flow']):.3e} off)")
    assert r <= REFINE_FLOW_TOL
    assert rel_l2(plain["flow"], ref["flow"]) > 0.1  # the perturbed buffers matter
    assert (out.covisibility.mask.cpu() - ref["covisibility"]).abs().mean().item() < 1e-2
