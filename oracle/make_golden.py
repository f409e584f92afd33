"""Generate ``tests/golden/*.npz`` from the UNMODIFIED reference (build container only).

    python -m oracle.make_golden [case ...]

For every case in :data:`oracle.golden_util.GOLDEN_CASES` this builds the reference model from
``/root/reference`` (``oracle.ref_harness``), loads the deterministic ``state_dict``
(:func:`ufm_b200.weights.make_state_dict`), runs ``predict_correspondences_batched`` on seeded uint8
pairs on CPU fp32 and records (a) the API outputs and (b) stage-boundary tensors captured with forward
hooks, each as a strided subsample plus [mean, mean|x|, L2, max|x|] statistics of the FULL tensor.
The fixtures travel to the GPU box; the reference does not.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle.golden_util import (GOLDEN_CASES, SHAPE_CASES, SHAPE_GOLDEN, case_config, make_inputs, make_shape_inputs,  # noqa: E402
                                pack_stages)
from oracle.ref_harness import build_reference  # noqa: E402
from ufm_b200.weights import make_state_dict  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def run_reference(name: str):
    case = GOLDEN_CASES[name]
    cfg = case_config(name)
    sd = make_state_dict(cfg, seed=0, out_gain=case["out_gain"])
    model = build_reference(cfg, sd)
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(case["batch"], H, W)
    B = case["batch"]
    st = {}
    hooks = []

    def hk(mod, fn):
        hooks.append(mod.register_forward_hook(lambda m, i, o: fn(o)))

    def enc(o):
        for k, e in enumerate(o):
            st[f"enc{k}"] = e.features

    def info(o):
        final, inter = o
        st["info_final"] = torch.stack(final.features, 0)
        for k, t in enumerate(inter):
            st[f"info_inter{k}"] = torch.stack(t.features, 0)

    hk(model.encoder, enc)
    hk(model.info_sharing, info)
    hk(model.head1[0][0], lambda o: st.__setitem__("dpt1", o.features_upsampled_8x))
    hk(model.head1[0], lambda o: st.__setitem__("head1_raw", o.decoded_channels))
    hk(model.head1[1], lambda o: st.__setitem__("flow_reg", o["flow"].value))
    if hasattr(model, "uncertainty_head"):
        hk(model.uncertainty_head[0][0], lambda o: st.__setitem__("dptu", o.features_upsampled_8x))
        hk(model.uncertainty_head[0], lambda o: st.__setitem__("headu_raw", o.decoded_channels))

        def unc(o):
            if "flow_cov" in o:
                st["flow_cov"], st["flow_cov_inv"] = o["flow_cov"].covariance, o["flow_cov"].inv_covariance
                st["flow_cov_log_det"] = o["flow_cov"].log_det
            if "keypoint_confidence" in o:
                st["keypoint_confidence"] = o["keypoint_confidence"].value

        hk(model.uncertainty_head[1], unc)
    if hasattr(model, "classification_head"):
        unet = getattr(model, "use_unet_feature", False)
        hk(model.classification_head, lambda o: st.__setitem__("desc_cls" if unet else "desc", o.decoded_channels))
        if unet:
            calls = []

            def unet_out(o):  # called once per view (ufm.py:916-918)
                calls.append(o)
                st["unet_feat"] = torch.cat(calls, 0)

            hk(model.unet_feature, unet_out)
        orig = model.classification_refinement

        def wrapped(flow, feats):
            if unet:
                st["desc"] = feats  # the combined features (ufm.py:967-982)
            r, l = orig(flow, feats)
            st["residual"], st["log_softmax"] = r, l.permute(0, 3, 4, 1, 2)  # (B,P,P,H,W) so H,W trail
            return r, l

        model.classification_refinement = wrapped
    t0 = time.time()
    torch.manual_seed(case.get("rng_seed", 0))  # use_rand_idx_pe_for_non_reference_views draws from the global generator
    with torch.no_grad():
        out = model.predict_correspondences_batched(src, tgt)
    dt = time.time() - t0
    for h in hooks:
        h.remove()
    st["api_flow"] = out.flow.flow_output
    if out.covisibility is not None:
        st["api_covis"] = out.covisibility.mask
    if out.flow.flow_covariance is not None:
        st["api_flow_cov"] = out.flow.flow_covariance
    return st, dt


def run_reference_shapes():
    """The reference's API shell on source / target images of foreign and different sizes (golden_util.SHAPE_CASES)."""
    cfg = case_config("tiny_base")
    sd = make_state_dict(cfg, seed=0, out_gain=8.0)
    model = build_reference(cfg, sd)
    st = {}
    for i in range(len(SHAPE_CASES)):
        src, tgt = make_shape_inputs(i)
        with torch.no_grad():
            out = model.predict_correspondences_batched(src, tgt)
        st[f"flow{i}"] = out.flow.flow_output
        st[f"covis{i}"] = out.covisibility.mask
    return st


def main(argv):
    os.makedirs(OUT, exist_ok=True)
    names = argv or (list(GOLDEN_CASES) + [SHAPE_GOLDEN])
    torch.set_num_threads(len(os.sched_getaffinity(0)))
    if SHAPE_GOLDEN in names:
        names = [n for n in names if n != SHAPE_GOLDEN]
        path = os.path.join(OUT, SHAPE_GOLDEN + ".npz")
        st = run_reference_shapes()
        np.savez_compressed(path, **pack_stages(st))
        print(f"{SHAPE_GOLDEN}: {len(st)} outputs, {os.path.getsize(path) / 1e6:.2f} MB")
    for name in names:
        st, dt = run_reference(name)
        packed = pack_stages(st)
        packed["meta:ref_seconds"] = np.array([dt])
        path = os.path.join(OUT, name + ".npz")
        np.savez_compressed(path, **packed)
        print(f"{name}: {len(st)} stages, reference {dt:.2f}s, {os.path.getsize(path) / 1e6:.2f} MB")


if __name__ == "__main__":
    main(sys.argv[1:])
