"""The CPU oracle reproduces the reference's own outputs (fixtures made by oracle/make_golden.py from the
unmodified reference).  fp32 vs fp32 on CPU: the bar is rounding-level agreement, rel-L2 < 1e-5 (the
reference's own parity convention for fp32 is abs<1e-2 & rel<1e-3, UC/examples/models/dust3r/dust3r.py:223-230)."""
import os

import numpy as np
import pytest
import torch

from oracle import ufm_oracle as O
from oracle.golden_util import (GOLDEN_CASES, SHAPE_CASES, SHAPE_GOLDEN, case_config, make_inputs, make_shape_inputs, rel_l2,
                                stats, subsample)
from ufm_b200.weights import make_state_dict

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _run_oracle(name):
    case = GOLDEN_CASES[name]
    cfg = case_config(name)
    sd = make_state_dict(cfg, seed=0, out_gain=case["out_gain"])
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(case["batch"], H, W)
    with torch.no_grad():
        s0, s1 = O.canonicalize_inputs(src, tgt, cfg["encoder_kwargs"]["data_norm_type"], None)
        torch.manual_seed(case.get("rng_seed", 0))
        st = O.forward_stages(s0, s1, sd, cfg)
        torch.manual_seed(case.get("rng_seed", 0))
        api = O.predict_correspondences_batched(src, tgt, sd, cfg) if name.startswith("tiny") else None
    if "log_softmax" in st:
        st["log_softmax"] = st["log_softmax"].permute(0, 3, 4, 1, 2)
    return st, api


@pytest.mark.parametrize("name", list(GOLDEN_CASES))
def test_oracle_matches_reference_golden(name):
    torch.set_num_threads(len(os.sched_getaffinity(0)))
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    st, api = _run_oracle(name)
    checked = 0
    for key in g.files:
        if not key.startswith("sub:"):
            continue
        stage = key[4:]
        if stage == "api_flow":
            mine = api["flow"] if api is not None else st["flow"]  # native size: un-map is the identity map
        elif stage == "api_covis":
            mine = api["covisibility"] if api is not None else st["covis"].squeeze(1)
        elif stage == "api_flow_cov":
            mine = api["flow_covariance"] if api is not None else st["flow_cov"]
        else:
            mine = st[stage]
        ref_sub = torch.from_numpy(g[key])
        r = rel_l2(subsample(mine), ref_sub)
        tol = 1e-5 if name.startswith("tiny") or not stage.startswith("api_") else 1e-4
        assert r < tol, f"{name}:{stage} rel-L2 {r:.3e}"
        np.testing.assert_allclose(stats(mine), g["stat:" + stage], rtol=2e-4, atol=1e-6, err_msg=f"{name}:{stage}")
        checked += 1
    assert checked >= 10


def test_oracle_api_shell_on_foreign_and_different_sizes():
    """Source and target images of foreign and DIFFERENT sizes (the shapes of the reference's own example pairs, scaled
    down): the oracle's API shell (resize in, un-map out) against the unmodified reference's outputs."""
    torch.set_num_threads(len(os.sched_getaffinity(0)))
    g = np.load(os.path.join(GOLDEN, SHAPE_GOLDEN + ".npz"))
    cfg = case_config("tiny_base")
    sd = make_state_dict(cfg, seed=0, out_gain=8.0)
    for i, (h0, w0, h1, w1) in enumerate(SHAPE_CASES):
        src, tgt = make_shape_inputs(i)
        with torch.no_grad():
            api = O.predict_correspondences_batched(src, tgt, sd, cfg)
        assert api["flow"].shape == (1, 2, h0, w0) and api["covisibility"].shape == (1, h0, w0)
        for key, mine in ((f"flow{i}", api["flow"]), (f"covis{i}", api["covisibility"])):
            assert rel_l2(subsample(mine), torch.from_numpy(g["sub:" + key])) < 1e-5, key
            np.testing.assert_allclose(stats(mine), g["stat:" + key], rtol=2e-4, atol=1e-6, err_msg=key)
