"""Checkpoint ingestion (uniflowmatch/models/ufm.py:85-117, 197-240, 1254-1256): CPU only -- the model
classes hold the reference-named tensors as buffers, an engine is only built when a forward is requested."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from ufm_b200 import UniFlowMatchClassificationRefinement, UniFlowMatchConfidence, get_config  # noqa: E402
from ufm_b200.models import modify_state_dict  # noqa: E402
from ufm_b200.weights import make_state_dict  # noqa: E402


def _same(a, b):
    assert set(a) == set(b)
    for k in a:
        assert torch.equal(a[k], b[k]), k


def test_modify_state_dict_semantics():
    sd = {"model.a.w": 1, "feature_matching_proj.w": 2, "encoder.model.mask_token": 3, "x.old.y": 4}
    out = modify_state_dict(sd, {"feature_matching_proj": None, "encoder.model.mask_token": None, "old": "new", "x.": "z."})
    assert out == {"model.a.w": 1, "x.new.y": 4}  # first matching rule wins, None drops


@pytest.mark.parametrize("name,cls", [("tiny_base", UniFlowMatchConfidence), ("tiny_refine", UniFlowMatchClassificationRefinement)])
def test_save_and_from_pretrained_round_trip(tmp_path, name, cls):
    cfg = get_config(name)
    m = cls(**cfg)
    sd = make_state_dict(cfg, seed=3)
    m.load_state_dict(sd)
    m.save_pretrained(str(tmp_path / "ckpt"))
    assert sorted(os.listdir(tmp_path / "ckpt")) == ["config.json", "model.safetensors"]
    m2 = cls.from_pretrained(str(tmp_path / "ckpt"))
    _same(m2.state_dict(), sd)
    assert m2.config["encoder_kwargs"] == m.config["encoder_kwargs"]
    with pytest.raises(ValueError):
        cls.from_pretrained("infinity1096/UFM-Base-not-in-the-local-cache")


def test_from_pretrained_ckpt_and_backbone_checkpoint(tmp_path):
    cfg = get_config("tiny_base")
    sd = make_state_dict(cfg, seed=5)
    p = str(tmp_path / "export.pth")
    torch.save({"model_args": cfg, "model": sd}, p)
    m = UniFlowMatchConfidence.from_pretrained_ckpt(p)
    _same(m.state_dict(), sd)
    with pytest.raises(ValueError):
        UniFlowMatchConfidence.from_pretrained_ckpt(str(tmp_path / "missing.pth"))
    # training checkpoint of the backbone: "model." prefix, feature-matching projection and mask token are dropped
    backbone = {k: v for k, v in sd.items() if k.split(".")[0] in ("encoder", "info_sharing", "head1")}
    train = {"model." + k: v for k, v in backbone.items()}
    train["model.feature_matching_proj.weight"] = torch.zeros(3)
    train["model.encoder.model.mask_token"] = torch.zeros(1, 384)
    train["optimizer.state"] = torch.zeros(1)
    pt = str(tmp_path / "train.ckpt")
    torch.save({"state_dict": train}, pt)
    m2 = UniFlowMatchConfidence(pretrained_backbone_checkpoint_path=pt, seed=9, **cfg)
    got = m2.state_dict()
    for k, v in backbone.items():
        assert torch.equal(got[k], v), k
    fresh = make_state_dict(cfg, seed=9)
    assert torch.equal(got["uncertainty_head.0.1.conv1.weight"], fresh["uncertainty_head.0.1.conv1.weight"])
    # exported backbone checkpoint with a missing key -> assertion like ufm.py:217
    bad = dict(backbone)
    bad.pop("head1.0.1.conv1.weight")
    pb = str(tmp_path / "bad.pth")
    torch.save({"model": bad}, pb)
    with pytest.raises(AssertionError):
        UniFlowMatchConfidence(pretrained_backbone_checkpoint_path=pb, **cfg)
    with pytest.raises(AssertionError):
        UniFlowMatchConfidence(pretrained_checkpoint_path=pb, **cfg)  # ufm.py:558


def _has_reference():
    from oracle.ref_harness import reference_available

    return reference_available()


def test_flow_adaptor_buffers_win_over_kwargs():
    """FlowAdaptor.flow_mean / flow_std are registered buffers (adaptors.py:70-71): the engine's affine comes from the
    state_dict, the constructor kwargs are only the fallback."""
    from ufm_b200.packing import engine_config

    cfg = get_config("tiny_base")
    sd = make_state_dict(cfg, seed=0)
    c0 = engine_config(cfg, sd)
    assert (c0.flow_scale[0], c0.flow_scale[1], c0.flow_shift[0], c0.flow_shift[1]) == (1.0, 1.0, 0.0, 0.0)
    sd["head1.1.adaptors.flow.flow_mean"] = torch.tensor([3.0, -2.0]).view(1, 2, 1, 1)
    sd["head1.1.adaptors.flow.flow_std"] = torch.tensor([2.0, 0.5]).view(1, 2, 1, 1)
    c1 = engine_config(cfg, sd)
    assert (c1.flow_scale[0], c1.flow_scale[1], c1.flow_shift[0], c1.flow_shift[1]) == (2.0, 0.5, 3.0, -2.0)
    c2 = engine_config(cfg)
    assert (c2.flow_scale[0], c2.flow_shift[0]) == (1.0, 0.0)


def test_config_rejections_are_loud():
    """Configurations the engine does not implement raise at construction instead of loading and mis-computing."""
    from ufm_b200.packing import engine_config

    cfg = get_config("tiny_base")
    cfg["feature_head_kwargs"]["dpt_feature"]["nonlinearity"] = "gelu"
    with pytest.raises(NotImplementedError):
        engine_config(cfg)
    cfg = get_config("tiny_base")
    cfg["encoder_kwargs"]["data_norm_type"] = "no_such_norm"
    with pytest.raises(ValueError):
        engine_config(cfg)
    c = engine_config(get_config("tiny_nonorm"))
    assert c.enc_norm_intermediate == 0 and c.info_norm_intermediate == 0 and abs(c.img_mean[0] - 0.5) < 1e-7
    c = engine_config(get_config("tiny_base"))
    assert c.enc_norm_intermediate == 1 and c.info_norm_intermediate == 1 and abs(c.img_std[2] - 0.225) < 1e-7


@pytest.mark.skipif(not _has_reference(), reason="no reference tree (/root/reference or the staged oracle/_ref)")
def test_from_pretrained_reads_what_the_reference_saves(tmp_path):
    """The unmodified reference class (hub mixin) writes config.json + model.safetensors; ours must read it back."""
    from oracle.ref_harness import build_reference

    cfg = get_config("tiny_base")
    sd = make_state_dict(cfg, seed=0)
    ref = build_reference(cfg, sd)
    ref.save_pretrained(str(tmp_path / "hub"))
    m = UniFlowMatchConfidence.from_pretrained(str(tmp_path / "hub"))
    _same(m.state_dict(), sd)
