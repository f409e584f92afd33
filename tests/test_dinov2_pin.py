"""Pin the restated DINOv2 arithmetic (oracle/ref_harness.py::HubDinoVisionTransformer) against an
INDEPENDENT implementation: ``transformers.Dinov2Model`` (installed in this image; weights copied across).

The reference loads ``facebookresearch/dinov2`` through ``torch.hub`` (UC/models/encoders/dinov2.py:85-91), which is
neither vendored nor pinned, so oracle, reference harness and engine all rest on the restatement.  This test removes
"from memory" from everything except one constant:
  * 518x518 (the native grid, no interpolation): patch embedding, cls token, position embedding, every block
    (LayerNorm eps, fused qkv split order, softmax scale, LayerScale, GELU MLP) and the final norm agree to fp32 rounding;
  * 420x560 (the UFM resolution, a NON-square 30x40 grid): the bicubic, non-antialiased position-embedding
    interpolation, its (rows, cols) ordering and the cls-token handling agree with transformers' implementation when the
    restatement's ``interpolate_offset`` is 0 (transformers resamples to the exact grid size, which is the same map);
  * what stays unpinned is upstream's ``interpolate_offset = 0.1`` itself (scale_factor = (n + 0.1) / 37): its effect on
    the features is measured and bounded here.
"""
import pytest
import torch

transformers = pytest.importorskip("transformers")

from oracle.ref_harness import HubDinoVisionTransformer  # noqa: E402

DIM, DEPTH, HEADS = 384, 12, 6  # dinov2_vits14


def _pair(seed=0):
    from transformers import Dinov2Config, Dinov2Model

    torch.manual_seed(seed)
    hub = HubDinoVisionTransformer(DIM, DEPTH, HEADS).eval()
    with torch.no_grad():
        for p in hub.parameters():  # non-trivial everything (pos_embed, cls, LayerScale, biases, norms)
            p.copy_(torch.randn_like(p) * 0.05)
        for n, p in hub.named_parameters():
            if n.endswith("norm1.weight") or n.endswith("norm2.weight") or n == "norm.weight" or n.endswith("gamma"):
                p.add_(1.0)
    cfg = Dinov2Config(hidden_size=DIM, num_hidden_layers=DEPTH, num_attention_heads=HEADS, mlp_ratio=4, patch_size=14,
                       image_size=518, layerscale_value=1.0, layer_norm_eps=1e-6, hidden_act="gelu", qkv_bias=True,
                       use_swiglu_ffn=False, attn_implementation="eager")
    hf = Dinov2Model(cfg).eval()
    sd = hub.state_dict()
    m = {"embeddings.cls_token": sd["cls_token"], "embeddings.mask_token": sd["mask_token"],
         "embeddings.position_embeddings": sd["pos_embed"],
         "embeddings.patch_embeddings.projection.weight": sd["patch_embed.proj.weight"],
         "embeddings.patch_embeddings.projection.bias": sd["patch_embed.proj.bias"],
         "layernorm.weight": sd["norm.weight"], "layernorm.bias": sd["norm.bias"]}
    for i in range(DEPTH):
        a, b = f"encoder.layer.{i}.", f"blocks.{i}."
        qw, qb = sd[b + "attn.qkv.weight"], sd[b + "attn.qkv.bias"]
        for j, nm in enumerate(("query", "key", "value")):  # fused qkv rows are ordered q | k | v
            m[a + f"attention.attention.{nm}.weight"] = qw[j * DIM:(j + 1) * DIM]
            m[a + f"attention.attention.{nm}.bias"] = qb[j * DIM:(j + 1) * DIM]
        m[a + "attention.output.dense.weight"], m[a + "attention.output.dense.bias"] = sd[b + "attn.proj.weight"], sd[b + "attn.proj.bias"]
        for k in ("norm1", "norm2"):
            m[a + k + ".weight"], m[a + k + ".bias"] = sd[b + k + ".weight"], sd[b + k + ".bias"]
        m[a + "layer_scale1.lambda1"], m[a + "layer_scale2.lambda1"] = sd[b + "ls1.gamma"], sd[b + "ls2.gamma"]
        for k in ("fc1", "fc2"):
            m[a + f"mlp.{k}.weight"], m[a + f"mlp.{k}.bias"] = sd[b + f"mlp.{k}.weight"], sd[b + f"mlp.{k}.bias"]
    missing, unexpected = hf.load_state_dict(m, strict=False)
    assert not unexpected and not missing, (missing, unexpected)
    return hub, hf


def _rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm()).item()


def test_blocks_and_embeddings_at_native_grid():
    hub, hf = _pair()
    x = torch.randn(1, 3, 518, 518, generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        a = hub.forward_features(x)
        b = hf(pixel_values=x).last_hidden_state
    assert _rel(a["x_norm_patchtokens"], b[:, 1:]) < 1e-5
    assert _rel(a["x_norm_clstoken"], b[:, 0]) < 1e-5
    with torch.no_grad():  # the intermediate-layer path the reference wrapper uses (dinov2.py:259-261)
        inter = hub.get_intermediate_layers(x, n=[5, 11], reshape=True, norm=True)
        hs = hf(pixel_values=x, output_hidden_states=True).hidden_states
    for t, k in zip(inter, (6, 12)):  # hidden_states[k] = output of block k-1, before the final norm
        ref = hf.layernorm(hs[k])[:, 1:].reshape(1, 37, 37, DIM).permute(0, 3, 1, 2)
        assert _rel(t, ref) < 1e-5


def test_position_embedding_interpolation_on_the_ufm_grid():
    hub, hf = _pair(seed=3)
    x = torch.randn(2, 3, 420, 560, generator=torch.Generator().manual_seed(2))  # 30 x 40 patches
    with torch.no_grad():
        ref = hf(pixel_values=x, interpolate_pos_encoding=True).last_hidden_state[:, 1:]
        hub.interpolate_offset = 0.0
        same_map = hub.forward_features(x)["x_norm_patchtokens"]
        hub.interpolate_offset = 0.1
        upstream = hub.forward_features(x)["x_norm_patchtokens"]
    # identical interpolation map (exact grid size): everything but the offset constant is pinned
    assert _rel(same_map, ref) < 1e-5
    # the 0.1 offset of upstream's hub models moves the features only slightly; this is the unpinned residue
    d = _rel(upstream, ref)
    print(f"\n[dinov2 pin] effect of interpolate_offset=0.1 on the final features at 420x560: rel-L2 {d:.3e}")
    assert 0.0 < d < 5e-2


def test_register_variant_against_transformers():
    """hub dinov2_vit*14_reg (with_registers=True, UC dinov2.py:24,49,74-79): 4 register tokens between cls and the patch
    tokens, no positional embedding on them, interpolate_antialias=True and interpolate_offset=0 -- against
    ``transformers.Dinov2WithRegistersModel`` at the native grid and on the 30 x 40 UFM grid."""
    from transformers import Dinov2WithRegistersConfig, Dinov2WithRegistersModel

    torch.manual_seed(5)
    hub = HubDinoVisionTransformer(DIM, 4, HEADS, num_register_tokens=4).eval()
    with torch.no_grad():
        for n, p in hub.named_parameters():
            p.copy_(torch.randn_like(p) * 0.05)
            if n.endswith("norm1.weight") or n.endswith("norm2.weight") or n == "norm.weight" or n.endswith("gamma"):
                p.add_(1.0)
    cfg = Dinov2WithRegistersConfig(hidden_size=DIM, num_hidden_layers=4, num_attention_heads=HEADS, mlp_ratio=4, patch_size=14,
                                    image_size=518, layerscale_value=1.0, layer_norm_eps=1e-6, hidden_act="gelu", qkv_bias=True,
                                    use_swiglu_ffn=False, num_register_tokens=4, attn_implementation="eager")
    hf = Dinov2WithRegistersModel(cfg).eval()
    sd = hub.state_dict()
    m = {"embeddings.cls_token": sd["cls_token"], "embeddings.mask_token": sd["mask_token"],
         "embeddings.register_tokens": sd["register_tokens"], "embeddings.position_embeddings": sd["pos_embed"],
         "embeddings.patch_embeddings.projection.weight": sd["patch_embed.proj.weight"],
         "embeddings.patch_embeddings.projection.bias": sd["patch_embed.proj.bias"],
         "layernorm.weight": sd["norm.weight"], "layernorm.bias": sd["norm.bias"]}
    for i in range(4):
        a, b = f"encoder.layer.{i}.", f"blocks.{i}."
        qw, qb = sd[b + "attn.qkv.weight"], sd[b + "attn.qkv.bias"]
        for j, nm in enumerate(("query", "key", "value")):
            m[a + f"attention.attention.{nm}.weight"] = qw[j * DIM:(j + 1) * DIM]
            m[a + f"attention.attention.{nm}.bias"] = qb[j * DIM:(j + 1) * DIM]
        m[a + "attention.output.dense.weight"], m[a + "attention.output.dense.bias"] = sd[b + "attn.proj.weight"], sd[b + "attn.proj.bias"]
        for k in ("norm1", "norm2"):
            m[a + k + ".weight"], m[a + k + ".bias"] = sd[b + k + ".weight"], sd[b + k + ".bias"]
        m[a + "layer_scale1.lambda1"], m[a + "layer_scale2.lambda1"] = sd[b + "ls1.gamma"], sd[b + "ls2.gamma"]
        for k in ("fc1", "fc2"):
            m[a + f"mlp.{k}.weight"], m[a + f"mlp.{k}.bias"] = sd[b + f"mlp.{k}.weight"], sd[b + f"mlp.{k}.bias"]
    missing, unexpected = hf.load_state_dict(m, strict=False)
    assert not unexpected and not missing, (missing, unexpected)
    for H, W in ((518, 518), (420, 560)):
        x = torch.randn(1, 3, H, W, generator=torch.Generator().manual_seed(H))
        with torch.no_grad():
            a = hub.forward_features(x)["x_norm_patchtokens"]
            b = hf(pixel_values=x).last_hidden_state[:, 5:]
        r = _rel(a, b)
        print(f"\n[dinov2 reg pin] {H}x{W}: rel-L2 {r:.3e}")
        assert r < 1e-5, (H, W, r)
