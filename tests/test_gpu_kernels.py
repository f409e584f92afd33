"""Kernel-level parity on a B200: each hand-written kernel, called through the C ABI, against a plain PyTorch fp32
computation of the same op on the same (bf16-rounded) operands.  Tolerances: fp32-accumulated bf16 products differ from
torch only by summation order (<= 2e-3 relative to the output scale); bf16 outputs add one rounding (2^-9)."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from ufm_b200 import _lib  # noqa: E402


@pytest.fixture(scope="module", autouse=True)
def _built():
    import __graft_entry__ as g

    g.build()
    assert torch.cuda.is_available(), "GPU tests need a B200"
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False


def rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


def _rand(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).cuda()


@pytest.mark.parametrize("M,K,N", [(300, 128, 128), (1000, 384, 384), (2402, 1024, 256), (257, 640, 1536), (4096, 768, 3072)])
def test_gemm_bias_bf16_out(M, K, N):
    import gpu_ops as ops

    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    bias = _rand(N, seed=3)
    out = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    outf = torch.zeros(M, N, dtype=torch.float32, device="cuda")
    ops.linear(a, w, bias=bias, out_bf16=out, out_f32=outf)
    torch.cuda.synchronize()
    ref = a.float() @ w.float().t() + bias
    assert rel(outf, ref) < 1e-5, rel(outf, ref)
    assert rel(out.float(), ref) < 4e-3


def test_gemm_gelu_scale_residual_inplace():
    import gpu_ops as ops

    M, K, N = 1201 * 2, 384, 1536
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    bias = _rand(N, seed=3)
    h = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    ops.linear(a, w, bias=bias, out_bf16=h, flags=_lib.GEMM_GELU)
    ref_h = F.gelu(a.float() @ w.float().t() + bias)
    torch.cuda.synchronize()
    assert rel(h.float(), ref_h) < 4e-3
    # fc2-style: x = x + gamma * (h W2^T + b2), in place on the fp32 residual stream
    w2 = _rand(K, N, scale=N ** -0.5, seed=4).bfloat16()
    b2 = _rand(K, seed=5)
    gamma = _rand(K, seed=6)
    x = _rand(M, K, seed=7)
    ref = x + gamma * (h.float() @ w2.float().t() + b2)
    ops.linear(h, w2, bias=b2, scale=gamma, res1=x, out_f32=x)
    torch.cuda.synchronize()
    assert rel(x, ref) < 1e-5


# 512 x 256 pair tiles (gemm.cu R2; engine.cu rows2_rows): on a 148-SM part the whole waves of a row-matrix GEMM with N % 256
# == 0 run with two 128-row blocks per CTA and one accumulator set, the remaining rows as 256-row half items (column-sliced
# when they fill less than half a wave) in the SAME launch.
# (M, K, N): the UFM-Base encoder shapes (74 x 512 rows + 544 -> 12 half tiles in quarter slices, 36 in half slices), an
# exact multiple (37 x 512 rows x 2 column tiles: one wave, no half items), 77 rows behind a whole wave (2 half tiles in
# 64-column quarter slices taken by one warp half), a case too small for any 512-row wave (falls back entirely), and a
# long-K shape that the shape rule of mode bit 3 admits (the short-K ones then run on the 256-row kernel), and 12 pairs' worth
# of rows, where more than a wave of half items would remain behind the one whole wave (falls back entirely).
@pytest.mark.parametrize("M,K,N", [(32 * 1201, 256, 1024), (32 * 1201, 128, 3072), (37 * 512, 192, 512), (37 * 512 + 77, 64, 512),
                                   (3000, 128, 512), (74 * 512 + 300, 2048, 256), (32 * 1201, 128, 4096), (24 * 1201, 128, 1024)])
@pytest.mark.parametrize("mode,shift", [(3, 0), (1, 0), (3, 4), (7, 0), (11, 0)])
def test_gemm_rows2_pair_tiles(M, K, N, mode, shift):
    """bf16 output with bias and the in-place fp32 residual (TMA reduce-add) through the 512-row tiles: against fp32 torch
    and bit-identical to the 256-row path (same MMA instruction shapes and K order, so the accumulators are the same)."""
    import gpu_ops as ops

    lib = _lib.load()
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    bias = _rand(N, seed=3)
    gamma = _rand(N, seed=4)
    x0 = _rand(M, N, seed=5)
    outs = {}
    try:
        for m in (0, mode):
            lib.ufm_debug_flag(6, m)
            lib.ufm_debug_flag(7, shift if m else 0)
            ob = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
            ops.linear(a, w, bias=bias, out_bf16=ob)
            x = x0.clone()
            ops.linear(a, w, bias=bias, scale=gamma, res1=x, out_f32=x)
            og = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
            ops.linear(a, w, bias=bias, out_bf16=og, flags=_lib.GEMM_GELU)  # 512-row tiles only with mode bit 2
            torch.cuda.synchronize()
            outs[m] = (ob, x, og)
    finally:
        lib.ufm_debug_flag(6, -1)  # back to the library default (UFM_ROWS2 or the built-in one)
        lib.ufm_debug_flag(7, 0)
    ref = a.float() @ w.float().t() + bias
    assert rel(outs[mode][0].float(), ref) < 4e-3
    assert rel(outs[mode][1], x0 + gamma * ref) < 1e-5
    assert torch.equal(outs[mode][0], outs[0][0])
    assert torch.equal(outs[mode][1], outs[0][1])
    assert torch.equal(outs[mode][2], outs[0][2])
    assert rel(outs[mode][2].float(), F.gelu(ref)) < 4e-3


def test_gemm_gelu_packed_bits():
    """The packed f32x2 GELU of the bf16 TMA epilogue (ufm_debug_flag 9) gives the bits of the scalar one."""
    import gpu_ops as ops

    lib = _lib.load()
    M, K, N = 5000, 192, 1024
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=4.0 * K ** -0.5, seed=2).bfloat16()
    bias = _rand(N, seed=3)
    outs = []
    try:
        for m in (0, 1):
            lib.ufm_debug_flag(9, m)
            og = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
            ops.linear(a, w, bias=bias, out_bf16=og, flags=_lib.GEMM_GELU)
            torch.cuda.synchronize()
            outs.append(og)
    finally:
        lib.ufm_debug_flag(9, -1)
    assert torch.equal(outs[0], outs[1])
    assert rel(outs[1].float(), F.gelu(a.float() @ w.float().t() + bias)) < 4e-3


# Last-wave column slices (gemm.cu tile_ref / engine.cu build_gemm) on a 148-SM part: pair tiles 256 x 256 on 74 units,
# single-CTA tiles 128 x 128 on 148.  (M, K, N): 86 row tiles x 1 -> R = 12 -> quarter slices; 104 -> R = 30 -> halves;
# 151 x 4 = 604 -> R = 12 (the UFM-Base encoder proj shape, ragged last row tile); N = 384: BN = 128, 60 x 3 = 180 ->
# R = 32 -> 64-column halves taken by one warp half.
SLICED = [(86 * 256, 128, 256), (104 * 256 - 24, 192, 256), (32 * 1201, 256, 1024), (60 * 128, 128, 384)]


@pytest.mark.parametrize("M,K,N", SLICED)
def test_gemm_last_wave_slices(M, K, N):
    import gpu_ops as ops

    lib = _lib.load()
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    bias, gamma = _rand(N, seed=3), _rand(N, seed=4)
    x0 = _rand(M, N, seed=5)
    ref_lin = a.float() @ w.float().t() + bias
    res = {}
    for split in (1, 0):
        assert lib.ufm_debug_flag(5, split) == 0
        ob = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
        ops.linear(a, w, bias=bias, out_bf16=ob, flags=_lib.GEMM_GELU)  # TMA bf16 epilogue
        x = x0.clone()
        ops.linear(a, w, bias=bias, scale=gamma, res1=x, out_f32=x)  # TMA fp32 epilogue, in-place residual
        of = torch.zeros(M, N, device="cuda")
        ob2 = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
        ops.linear(a, w, bias=bias, res1=x0, out_f32=of, out_bf16=ob2)  # direct epilogue (two outputs)
        torch.cuda.synchronize()
        res[split] = (ob, x, of, ob2)
    lib.ufm_debug_flag(5, 1)
    ob, x, of, ob2 = res[1]
    assert rel(ob.float(), F.gelu(ref_lin)) < 4e-3
    assert rel(x, x0 + gamma * ref_lin) < 1e-5
    assert rel(of, x0 + ref_lin) < 1e-5
    assert rel(ob2.float(), x0 + ref_lin) < 4e-3
    for t1, t0 in zip(res[1], res[0]):  # a slice accumulates the same K sequence as the whole tile: identical bits
        assert torch.equal(t1, t0)


# EPI_TMA_MIX (gemm.cu): the layers with two residuals and / or a bf16 copy next to the fp32 result (RCU convs of the DPT
# fusion blocks, ref:dpt_block.py:156-177, 225-255) take res1 / the fp32 output through TMA slabs and res2 / the bf16 output
# through a coalesced second pass; the adds keep the direct epilogue's order, so ufm_debug_flag(0, 0) gives the same bits.
@pytest.mark.parametrize("n,H,W,Ci,Co,mode", [(2, 30, 40, 256, 256, "rcu1"), (1, 60, 80, 256, 256, "rcu1"),
                                               (2, 15, 20, 256, 256, "rcu1"), (2, 30, 40, 128, 256, "rcu2"),
                                               (1, 24, 32, 64, 64, "rcu1"), (1, 42, 56, 96, 128, "dual")])
def test_conv3x3_mixed_epilogue(n, H, W, Ci, Co, mode):
    import gpu_ops as ops

    lib = _lib.load()
    x = _rand(n, H, W, Ci, seed=1).bfloat16()
    w = _rand(Co, Ci, 3, 3, scale=(9 * Ci) ** -0.5, seed=2).bfloat16().float()
    bias = _rand(Co, seed=3)
    r1, r2 = _rand(n, H, W, Co, seed=4), _rand(n, H, W, Co, seed=5)
    conv = F.conv2d(x.float().permute(0, 3, 1, 2), w, bias, padding=1).permute(0, 2, 3, 1)
    res = {}
    for variant in (1, 0):
        assert lib.ufm_debug_flag(0, variant) == 0
        of = torch.zeros(n, H, W, Co, device="cuda")
        ob = torch.zeros(n, H, W, Co, dtype=torch.bfloat16, device="cuda")
        wp = _pack3(w)
        if mode == "rcu1":  # path + skip + conv: fp32 sum and its ReLU'd bf16 copy
            ops.conv3x3(x, wp, bias=bias, res1=r1, res2=r2, r2=(H * W * Co, W * Co, Co), out_f32=of, out_bf16=ob,
                        flags=_lib.GEMM_RELU_BF16)
        elif mode == "rcu2":  # x + conv, bf16 only
            ops.conv3x3(x, wp, bias=bias, res1=r1, out_bf16=ob)
        else:  # two outputs, no residual
            ops.conv3x3(x, wp, bias=bias, out_f32=of, out_bf16=ob)
        torch.cuda.synchronize()
        res[variant] = (of, ob)
    lib.ufm_debug_flag(0, 1)
    of, ob = res[1]
    want = conv + (r1 if mode != "dual" else 0) + (r2 if mode == "rcu1" else 0)
    if mode != "rcu2":
        assert rel(of, want) < 1e-5
    assert rel(ob.float(), F.relu(want) if mode == "rcu1" else want) < 4e-3
    for t1, t0 in zip(res[1], res[0]):
        assert torch.equal(t1, t0)


def test_gemm_mixed_epilogue_broadcast_residual():
    """Linear with a row-broadcast res2 (view embedding, ref:global_attention_transformer.py:168-178) and in-place res1."""
    import gpu_ops as ops

    lib = _lib.load()
    M, K, N = 2 * 1200 + 37, 256, 768
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    bias, row = _rand(N, seed=3), _rand(N, seed=4)
    x0 = _rand(M, N, seed=5)
    outs = []
    for variant in (1, 0):
        lib.ufm_debug_flag(0, variant)
        x = x0.clone()
        ob = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
        ops.linear(a, w, bias=bias, res1=x, res2=row, r2=(0, 0, 0), out_f32=x, out_bf16=ob)
        torch.cuda.synchronize()
        outs.append((x, ob))
    lib.ufm_debug_flag(0, 1)
    want = a.float() @ w.float().t() + bias + x0 + row
    assert rel(outs[0][0], want) < 1e-5
    assert rel(outs[0][1].float(), want) < 4e-3
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize("M,K,N", [(32 * 1201, 256, 1024), (2400 * 3 + 5, 192, 768), (300, 128, 128)])
def test_gemm_inplace_residual_reduce_add(M, K, N):
    """The in-place residual add x += gamma * (a W^T + b) (ref:transformer_blocks.py:386-403) is done by the L2 through a
    TMA reduce-add store; ufm_debug_flag(0, 5) switches back to the residual slab (TMA load, add in the SM).  One fp32
    rounding either way -> same bits."""
    import gpu_ops as ops

    lib = _lib.load()
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    bias, gamma = _rand(N, seed=3), _rand(N, seed=4)
    x0 = _rand(M, N, seed=5)
    outs = []
    for variant in (1, 5):
        lib.ufm_debug_flag(0, variant)
        x = x0.clone()
        ops.linear(a, w, bias=bias, scale=gamma, res1=x, out_f32=x)
        torch.cuda.synchronize()
        outs.append(x)
    lib.ufm_debug_flag(0, 1)
    assert rel(outs[0], x0 + gamma * (a.float() @ w.float().t() + bias)) < 1e-5
    assert torch.equal(outs[0], outs[1])


def test_gemm_patch_embed_addressing():
    """3-D A view (K, tokens, images), output shifted one row (cls slot) and a row-broadcast positional add."""
    import gpu_ops as ops

    imgs, n, K, D = 3, 192, 640, 384
    a = _rand(imgs * n, K, seed=1).bfloat16()
    w = _rand(D, K, scale=K ** -0.5, seed=2).bfloat16()
    bias = _rand(D, seed=3)
    pos = _rand(n + 1, D, seed=4)
    x = torch.full((imgs, n + 1, D), 7.0, device="cuda")
    ops.gemm(a, w, X=n, Y=1, Nimg=imgs, C_=K, a_strides=(K, n * K, n * K), bias=bias, res2=pos[1:].contiguous(),
             r2=(0, 0, D), out_f32=x, out_strides=(D, 0, (n + 1) * D), out_off=D)
    torch.cuda.synchronize()
    ref = (a.float() @ w.float().t() + bias).view(imgs, n, D) + pos[1:]
    assert rel(x[:, 1:], ref) < 1e-5
    assert torch.all(x[:, 0] == 7.0)


def _pack3(w):
    from ufm_b200.packing import _conv3_pack

    return _conv3_pack(w.cpu()).cuda()


@pytest.mark.parametrize("n,H,W,Ci,Co", [(2, 30, 40, 96, 64), (1, 15, 20, 64, 64), (2, 24, 32, 256, 256), (1, 42, 56, 32, 128)])
def test_conv3x3_implicit_gemm(n, H, W, Ci, Co):
    import gpu_ops as ops

    x = _rand(n, H, W, Ci, seed=1).bfloat16()
    w = _rand(Co, Ci, 3, 3, scale=(9 * Ci) ** -0.5, seed=2).bfloat16().float()
    bias = _rand(Co, seed=3)
    r1 = _rand(n, H, W, Co, seed=4)
    r2 = _rand(n, H, W, Co, seed=5)
    of = torch.zeros(n, H, W, Co, device="cuda")
    ob = torch.zeros(n, H, W, Co, dtype=torch.bfloat16, device="cuda")
    ops.conv3x3(x, _pack3(w), bias=bias, res1=r1, res2=r2, r2=(H * W * Co, W * Co, Co), out_f32=of, out_bf16=ob,
                flags=_lib.GEMM_RELU_BF16)
    torch.cuda.synchronize()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w, bias, padding=1).permute(0, 2, 3, 1) + r1 + r2
    assert rel(of, ref) < 1e-5, rel(of, ref)
    assert rel(ob.float(), F.relu(ref)) < 4e-3


def test_conv3x3_tail_epilogue():
    import gpu_ops as ops

    n, H, W, Ci, hid, out = 2, 42, 56, 128, 128, 2
    x = _rand(n, H, W, Ci, seed=1).bfloat16()
    w = _rand(hid, Ci, 3, 3, scale=(9 * Ci) ** -0.5, seed=2).bfloat16().float()
    bias = _rand(hid, seed=3)
    w2 = _rand(out, hid, scale=hid ** -0.5, seed=4)
    b2, sc, sh = [0.3, -0.2], [2.0, 3.0], [0.5, -1.0]
    y = torch.zeros(n, out, H, W, device="cuda")
    raw = torch.zeros(n, out, H, W, device="cuda")
    ops.conv3x3(x, _pack3(w), bias=bias, out_f32=y, out2=raw, flags=_lib.GEMM_TAIL, tail=(w2, b2, sc, sh))
    torch.cuda.synchronize()
    hdn = F.relu(F.conv2d(x.float().permute(0, 3, 1, 2), w, bias, padding=1))
    ref_raw = F.conv2d(hdn, w2.view(out, hid, 1, 1), torch.tensor(b2).cuda())
    ref = ref_raw * torch.tensor(sc).cuda().view(1, 2, 1, 1) + torch.tensor(sh).cuda().view(1, 2, 1, 1)
    assert rel(raw, ref_raw) < 1e-5
    assert rel(y, ref) < 1e-5
    # sigmoid variant, one output, narrow hidden (tiny config)
    w2 = _rand(1, hid, scale=hid ** -0.5, seed=5)
    y1 = torch.zeros(n, 1, H, W, device="cuda")
    ops.conv3x3(x, _pack3(w), bias=bias, out_f32=y1, flags=_lib.GEMM_TAIL | _lib.GEMM_TAIL_SIGMOID,
                tail=(w2, [0.1], [1.0], [0.0]))
    torch.cuda.synchronize()
    ref1 = torch.sigmoid(F.conv2d(hdn, w2.view(1, hid, 1, 1), torch.tensor([0.1]).cuda()))
    assert (y1 - ref1).abs().max().item() < 1e-5


@pytest.mark.parametrize("n_seq,seq_len,heads", [(2, 193, 6), (3, 1201, 2), (1, 2400, 3), (2, 130, 2), (1, 256, 1), (2, 384, 2),
                                                   (1, 4801, 2), (1, 9600, 1)])  # the last two: 1120x840 encoder / info sharing
def test_attention(n_seq, seq_len, heads):
    import gpu_ops as ops

    D = heads * 64
    qkv = _rand(n_seq * seq_len, 3 * D, seed=1).bfloat16()
    qkv[:, :D] *= 2.0  # sharper softmax: exercises the lazy-rescale path
    out = ops.attention(qkv, n_seq, seq_len, heads, 0.125)
    torch.cuda.synchronize()
    t = qkv.float().view(n_seq, seq_len, 3, heads, 64).permute(2, 0, 3, 1, 4)
    ref = F.scaled_dot_product_attention(t[0], t[1], t[2], scale=0.125).transpose(1, 2).reshape(n_seq * seq_len, D)
    r = rel(out.float(), ref)
    assert r < 1e-2, r  # P is rounded to bf16 before the PV product (as in every flash kernel)


def test_layernorm_remap():
    lib = _lib.load()
    B, S, D = 2, 13, 384
    x = _rand(2 * B * S, D, seed=1)
    w, b = _rand(D, seed=2), _rand(D, seed=3)
    ob = torch.zeros(B * 2 * (S - 1), D + 128, dtype=torch.bfloat16, device="cuda")
    of = torch.zeros(B * 2 * (S - 1), D, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert lib.ufm_op_layernorm(x.data_ptr(), 2 * B, S, 1, S - 1, D, w.data_ptr(), b.data_ptr(), 1e-6, B, ob.data_ptr(),
                                D + 128, 128, of.data_ptr(), D, 0, st) == 0
    torch.cuda.synchronize()
    ref = F.layer_norm(x.view(2, B, S, D)[:, :, 1:], (D,), w, b, 1e-6).permute(1, 0, 2, 3).reshape(-1, D)
    assert rel(of, ref) < 1e-6
    assert rel(ob[:, 128:].float(), ref) < 3e-3
    assert torch.all(ob[:, :128] == 0)


def test_qk_norm():
    lib = _lib.load()
    rows, heads = 77, 3
    qkv = _rand(rows, 3 * heads * 64, seed=1).bfloat16()
    orig = qkv.clone()
    ws = [_rand(64, seed=s) for s in (2, 3, 4, 5)]
    st = torch.cuda.current_stream().cuda_stream
    assert lib.ufm_op_qk_norm(qkv.data_ptr(), rows, heads, *[t.data_ptr() for t in ws], 1e-6, st) == 0
    torch.cuda.synchronize()
    o = orig.float().view(rows, 3, heads, 64)
    q = F.layer_norm(o[:, 0], (64,), ws[0], ws[1], 1e-6)
    k = F.layer_norm(o[:, 1], (64,), ws[2], ws[3], 1e-6)
    got = qkv.float().view(rows, 3, heads, 64)
    assert rel(got[:, 0], q) < 4e-3 and rel(got[:, 1], k) < 4e-3
    assert torch.equal(got[:, 2], o[:, 2])


def test_rope2d_vs_oracle():
    """RoPE2D on q and k of a fused qkv buffer (2 sequences x 2 views x a 30 x 40 patch grid) vs the oracle's restatement
    of pos_embed.py:108-155; v untouched."""
    from oracle import ufm_oracle as O

    lib = _lib.load()
    h, w, heads, B = 30, 40, 3, 2
    n = h * w
    rows = B * 2 * n
    qkv = _rand(rows, 3 * heads * 64, seed=7).bfloat16()
    orig = qkv.clone()
    assert lib.ufm_op_rope(qkv.data_ptr(), rows, heads, n, w, torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    o = orig.float().cpu().view(B, 2 * n, 3, heads, 64).permute(2, 0, 3, 1, 4)  # (3, B, heads, 2n, 64)
    pos = O.grid_positions(B, h, w, views=2)
    got = qkv.float().cpu().view(B, 2 * n, 3, heads, 64).permute(2, 0, 3, 1, 4)
    for i in (0, 1):
        ref = O.rope2d(o[i], pos)
        assert rel(got[i], ref) < 4e-3, (i, rel(got[i], ref))
        assert rel(got[i], o[i]) > 0.5  # the rotation is far from the identity on this grid
    assert torch.equal(got[2], o[2])


@pytest.mark.parametrize("h,w,rh,rw,oh,ow", [(15, 20, 30, 40, 30, 40), (8, 10, 16, 20, 15, 19), (96, 128, 168, 224, 168, 224),
                                             (30, 40, 15, 20, 15, 20), (20, 30, 25, 37, 25, 37), (48, 64, 84, 112, 84, 112),
                                             (6, 7, 40, 61, 40, 61)])
def test_resize_align_corners(h, w, rh, rw, oh, ow):
    lib = _lib.load()
    n, Cc = 2, 64
    x = _rand(n, h, w, Cc, seed=1)
    of = torch.zeros(n, oh, ow, Cc, device="cuda")
    ob = torch.zeros(n, oh, ow, Cc, dtype=torch.bfloat16, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert lib.ufm_op_resize_ac(x.data_ptr(), None, n, h, w, Cc, rh, rw, oh, ow, of.data_ptr(), ob.data_ptr(), st) == 0
    xb = x.bfloat16()
    of2 = torch.zeros_like(of)
    assert lib.ufm_op_resize_ac(None, xb.data_ptr(), n, h, w, Cc, rh, rw, oh, ow, of2.data_ptr(), None, st) == 0
    torch.cuda.synchronize()
    ref = F.interpolate(x.permute(0, 3, 1, 2), size=(rh, rw), mode="bilinear", align_corners=True)[:, :, :oh, :ow]
    ref = ref.permute(0, 2, 3, 1)
    assert (of - ref).abs().max().item() < 5e-5
    assert rel(ob.float(), ref) < 3e-3
    ref2 = F.interpolate(xb.float().permute(0, 3, 1, 2), size=(rh, rw), mode="bilinear", align_corners=True)
    # the kernel interpolates vertically first (column values are shared along a strip), torch horizontally first:
    # fp32 rounding-order differences only
    assert (of2 - ref2[:, :, :oh, :ow].permute(0, 2, 3, 1)).abs().max().item() < 5e-5


def test_resize_upsampling_kernel_matches_generic(monkeypatch):
    """The up-front-load kernel used for 7*sx < 4 performs the generic strip kernel's arithmetic: identical bits."""
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    for (h, w, oh, ow, Cc) in ((15, 20, 30, 40, 64), (24, 32, 42, 56, 128), (9, 11, 18, 21, 32), (5, 3, 40, 23, 8)):
        x = _rand(2, h, w, Cc, seed=3)
        xb = x.bfloat16()
        outs = []
        for generic in (False, True):
            if generic:
                monkeypatch.setenv("UFM_RESIZE_GENERIC", "1")
            else:
                monkeypatch.delenv("UFM_RESIZE_GENERIC", raising=False)
            of = torch.zeros(2, oh, ow, Cc, device="cuda")
            ob = torch.zeros(2, oh, ow, Cc, dtype=torch.bfloat16, device="cuda")
            ob2 = torch.zeros_like(ob)
            assert lib.ufm_op_resize_ac(x.data_ptr(), None, 2, h, w, Cc, oh, ow, oh, ow, of.data_ptr(), ob.data_ptr(), st) == 0
            assert lib.ufm_op_resize_ac(None, xb.data_ptr(), 2, h, w, Cc, oh, ow, oh, ow, None, ob2.data_ptr(), st) == 0
            torch.cuda.synchronize()
            outs.append((of, ob, ob2))
        for a, b in zip(*outs):
            assert torch.equal(a, b)
        assert outs[0][0].abs().max().item() > 0


def test_shuffle_and_convt_fold():
    """1x1 conv -> ConvTranspose2d(k==s) folded into one GEMM + pixel shuffle (dpt.py:107-125)."""
    import gpu_ops as ops
    from ufm_b200.packing import _fold_convt

    lib = _lib.load()
    n, h, w, Ci, m, k = 2, 12, 16, 384, 64, 4
    x = _rand(n, h, w, Ci, seed=1).bfloat16()
    w1, b1 = _rand(m, Ci, 1, 1, scale=Ci ** -0.5, seed=2), _rand(m, seed=3)
    wt, bt = _rand(m, m, k, k, scale=m ** -0.5, seed=4), _rand(m, seed=5)
    wf, bf_ = _fold_convt(w1.cpu(), b1.cpu(), wt.cpu(), bt.cpu())
    wf, bf_ = wf.cuda(), bf_.cuda()
    G = torch.zeros(n * h * w, k * k * m, dtype=torch.bfloat16, device="cuda")
    ops.linear(x.view(-1, Ci), wf, bias=bf_, out_bf16=G)
    T = torch.zeros(n, h * k, w * k, m, dtype=torch.bfloat16, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert lib.ufm_op_shuffle(G.data_ptr(), T.data_ptr(), 0, n, h, w, k, m, st) == 0
    torch.cuda.synchronize()
    ref = F.conv_transpose2d(F.conv2d(x.float().permute(0, 3, 1, 2), w1, b1), wt, bt, stride=k).permute(0, 2, 3, 1)
    assert rel(T.float(), ref) < 6e-3


def test_im2col_s2_conv():
    import gpu_ops as ops

    lib = _lib.load()
    n, h, w, Cc, Co = 2, 12, 15, 128, 128
    oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    x = _rand(n, h, w, Cc, seed=1).bfloat16()
    wgt = _rand(Co, Cc, 3, 3, scale=(9 * Cc) ** -0.5, seed=2).bfloat16()
    bias = _rand(Co, seed=3)
    col = torch.zeros(n * oh * ow, 9 * Cc, dtype=torch.bfloat16, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert lib.ufm_op_im2col_s2(x.data_ptr(), col.data_ptr(), n, h, w, Cc, oh, ow, st) == 0
    out = torch.zeros(n * oh * ow, Co, device="cuda")
    ops.linear(col, wgt.permute(0, 2, 3, 1).reshape(Co, -1).contiguous(), bias=bias, out_f32=out)
    torch.cuda.synchronize()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wgt.float(), bias, stride=2, padding=1).permute(0, 2, 3, 1)
    assert tuple(ref.shape[1:3]) == (oh, ow)
    assert rel(out.view(n, oh, ow, Co), ref) < 1e-5


def test_local_refinement_vs_oracle():
    from oracle import ufm_oracle as O

    lib = _lib.load()
    B, H, W, Cc = 2, 56, 70, 16
    g = torch.Generator().manual_seed(0)
    desc = torch.randn(B, 2, H, W, Cc, generator=g).bfloat16().float()  # descriptors are bf16 in the engine
    # smooth field U(-40,40) + N(0,1), with some targets far outside the image
    base = F.interpolate(torch.rand(B, 2, 4, 5, generator=g) * 80 - 40, size=(H, W), mode="bilinear", align_corners=True)
    flow = base + torch.randn(B, 2, H, W, generator=g)
    flow[0, :, :2, :2] = 1e4
    bias = torch.randn(25, generator=g) * 0.3
    # engine layout: channel-group planar [B, view, C/8, H, W, 8] bf16
    d_desc = desc.view(B, 2, H, W, Cc // 8, 8).permute(0, 1, 4, 2, 3, 5).contiguous().bfloat16().cuda()
    d_flow, d_bias = flow.cuda(), bias.cuda()
    out = torch.zeros(B, 2, H, W, device="cuda")
    res = torch.zeros(B, 2, H, W, device="cuda")
    lsm = torch.zeros(B, H, W, 25, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert lib.ufm_op_refine(d_desc.data_ptr(), B, H, W, d_flow.data_ptr(), d_bias.data_ptr(), 0.25, out.data_ptr(),
                             res.data_ptr(), lsm.data_ptr(), st) == 0
    torch.cuda.synchronize()
    f1 = desc[:, 0].permute(0, 3, 1, 2).contiguous()
    f2 = desc[:, 1].permute(0, 3, 1, 2).contiguous()
    r_ref, l_ref = O.local_refinement(flow, f1, f2, bias, 4.0, 5)
    assert (res.cpu() - r_ref).abs().max().item() < 2e-4
    assert (lsm.cpu().view(B, H, W, 5, 5) - l_ref).abs().max().item() < 2e-3
    assert (out.cpu() - (flow + r_ref)).abs()[..., 2:, 2:].max().item() < 1e-3


# ---------------------------------------------------------------------------------------------------------------------
# non-native sizes: normalise + anti-alias resize in, un-map out
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("H0,W0,th,tw", [(300, 400, 42, 56), (97, 131, 56, 70), (640, 853, 56, 70), (56, 70, 56, 70),
                                         (40, 33, 84, 112)])
def test_prep_resize_vs_torch_antialias(H0, W0, th, tw):
    """normalise (base.py:215-229) + F.interpolate(bilinear, antialias=True) (flow_resizing.py:276-354), fp32 CPU torch as
    the checker; uint8 BHWC / BCHW and fp32 (identity and re-normalised)."""
    from ufm_b200 import prepost

    g = torch.Generator().manual_seed(H0 * 7 + W0)
    u8 = torch.randint(0, 256, (2, H0, W0, 3), dtype=torch.uint8, generator=g)
    mean, std = prepost.norm_tensors("dinov2", "cpu")
    x = (u8.permute(0, 3, 1, 2).float() / 255.0 - mean) / std
    ref = F.interpolate(x, size=(th, tw), mode="bilinear", align_corners=False, antialias=True)
    a = prepost.prep_resize(u8.cuda(), True, th, tw, "dinov2").cpu()
    b = prepost.prep_resize(u8.permute(0, 3, 1, 2).contiguous().cuda(), False, th, tw, "dinov2").cpu()
    assert torch.equal(a, b)
    assert (a - ref).abs().max().item() < 2e-5, (a - ref).abs().max().item()
    c = prepost.prep_resize(x.cuda(), False, th, tw, "dinov2", "dinov2").cpu()
    assert (c - ref).abs().max().item() < 2e-5
    # fp32 image normalised for another model family -> re-normalised to the encoder's statistics
    pm, ps = prepost.norm_tensors("dust3r", "cpu")
    xd = (u8.permute(0, 3, 1, 2).float() / 255.0 - pm) / ps
    refd = F.interpolate(xd * (ps / std) + (pm - mean) / std, size=(th, tw), mode="bilinear", align_corners=False, antialias=True)
    d = prepost.prep_resize(xd.permute(0, 2, 3, 1).contiguous().cuda(), True, th, tw, "dinov2", "dust3r").cpu()
    assert (d - refd).abs().max().item() < 2e-5


@pytest.mark.parametrize("H0,W0,H1,W1,H,W", [(300, 400, 300, 400, 42, 56), (97, 131, 120, 90, 56, 70), (30, 40, 30, 40, 56, 70),
                                             (56, 70, 56, 70, 56, 70)])
def test_unmap_vs_torch_shell(H0, W0, H1, W1, H, W):
    """unmap_predicted_flow / unmap_predicted_channels (flow_resizing.py:749-877, 955-1010) against their torch
    restatement (oracle.ufm_oracle.unmap_flow / unmap_channels) on the CPU."""
    from oracle import ufm_oracle as O
    from ufm_b200 import prepost

    g = torch.Generator().manual_seed(H0 + W1)
    flow = torch.randn(2, 2, H, W, generator=g) * 10
    cov = torch.rand(2, H, W, generator=g)
    rep0 = prepost.scaled_region(H, W, H0, W0)
    src0, src1 = [0, H0, 0, W0], [0, H1, 0, W1]
    rf = O.unmap_flow(flow, rep0, src0, src1, (H0, W0))
    rc = O.unmap_channels(cov.unsqueeze(1), rep0, src0, (H0, W0)).squeeze(1)
    f, c = prepost.unmap(flow.cuda(), cov.cuda(), rep0, src0, src1, (H0, W0))
    assert f.shape == rf.shape and c.shape == rc.shape
    assert torch.equal(c.cpu(), rc)
    assert (f.cpu() - rf).abs().max().item() < 2e-4, (f.cpu() - rf).abs().max().item()  # fp32 cancellation at |coord| ~ W
    # a source region smaller than the image: everything outside stays zero
    src0b = [3, H0 - 2, 5, W0 - 4]
    rep0b = prepost.scaled_region(H, W, H0, W0)
    rf2 = O.unmap_flow(flow, rep0b, src0b, src1, (H0, W0))
    f2, _ = prepost.unmap(flow.cuda(), None, rep0b, src0b, src1, (H0, W0))
    assert (f2.cpu() - rf2).abs().max().item() < 2e-4
    assert f2[:, :, :3].abs().max().item() == 0 and f2[:, :, :, :5].abs().max().item() == 0
