"""The allocation-free entry point `ufm_op_conv3x3_s2_static` (include/ufm_b200.h; no reference counterpart)
against the existing pair ufm_op_im2col_s2 + ufm_op_gemm, through the C ABI, on
two calls in a row that share the library's static device scratch.  The same two kernels run on the same operands, so
the results are expected to be bit-identical; the assertion uses the tolerance of test_im2col_s2_conv (1e-5 relative).
Named test_zz_* so that pytest collects it last."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from ufm_b200 import _lib  # noqa: E402

TOL = 1e-5


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


def _case(n, h, w, Cc, Co, seed):
    x = _rand(n, h, w, Cc, seed=seed).bfloat16()
    wgt = _rand(Co, Cc, 3, 3, scale=(9 * Cc) ** -0.5, seed=seed + 1).bfloat16()
    bias = _rand(Co, seed=seed + 2)
    wp = wgt.permute(0, 2, 3, 1).reshape(Co, -1).contiguous()
    return x, wgt, wp, bias


def _existing(x, wp, bias):
    """ufm_op_im2col_s2 into a caller-allocated buffer, then ufm_op_gemm (as test_im2col_s2_conv does)."""
    import gpu_ops as ops

    lib = _lib.load()
    n, h, w, Cc = x.shape
    Co = wp.shape[0]
    oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    col = torch.zeros(n * oh * ow, 9 * Cc, dtype=torch.bfloat16, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert lib.ufm_op_im2col_s2(x.data_ptr(), col.data_ptr(), n, h, w, Cc, oh, ow, st) == 0
    out = torch.zeros(n * oh * ow, Co, device="cuda")
    ops.linear(col, wp, bias=bias, out_f32=out)
    torch.cuda.synchronize()
    return out.view(n, oh, ow, Co)


def _static(x, wp, bias, out):
    lib = _lib.load()
    n, h, w, Cc = x.shape
    err = C.create_string_buffer(1024)
    rc = lib.ufm_op_conv3x3_s2_static(x.data_ptr(), n, h, w, Cc, wp.data_ptr(), wp.shape[0],
                                      None if bias is None else bias.data_ptr(), out.data_ptr(),
                                      torch.cuda.current_stream().cuda_stream, err, 1024)
    return rc, err.value.decode()


def test_static_scratch_conv_matches_existing_entry_points_twice():
    # call 1 fills more of the shared scratch than call 2 uses, with different data: stale contents would show in call 2
    cases = [_case(3, 30, 40, 256, 256, seed=11), _case(2, 12, 15, 128, 192, seed=21)]
    expected = [_existing(x, wp, b) for x, _, wp, b in cases]
    outs = [torch.full_like(e, float("nan")) for e in expected]
    torch.cuda.synchronize()
    for (x, _, wp, b), o in zip(cases, outs):  # two calls in a row on one stream, no synchronisation between them
        rc, msg = _static(x, wp, b, o)
        assert rc == 0, msg
    torch.cuda.synchronize()
    for (x, wgt, _, b), o, e in zip(cases, outs, expected):
        assert torch.isfinite(o).all()
        assert rel(o, e) < TOL, rel(o, e)
        assert torch.equal(o, e), "same kernels on the same operands: expected bit-identical results"
        ref = F.conv2d(x.float().permute(0, 3, 1, 2), wgt.float(), b, stride=2, padding=1).permute(0, 2, 3, 1)
        assert rel(o, ref) < TOL, rel(o, ref)
    # and once more in the reverse order (a larger call after a smaller one).  The static array itself (16 MiB) appears
    # when the CUDA module that holds it is loaded -- with lazy loading that is the first call above -- so the free memory
    # of the device is compared around THESE calls: the entry point takes no device memory of its own
    free0 = torch.cuda.mem_get_info()[0]
    for (x, _, wp, b), o, e in reversed(list(zip(cases, outs, expected))):
        o.fill_(float("nan"))
        rc, msg = _static(x, wp, b, o)
        assert rc == 0, msg
        torch.cuda.synchronize()
        assert torch.equal(o, e)
    free1 = torch.cuda.mem_get_info()[0]
    assert free1 >= free0, (free0, free1)


def test_static_scratch_conv_without_bias_and_odd_sizes():
    x, wgt, wp, _ = _case(1, 31, 17, 64, 128, seed=31)
    e = _existing(x, wp, None)
    o = torch.full_like(e, float("nan"))
    rc, msg = _static(x, wp, None, o)
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert rel(o, e) < TOL and torch.equal(o, e)


def test_static_scratch_conv_rejects_sizes_beyond_the_maximum():
    # 8 x 16 x 20 output pixels x 9 x 1024 channels x 2 bytes = 47 MB > 16 MiB; rejected before anything is launched
    n, h, w, Cc, Co = 8, 32, 40, 1024, 128
    assert n * 16 * 20 * 9 * Cc * 2 > _lib.CONV3X3_S2_STATIC_MAX_BYTES
    x = torch.zeros(n, h, w, Cc, dtype=torch.bfloat16, device="cuda")
    wp = torch.zeros(Co, 9 * Cc, dtype=torch.bfloat16, device="cuda")
    o = torch.full((n, 16, 20, Co), 7.0, device="cuda")
    rc, msg = _static(x, wp, None, o)
    torch.cuda.synchronize()
    assert rc == 2
    assert "UFM_CONV3X3_S2_STATIC_MAX_BYTES" in msg and "needs" in msg, msg
    assert (o == 7.0).all()
    # a size just under the maximum is accepted: 1 x 55 x 129 -> 28 x 65 = 1820 rows x 9 x 512 x 2 B = 16 773 120 <= 16 MiB
    n, h, w, Cc = 1, 55, 129, 512
    assert 28 * 65 * 9 * Cc * 2 <= _lib.CONV3X3_S2_STATIC_MAX_BYTES < 28 * 66 * 9 * Cc * 2
    x, wgt, wp, b = _case(n, h, w, Cc, 128, seed=41)
    e = _existing(x, wp, b)
    o = torch.full_like(e, float("nan"))
    rc, msg = _static(x, wp, b, o)
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert torch.equal(o, e)


def test_static_scratch_conv_rejects_bad_arguments():
    x = torch.zeros(1, 8, 8, 96, dtype=torch.bfloat16, device="cuda")  # C not a multiple of 64
    wp = torch.zeros(64, 9 * 96, dtype=torch.bfloat16, device="cuda")
    o = torch.zeros(1, 4, 4, 64, device="cuda")
    rc, msg = _static(x, wp, None, o)
    assert rc == 1 and "multiple of 64" in msg, (rc, msg)
