"""tkb_qnn_conv2d_noalloc (include/tkb200.h): the allocation-free convolution entry point, called through the C ABI
and compared BIT FOR BIT (integer path: np.testing.assert_equal, the tolerance of every integer test here) with the
existing entry point (tkb_run on a one-node qnn.conv2d graph) and with the CPU oracle, on calls in a row that share
the static scratch array.  The file name sorts last on purpose."""
import ctypes

import numpy as np
import pytest

from oracle import interp
from tachikoma_b200 import relay as R, runtime as RT

pytestmark = pytest.mark.gpu


def _case(seed, shape, oc, ker, stride, pad, dil, zp_x, zp_w, x_dtype="uint8", w_dtype="int8"):
    rng = np.random.default_rng(seed)
    n, c, h, w = shape
    xinfo, winfo = np.iinfo(x_dtype), np.iinfo(w_dtype)
    x = rng.integers(xinfo.min, xinfo.max + 1, shape).astype(x_dtype)
    wgt = rng.integers(winfo.min, winfo.max + 1, (oc, c, ker[0], ker[1])).astype(w_dtype)
    d = RT.Conv2dDesc(n=n, c=c, h=h, w=w, o=oc, kh=ker[0], kw=ker[1], stride_h=stride[0], stride_w=stride[1],
                      dilation_h=dil[0], dilation_w=dil[1], pad_top=pad[0], pad_left=pad[1], pad_bottom=pad[2],
                      pad_right=pad[3], data_dtype_code=0 if x_dtype == "int8" else 1,
                      weight_dtype_code=0 if w_dtype == "int8" else 1, input_zero_point=zp_x, kernel_zero_point=zp_w)
    var = R.var("in_0", shape, x_dtype)
    func = R.Function([var], R.qnn_conv2d(var, R.const(wgt), zp_x, zp_w, 0.5, 0.25, ker, oc, strides=stride, padding=pad,
                                          dilation=dil))
    return d, x, wgt, func


def _existing_entry_point(func, x, conv_impl):
    mod, run = RT.create_executor(func, conv_impl=conv_impl)
    return run(in_0=x)


def _noalloc(d, x, wgt, stream=None, fill=None):
    import torch
    xd, wd = torch.from_numpy(x.view(np.uint8)).cuda(), torch.from_numpy(wgt.view(np.uint8)).cuda()
    oh = (d.h + d.pad_top + d.pad_bottom - d.dilation_h * (d.kh - 1) - 1) // d.stride_h + 1
    ow = (d.w + d.pad_left + d.pad_right - d.dilation_w * (d.kw - 1) - 1) // d.stride_w + 1
    out = torch.full((d.n, d.o, oh, ow), -7 if fill is None else fill, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    RT.qnn_conv2d_noalloc(d, xd.data_ptr(), wd.data_ptr(), out.data_ptr(), stream=stream)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return out.cpu().numpy(), free0 - free1


CASES = [
    # ResNet-style 3x3 with padding and an input zero point, then a smaller ragged case that reuses the scratch
    dict(seed=1, shape=(2, 16, 14, 14), oc=40, ker=(3, 3), stride=(1, 1), pad=(1, 1, 1, 1), dil=(1, 1), zp_x=3, zp_w=0),
    dict(seed=2, shape=(1, 5, 9, 11), oc=7, ker=(3, 2), stride=(2, 1), pad=(0, 1, 2, 0), dil=(1, 2), zp_x=17, zp_w=0),
    dict(seed=3, shape=(3, 3, 20, 20), oc=65, ker=(7, 7), stride=(2, 2), pad=(3, 3, 3, 3), dil=(1, 1), zp_x=0, zp_w=0),
    dict(seed=4, shape=(2, 8, 6, 6), oc=16, ker=(1, 1), stride=(1, 1), pad=(0, 0, 0, 0), dil=(1, 1), zp_x=-5, zp_w=0,
         x_dtype="int8"),
]


def test_noalloc_conv_matches_existing_entry_point_on_calls_in_a_row():
    """Calls back to back on the shared scratch (different shapes, so every call sees the previous call's bytes), then
    the first case again: each equals tkb_run's explicit-im2col convolution, its default route, and the oracle."""
    wants = []
    for kw in CASES:
        d, x, wgt, func = _case(**kw)
        want = _existing_entry_point(func, x, "tc_explicit")
        np.testing.assert_equal(_existing_entry_point(func, x, "auto"), want)
        ref, _ = interp.run_expr(func, {"in_0": x})
        np.testing.assert_equal(want, ref)
        wants.append((d, x, wgt, want))
    leaks = []
    for rep in range(2):
        for d, x, wgt, want in wants:
            got, leaked = _noalloc(d, x, wgt)
            assert got.shape == want.shape and got.dtype == np.int32
            np.testing.assert_equal(got, want)
            leaks.append(leaked)
    # after the first pass (lazy module loading) a call must not change the device's free memory
    assert all(v == 0 for v in leaks[len(wants):]), leaks


def test_noalloc_conv_kernel_zero_point_and_stream():
    import torch
    d, x, wgt, func = _case(seed=9, shape=(2, 6, 8, 8), oc=12, ker=(3, 3), stride=(1, 1), pad=(1, 1, 1, 1), dil=(1, 1),
                            zp_x=4, zp_w=3)
    ref, _ = interp.run_expr(func, {"in_0": x})
    s = torch.cuda.Stream()
    for _ in range(2):
        got, _ = _noalloc(d, x, wgt, stream=s.cuda_stream)
        np.testing.assert_equal(got, ref)
    try:
        want = _existing_entry_point(func, x, "auto")
    except RT.TachikomaError:
        return  # the JSON runtime does not take a kernel zero point on this route; the oracle check above stands
    np.testing.assert_equal(want, ref)


def test_noalloc_conv_rejects_what_does_not_fit():
    import torch
    buf = torch.zeros(64, dtype=torch.int32, device="cuda")
    p = buf.data_ptr()
    big, _, _, _ = _case(seed=0, shape=(1, 1, 4, 4), oc=1, ker=(1, 1), stride=(1, 1), pad=(0, 0, 0, 0), dil=(1, 1), zp_x=0,
                         zp_w=0)
    big.n, big.h, big.w, big.c = 64, 224, 224, 64  # 64*224*224 rows of 64 bytes = 196 MiB > 16 MiB
    with pytest.raises(RT.TachikomaError, match="static scratch holds 16777216"):
        RT.qnn_conv2d_noalloc(big, p, p, p)
    deep, _, _, _ = _case(seed=0, shape=(1, 1, 4, 4), oc=1, ker=(3, 3), stride=(1, 1), pad=(1, 1, 1, 1), dil=(1, 1),
                          zp_x=0, zp_w=0)
    deep.c = 2048  # K = 18432 > 12288
    with pytest.raises(RT.TachikomaError, match="exceeds the maximum of 12288"):
        RT.qnn_conv2d_noalloc(deep, p, p, p)
    deep.c, deep.input_zero_point = 4, 300
    with pytest.raises(RT.TachikomaError, match="representable"):
        RT.qnn_conv2d_noalloc(deep, p, p, p)
    with pytest.raises(RT.TachikomaError, match="null argument"):
        RT.qnn_conv2d_noalloc(deep, 0, p, p)
    assert RT.NOALLOC_SCRATCH_BYTES == 16 << 20 and RT.NOALLOC_MAX_K == 12288
    lib = RT.load_library()
    assert isinstance(lib.tkb_qnn_conv2d_noalloc, ctypes._CFuncPtr)
