"""Pins the CPU oracle against the reference's own golden vectors.

The fixtures were produced by executing the reference's unmodified test files
(tests/python/relay/test_op_qnn_{requantize,dense,add,batch_matmul,conv2d}.py) under a shim, see
tests/golden/gen_from_reference_tests.py.  Bit-exact (assert_equal), like the reference."""
import numpy as np
import pytest

from golden_util import load_cases
from oracle import interp

CASES = load_cases()


@pytest.mark.parametrize("case", CASES, ids=[c["id"] for c in CASES])
def test_oracle_matches_reference_golden(case):
    out, _ = interp.run_expr(case["func"], case["inputs"])
    np.testing.assert_equal(out, case["golden"])


def test_fixed_point_multiplier_pairs():
    """(multiplier, shift) <-> value pairs of tests/python/contrib/test_hexagon/test_fixed_point_multiply.py:96-100.
    The first two are exactly what GetFixedPointMultiplierShift (src/relay/qnn/utils.cc:33-57) returns for the float32
    scale (the QNN ops hand it double(float32)); asserted exactly.  The third pair of that file (1288490188, 0) is a
    hand-written, truncated significand for 0.6 (round(0.6 * 2^31) = 1288490189; float32(0.6) gives 1288490240), so it
    is not an output of the function: asserted here is the exact output for both roundings of 0.6, and that the
    file's pair denotes the same value to within one unit of the significand."""
    f32 = lambda v: float(np.float32(v))
    assert interp.fixed_point_multiplier_shift(f32(0.15)) == (1288490240, -2)
    assert interp.fixed_point_multiplier_shift(f32(1.3)) == (1395864320, 1)
    assert interp.fixed_point_multiplier_shift(0.6) == (1288490189, 0)
    assert interp.fixed_point_multiplier_shift(f32(0.6)) == (1288490240, 0)
    assert abs(1288490188 - interp.fixed_point_multiplier_shift(0.6)[0]) == 1
    # boundary cases of utils.cc:33-57: zero, exact powers of two, a significand that rounds up to 2^31
    assert interp.fixed_point_multiplier_shift(0.0) == (0, 0)
    assert interp.fixed_point_multiplier_shift(1.0) == (1 << 30, 1)
    assert interp.fixed_point_multiplier_shift(0.25) == (1 << 30, -1)
    assert interp.fixed_point_multiplier_shift(1.0 - 2.0 ** -33) == (1 << 30, 1)


def test_avg_pool_reduces_in_the_tensor_dtype():
    """include/tvm/topi/nn/pooling.h:599: the window sum is tvm::sum over the tensor's own dtype, so it wraps at 8 bits
    for a uint8 tensor; the windowed form divides the (promoted) wrapped sum by an int32 count (:612-648), the adaptive
    / global form multiplies its divisor up in the tensor dtype (:384-389).  Values derived by hand."""
    x = np.full((1, 1, 2, 2), 200, dtype=np.uint8)          # 4 * 200 = 800 = 3 * 256 + 32
    out = interp.pool2d(x, [2, 2], [2, 2], [0, 0, 0, 0], "NCHW", 1, False)
    assert out.reshape(-1).tolist() == [32 // 4]
    xs = np.full((1, 1, 2, 2), -100, dtype=np.int8)         # -400 wraps to 112 in int8
    out = interp.pool2d(xs, [2, 2], [2, 2], [0, 0, 0, 0], "NCHW", 1, False)
    assert out.reshape(-1).tolist() == [112 // 4]
    x32 = np.full((1, 1, 2, 2), 200, dtype=np.int32)        # int32 (the frontend's upcast form): no wrap
    assert interp.pool2d(x32, [2, 2], [2, 2], [0, 0, 0, 0], "NCHW", 1, False).reshape(-1).tolist() == [200]
    # adaptive form on uint8, 20 x 20 window: divisor 400 wraps to 144, sum 400 * 3 = 1200 wraps to 176
    xg = np.full((1, 1, 20, 20), 3, dtype=np.uint8)
    assert interp.pool2d(xg, [20, 20], [1, 1], [0, 0, 0, 0], "NCHW", 2, False).reshape(-1).tolist() == [176 // 144]
    # negative int32 sums: truncation toward zero (-7 / 4 = -1, not -2)
    xn = np.array([[[[-1, -2], [-2, -2]]]], dtype=np.int32)
    assert interp.pool2d(xn, [2, 2], [2, 2], [0, 0, 0, 0], "NCHW", 1, False).reshape(-1).tolist() == [-1]


def test_fixed_point_multiply_level3():
    # tests/python/relay/test_op_level3.py:129-144: 23 * (1/16) with M=1073741824, s=-3 -> 1 (atol 1)
    x = np.full((2, 3), 23, dtype=np.int32)
    y = interp.fixed_point_multiply(x, 1073741824, -3)
    assert np.all(np.abs(y - 23 * 0.5 * 2.0 ** -3) <= 1)
    assert np.all(y == 1)
