"""The arithmetic behind `rsqrt_` (pyseqm_b200/csrc/common.cuh) and `rsqrt_unit` (eigh_warp.cuh): a hardware seed of
~20 good bits followed by ONE third-order step y (1 + e/2 + 3 e^2 / 8), e = 1 - a y^2.  numpy restatement of the step
with a seed truncated the way MUFU.RSQ64H delivers it (the high 32 bits of the double: 20 mantissa bits): the result
is within an ulp or two of 1/sqrt(a) over the range of the arguments the pair pass produces (squared distances plus
squared charge separations, bohr^2)."""
import numpy as np


def _step(a, y):
    e = 1.0 - (a * y) * y          # fma(-(a y), y, 1) up to one rounding
    return y + (y * e) * (0.5 + 0.375 * e)


def _truncate_to_high_word(x):
    return (x.view(np.uint64) & np.uint64(0xFFFFFFFF00000000)).view(np.float64)


def test_one_cubic_step_from_a_20_bit_seed_is_exact_to_rounding():
    rng = np.random.RandomState(0)
    a = np.concatenate([10.0 ** rng.uniform(-3, 4, 200000), rng.uniform(1.0, 2.0, 50000)])
    exact = 1.0 / np.sqrt(a.astype(np.longdouble))
    # seed: the correctly rounded value cut to the high word, and the same perturbed by the documented 2^-20 or so
    for rel in (0.0, 2.0 ** -20, -2.0 ** -20, 2.0 ** -18):
        seed = _truncate_to_high_word((1.0 / np.sqrt(a)) * (1.0 + rel))
        y = _step(a, seed)
        err = np.abs((y.astype(np.longdouble) - exact) / exact).max()
        assert float(err) < 4.0 * 2.0 ** -53, (rel, float(err))


def test_a_worse_seed_would_show():
    """the test above is not vacuous: from a 12-bit seed one cubic step is visibly short of double precision"""
    a = np.linspace(1.0, 2.0, 1001)
    seed = (1.0 / np.sqrt(a)) * (1.0 + 2.0 ** -12)
    y = _step(a, seed)
    err = np.abs(y * np.sqrt(a) - 1.0).max()
    assert err > 1.0e-12
