"""The oracle's specified f32 routines (log, cos, exp2, pow: oracle/rto_math.h) against
float64 libm: well inside what the reference's own stack guarantees (Vulkan: log 3 ULP,
cos 2^-11 absolute, pow = exp2(y*log2 x))."""
import ctypes as C

import numpy as np


def _ulp_err(got, want64):
    want32 = want64.astype(np.float32)
    ulp = np.spacing(np.abs(want32)).astype(np.float64)
    return np.abs(got.astype(np.float64) - want64) / ulp


def _vec(fn, xs):
    return np.array([fn(C.c_float(float(x))) for x in xs], np.float32)


def test_log(oracle):
    L = oracle.lib()
    rs = np.random.RandomState(1)
    xs = np.concatenate([rs.uniform(0, 1, 20000), 2.0 ** rs.uniform(-40, 0, 5000), [1.0, 0.5, 2.0 ** -32, 1 - 2.0 ** -24]]).astype(np.float32)
    xs = xs[xs > 0]
    got = _vec(L.rto_log, xs)
    err = _ulp_err(got, np.log(xs.astype(np.float64)))
    assert err.max() <= 2.0, err.max()
    assert L.rto_log(C.c_float(0.0)) == -np.inf and np.isnan(L.rto_log(C.c_float(-1.0))) and L.rto_log(C.c_float(1.0)) == 0.0


def test_cos(oracle):
    L = oracle.lib()
    xs = np.concatenate([np.random.RandomState(2).uniform(0, 6.2831855, 30000), [0.0, 6.2831852, np.pi / 2, np.pi]]).astype(np.float32)
    got = _vec(L.rto_cos, xs)
    want = np.cos(xs.astype(np.float64))
    assert np.abs(got - want).max() <= 2.0 ** -22          # absolute: near the zeros of cos ulps are meaningless
    big = np.abs(want) > 0.1
    assert _ulp_err(got[big], want[big]).max() <= 2.5
    assert L.rto_cos(C.c_float(0.0)) == 1.0


def test_exp2_and_pow(oracle):
    L = oracle.lib()
    ts = np.random.RandomState(3).uniform(-126, 127, 20000).astype(np.float32)
    got = _vec(L.rto_exp2, ts)
    assert _ulp_err(got, np.exp2(ts.astype(np.float64))).max() <= 2.0
    assert L.rto_exp2(C.c_float(0.0)) == 1.0 and L.rto_exp2(C.c_float(10.0)) == 1024.0 and L.rto_exp2(C.c_float(-200.0)) == 0.0
    assert L.rto_exp2(C.c_float(128.0)) == np.inf
    # pow as the shader uses it: x in [0,1] ^ 0.35, x ^ 500
    xs = np.random.RandomState(4).uniform(0, 1, 5000).astype(np.float32)
    got = np.array([L.rto_pow(C.c_float(float(x)), C.c_float(0.35)) for x in xs], np.float32)
    want = xs.astype(np.float64) ** np.float64(np.float32(0.35))
    assert np.abs(got - want).max() <= 4e-7
    assert L.rto_pow(C.c_float(0.0), C.c_float(0.35)) == 0.0 and L.rto_pow(C.c_float(1.0), C.c_float(0.35)) == 1.0
    assert L.rto_pow(C.c_float(1.0), C.c_float(500.0)) == 1.0 and L.rto_pow(C.c_float(0.0), C.c_float(500.0)) == 0.0
    got = np.array([L.rto_pow(C.c_float(float(x)), C.c_float(500.0)) for x in xs[xs > 0.9]], np.float64)
    want = xs[xs > 0.9].astype(np.float64) ** 500.0
    assert (np.abs(got - want) <= 2e-4 * want + 1e-30).all()   # 500 x the log's rounding error, as GLSL's definition implies


def test_unit_disk_acceptance_without_the_square_root():
    """The trace kernel accepts a lens sample when dot(p, p) <= 1 + 2^-23 instead of wgsl:222's length(p) <= 1.0:
    with a correctly rounded square root the two tests agree for every float (sqrt is monotonic, sqrt(1 + 2^-23) rounds
    to 1, sqrt(1 + 2^-22) to 1 + 2^-23).  Checked on every float of [0.5, 2] and samples of the rest of the range."""
    one = np.float32(1.0)
    bound = np.nextafter(one, np.float32(2.0))
    assert float(bound) == 1.00000011920928955078125
    xs = np.arange(np.float32(0.5).view(np.uint32), np.float32(2.0).view(np.uint32) + 1, dtype=np.uint32).view(np.float32)
    assert np.array_equal(np.sqrt(xs) <= one, xs <= bound)
    rest = np.concatenate([np.float32([0.0, 1e-45, 1e-38, 1e-20, 0.25, 0.4999999, 2.0000002, 4.0, 8.0, 3e38, np.inf]),
                           np.random.RandomState(5).uniform(0, 8, 200000).astype(np.float32)])
    assert np.array_equal(np.sqrt(rest) <= one, rest <= bound)
    assert not (np.float32(np.nan) <= bound)                    # NaN is rejected by both forms
