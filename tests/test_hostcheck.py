"""The kernels' shared scalar header (bphot_math.cuh), compiled for the host, against the golden
vectors of the unmodified reference and against the oracle on adversarial inputs.  CPU only."""
import numpy as np
import pytest

from oracle import oracle
from tests import hostcheck
from tests.conftest import assert_close, load_golden


def test_pixwt_golden():
    g = load_golden("golden_pixwt.npz")
    for i in range(g["xc"].size):
        m = g["pid"] == i
        w = hostcheck.pixwt(g["xc"][i], g["yc"][i], g["r"][i], g["x"][m], g["y"][m])
        assert_close(w, g["w"][m], rtol=1e-12, atol=1e-13, what="pixwt %d" % i)
    assert_close(hostcheck.pixwt(3.5, 3.5, 2.0, g["half_x"], g["half_y"]), g["half_w"], rtol=1e-12, atol=1e-13)


def test_psfmodel_golden():
    g = load_golden("golden_psfmodel.npz")
    o = hostcheck.psfmodel(g["xx"], g["yy"], g["gauss"], g["psf"])
    for row, key, tol in ((0, "v", 1e-12), (1, "dvx", 1e-11), (2, "dvy", 1e-11), (3, "f", 1e-12),
                          (9, "z", 1e-12), (10, "dfdx", 1e-12), (11, "dfdy", 1e-12)):
        assert_close(o[row], g[key], rtol=tol, atol=tol, what=key)
    assert_close(o[4:9], g["pder"], rtol=1e-11, atol=1e-12, what="pder")


def test_mmm_golden():
    g = load_golden("golden_mmm.npz")
    off = 0
    for n, out in zip(g["lens"], g["out"]):
        v = g["vec"][off:off + n]
        off += n
        got = hostcheck.mmm(v, readnoise=out[3], minsky=int(out[4]))
        assert_close(got[:3], out[:3], rtol=1e-11, atol=1e-11, what="mmm n=%d" % n)


@pytest.mark.parametrize("seed", range(6))
def test_mmm_bracket_search_equals_index_walk(seed):
    """The binary-searched brackets (mmm_walk_min/max) against the oracle's literal one-index-at-a-time
    walk, on data full of ties (integer-valued, few distinct levels, outliers on either side)."""
    rng = np.random.default_rng(100 + seed)
    for trial in range(300):
        n = int(rng.integers(20, 400))
        kind = trial % 5
        if kind == 0:
            v = np.round(rng.normal(300, rng.uniform(0.5, 6), n))
        elif kind == 1:
            v = rng.integers(0, 4, n).astype(float) * rng.uniform(1, 50)
        elif kind == 2:
            v = np.round(rng.normal(100, 3, n)); v[rng.integers(0, n, n // 6 + 1)] += np.round(rng.exponential(200, n // 6 + 1))
        elif kind == 3:
            v = np.round(rng.normal(100, 3, n)); v[rng.integers(0, n, n // 5 + 1)] -= np.round(rng.exponential(80, n // 5 + 1))
        else:
            v = rng.normal(50, 2, n)
        rn = float(rng.choice([0.0, 0.0, 4.0]))
        ms = int(rng.choice([20, 20, 35]))
        want = oracle.mmm(v, readnoise=rn, minsky=ms, return_info=True)
        got = hostcheck.mmm(v, readnoise=rn, minsky=ms)
        assert got[3] == want[3], "iteration count differs (trial %d kind %d): %r vs %r" % (trial, kind, got, want)
        assert_close(got[:3], want[:3], rtol=1e-10, atol=1e-10, what="mmm trial %d kind %d" % (trial, kind))


def test_inv3_matches_oracle():
    rng = np.random.default_rng(5)
    for t in range(500):
        a = rng.normal(size=(3, 3)) * 10.0 ** rng.integers(-3, 4, size=(1, 3))
        c = a @ a.T if t % 2 == 0 else a  # SPD -> Cholesky route, general -> LU route
        if t % 2 == 0:
            c = (c + c.T) / 2
        if t % 50 == 49:
            c[2] = c[1] * 2.0  # singular
            c[:, 2] = c[:, 1] * 2.0
        got, st = hostcheck.inv3(c, exact=True)  # reference-order arithmetic (PK_RECIP=0): LAPACK's bits
        want, st2 = oracle.inv3(c)
        assert st == st2
        if st == 0:
            assert np.array_equal(got, want), "inv3 differs from the oracle bit-for-bit at trial %d" % t
        # the product build multiplies by three reciprocals instead of dividing nine times (PK_RECIP=1): same answer to
        # within the conditioning of the matrix
        got, st = hostcheck.inv3(c)
        assert st == st2
        if st == 0:
            err = np.abs(got - want).max() / np.abs(want).max()
            assert err <= 64 * np.finfo(float).eps * np.linalg.cond(c), "trial %d: %g (cond %g)" % (t, err, np.linalg.cond(c))


def test_daoerf_axis_factors_are_bit_identical():
    """The kernels evaluate daoerf per column and per row of the fitting box (daoerf.py:49-72 is separable);
    the product must reproduce the per-pixel formula bit for bit."""
    rng = np.random.default_rng(5)
    xx, yy = rng.uniform(-7, 7, 20000), rng.uniform(-7, 7, 20000)
    for gauss in ([1000.0, 0.0, 0.0, 1.76, 1.70], [6107.2, -0.0126, -0.0037, 1.76, 1.7689], [3.0, 0.3, -0.2, 0.9, 2.5]):
        assert hostcheck.daoerf_separable(xx, yy, gauss)


@pytest.mark.parametrize("seed", range(3))
def test_mmm_through_index_equals_direct_sums(seed):
    """aper_mmm_kernel's view of the sorted annulus (16-ary search tree + prefix sums at every 16th key, IndexedView)
    against the oracle: iteration counts exact, values 1e-10; rank queries exact on tie-ridden data."""
    rng = np.random.default_rng(500 + seed)
    for trial in range(400):
        n = int(rng.integers(20, 4096 if trial % 3 else 300))
        kind = trial % 5
        if kind == 0:
            v = np.round(rng.normal(300, rng.uniform(0.5, 6), n))
        elif kind == 1:
            v = rng.integers(0, 4, n).astype(float) * rng.uniform(1, 50)
        elif kind == 2:
            v = rng.normal(300, 9, n); v[rng.integers(0, n, n // 6 + 1)] += rng.exponential(3000, n // 6 + 1)
        elif kind == 3:
            v = np.round(rng.normal(100, 3, n)); v[rng.integers(0, n, n // 5 + 1)] -= np.round(rng.exponential(80, n // 5 + 1))
        else:
            v = rng.normal(50, 2, n)
        f32 = bool(trial % 2)
        if f32:
            v = v.astype(np.float32).astype(np.float64)
        rn = float(rng.choice([0.0, 0.0, 4.0]))
        ms = int(rng.choice([20, 20, 35]))
        want = oracle.mmm(v, readnoise=rn, minsky=ms, return_info=True)
        got = hostcheck.mmm_indexed(v, f32=f32, readnoise=rn, minsky=ms)
        assert got[3] == want[3], "iteration count differs (trial %d kind %d): %r vs %r" % (trial, kind, got, want)
        assert_close(got[:3], want[:3], rtol=1e-10, atol=1e-10, what="indexed mmm trial %d kind %d" % (trial, kind))
        if trial % 8 == 0:
            assert hostcheck.indexed_first_ok(v)


def test_mmm_through_index_golden():
    g = load_golden("golden_mmm.npz")
    off = 0
    for n, out in zip(g["lens"], g["out"]):
        v = g["vec"][off:off + n]
        off += n
        if np.isfinite(v).sum() > 4096:
            continue
        got = hostcheck.mmm_indexed(v, readnoise=out[3], minsky=int(out[4]))
        # north_star's bar, not 1e-10: the bimodal golden (n=85, half the keys near 300, half near 5000) has
        # mean^2 / variance = 1e6, which amplifies ANY change of summation order to ~4e-10 in sigma
        assert_close(got[:3], out[:3], rtol=1e-9, atol=1e-10, what="indexed mmm n=%d" % n)


def test_warp_sum16_emulation_has_the_butterfly_bits():
    """bphot_device.cuh warp_sum16 (the fits' 13-14 sums in one transposing reduction) claims the SAME additions in the same
    order as sixteen warp_sum() butterflies.  The device code cannot run here; this is its data flow, lane by lane, in numpy
    (fp64 addition is commutative, so a + b on one lane and b + a on its partner are the same bits)."""
    rng = np.random.default_rng(16)
    for trial in range(50):
        v = rng.normal(size=(32, 16)) * 10.0 ** rng.integers(-8, 9, size=(32, 16))  # v[lane, value]
        # reference: one butterfly per value, lane distances 16, 8, 4, 2, 1
        b = v.copy()
        for o in (16, 8, 4, 2, 1):
            b = b + b[np.arange(32) ^ o]
        assert all(np.array_equal(b[0], b[l]) for l in range(32))
        # warp_sum16: level h = 8, 4, 2, 1 at lane distance 2h; a lane whose bit is set keeps the upper half of its slots
        w = v.copy()
        lanes = np.arange(32)
        for h in (8, 4, 2, 1):
            bit = 2 * h
            up = (lanes & bit) != 0
            send = np.where(up[:, None], w[:, :h], w[:, h:2 * h])
            keep = np.where(up[:, None], w[:, h:2 * h], w[:, :h])
            w = w.copy()
            w[:, :h] = keep + send[lanes ^ bit]
        tot = w[:, 0] + w[lanes ^ 1, 0]
        got = np.array([tot[2 * k] for k in range(16)])  # value k is complete in lanes 2k and 2k + 1
        assert np.array_equal(tot[0::2], tot[1::2])
        assert np.array_equal(got, b[0]), "trial %d: the transposing reduction changed a sum's bits" % trial


def test_rcp_bounded_newton_steps_reach_one_ulp():
    """pkfit_kernel.cu rcp_bounded: a reciprocal estimate good to 2^-23 (what rcp.approx.ftz.f64 promises) followed by two
    Newton steps in fused multiply-adds.  Emulated with exact rational arithmetic rounded once per fma, over its stated domain
    [1, 2^17] and with the worst seeds the estimate may return: the result is within one ulp of the correctly rounded 1/d."""
    from fractions import Fraction

    def fma(a, b, c):
        return float(Fraction(a) * Fraction(b) + Fraction(c))  # one rounding

    rng = np.random.default_rng(23)
    worst = 0.0
    for d in np.concatenate([rng.uniform(1.0, 2.0, 300), 2.0 ** rng.uniform(0, 17, 300), [1.0, 65537.0, 131072.0]]):
        d = float(d)
        for rel in (2.0 ** -23, -2.0 ** -23, 0.0, 3e-8):
            r = float(np.float64((1.0 / d) * (1.0 + rel)))
            e = fma(-d, r, 1.0)
            r = fma(r, e, r)
            e = fma(-d, r, 1.0)
            r = fma(r, e, r)
            exact = Fraction(1) / Fraction(d)
            ulp = np.spacing(1.0 / d)
            worst = max(worst, abs(float(Fraction(r) - exact)) / ulp)
    assert worst <= 1.0, worst


def test_pixwt_one_atan_form_against_the_reference_order():
    """PW_ONE_ATAN (atan(a) - atan(b) as one atan where x^2 + y0 y1 > 0) against the header compiled with the reference's
    two-atan form, over circles of every size the API admits -- radii below 1/2 exercise the fall-back -- centred anywhere in a
    pixel: the overlap area moves by ~1.3 eps r^2, the rounding of the sector area r^2 atan / 2 itself (1.6e-14 at r = 8; the
    apertures are held to 1e-9 relative on fluxes of 1e4 - 1e6 ADU)."""
    rng = np.random.default_rng(41)
    for r in (0.2, 0.45, 0.7, 1.0, 2.5, 3.0, 5.0, 8.0, 20.0, 75.0):
        for _ in range(20):
            xc, yc = rng.uniform(100, 101, 2)
            n = int(r) + 3
            gx, gy = np.meshgrid(np.arange(100 - n, 102 + n, dtype=float), np.arange(100 - n, 102 + n, dtype=float))
            a = hostcheck.pixwt(xc, yc, r, gx.ravel(), gy.ravel())
            b = hostcheck.pixwt(xc, yc, r, gx.ravel(), gy.ravel(), exact=True)
            assert float(np.abs(a - b).max()) <= 4 * np.finfo(float).eps * max(1.0, r * r), (r, float(np.abs(a - b).max()))
            assert abs(a.sum() - np.pi * r * r) < 1e-9 * max(1.0, r * r)  # the weights tile the circle
