// bphot_math.cuh -- scalar fp64 building blocks of the photometry kernels.
//
// Every function here is the per-pixel / per-star arithmetic of one reference routine
// (djones1040/PythonPhot @ 7992eb6), written in the reference's own order of operations so
// that discrete decisions (pixel membership, clipping brackets, convergence) agree with it.
// The translation unit is compiled with --fmad=false: no a*b+c contraction, each operation
// rounds once like numpy does.  Transcendentals (erf, exp, atan, log10, sqrt) are CUDA libm.
//
//   pixwt()        PythonPhot/pixwt.py:8-235   (vector branch, the one aper.py:328 calls)
//   daoerf()       PythonPhot/daoerf.py:49-90
//   rinter()       PythonPhot/rinter.py:98-195 (deriv=True)
//   psf_value()    PythonPhot/dao_value.py:95-107 (per point; the edge guard is the caller's)
//   inv3()         scipy.linalg.inv on the 3x3 normal matrix (pkfit*.py) -- potrf/potri when
//                  symmetric positive definite, getrf/getri otherwise
//   mmm_*          bracket searches of PythonPhot/mmm.py:199-248 on a sorted vector
#pragma once
#include <math.h>
#include <stdint.h>

#ifndef PK_RSQRT
#define PK_RSQRT 1  // device fits: 1 / sqrt(s) of the Cholesky pivots and of chi's normalisation by rsqrt() instead of sqrt + division
#endif
#ifndef PK_RECIP
#define PK_RECIP 1  // fits: divisions whose divisor is fixed for a pixel / an iteration become multiplications by one reciprocal
#endif

#if defined(__CUDACC__)
#define BP_HD __host__ __device__ __forceinline__
#define BP_D __device__ __forceinline__
#define BP_HDM __host__ __device__ __forceinline__  // member functions
#define BP_HD_COLD __host__ __device__ __noinline__
#else
#define BP_HD static inline
#define BP_HDM inline
#define BP_HD_COLD static
#endif

namespace bphot {

// ------------------------------------------------------------------ pixwt
// pixwt.py:50-62.  Plain division: x == 0 gives +-inf and atan(+-inf) = +-pi/2 (IEEE), as numpy.
// PW_ONE_ATAN: atan(a) - atan(b) = atan((a - b) / (1 + a b)) while a b > -1, i.e. x^2 + y0 y1 > 0 -- every arc of a pixel
// the circle crosses (its end points lie on one side of the axis) and every pixel beyond a circle of radius > 1/2: one
// atan and one division instead of two of each (they were 28 % of aper_sums' instructions).  The value moves by a few
// 1e-16 (the difference of two rounded atans is no more exact than this form); x == 0 gives atan(0) = 0 = pi/2 - pi/2.
#ifndef PW_ONE_ATAN
#define PW_ONE_ATAN 1
#endif
#ifndef PW_FAST_RCP
#define PW_FAST_RCP 1
#endif
BP_HD double pw_arc(double x, double y0, double y1, double r) {
#if PW_ONE_ATAN
    const double den = x * x + y0 * y1;
#if defined(__CUDA_ARCH__) && PW_FAST_RCP
    if (den > 1e-6 && den < 1e12) {  // the quotient by the hardware's reciprocal estimate + two Newton steps (~1 ulp, 6 instructions for ~30)
        double q;
        asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(q) : "d"(den));
        double e = fma(-den, q, 1.0);
        q = fma(q, e, q);
        e = fma(-den, q, 1.0);
        q = fma(q, e, q);
        return 0.5 * r * r * atan((y1 - y0) * x * q);
    }
#endif
    if (den > 0.0) return 0.5 * r * r * atan((y1 - y0) * x / den);
#endif
    return 0.5 * r * r * (atan(y1 / x) - atan(y0 / x));
}
// pixwt.py:64-72
BP_HD double pw_chord(double x, double y0, double y1) { return 0.5 * x * (y1 - y0); }

// pixwt.py:112-213: area between the line X = x and the circle, from y0 to y1.
// The reference's 3 x 3 case table (zone of y0: <= -yh, inside, >= yh; zone of y1: <= -yh, <= yh,
// above) always reduces to  [arc] + [chord] + [arc]  with the terms added left to right; the table
// is folded into endpoint selects so that the device code has TWO arc call sites instead of
// eleven (the inlined fp64 atan pairs of the case table overflowed the instruction cache).
BP_HD double pw_oneside(double x, double y0, double y1, double r) {
    const bool beyond = fabs(x) >= r;  // pixwt.py:93,129: the line misses the circle -> one arc y0..y1
    const double yh = beyond ? 0.0 : sqrt(r * r - x * x);
    const int z0 = beyond ? 0 : ((y0 <= -yh) ? 0 : ((y0 < yh) ? 1 : 2));
    const int z1 = beyond ? 0 : ((y1 <= -yh) ? 0 : ((y1 <= yh) ? 1 : 2));
    const bool has_a1 = (z0 != 1);                // leading arc starts at y0
    const bool has_a2 = (z1 != 1) && (z1 != z0);  // trailing arc ends at y1
    const bool has_c = !(z0 == z1 && z0 != 1);    // chord between them
    // Arc end points.  Most calls need no arc or exactly one (leading OR trailing); both only when the unit segment
    // spans the whole chord (2*yh < 1, near tangency).  A warp runs a call site if ANY lane needs it, so the arcs go
    // through a PRIMARY site (the leading arc if there is one, else the trailing one) and a rarely taken SECONDARY
    // site (trailing arc when both exist): half the fp64 atan work of two always-taken sites, same values, same
    // left-to-right order of addition.
    const double a1_end = (z1 == z0) ? y1 : (z0 == 0 ? -yh : yh);
    const double a2_beg = (z1 == 2) ? yh : -yh;
    const bool any_arc = has_a1 || has_a2;
    double prim = 0.0, sec = 0.0;
    if (any_arc) prim = pw_arc(x, has_a1 ? y0 : a2_beg, has_a1 ? a1_end : y1, r);
    if (has_a1 && has_a2) sec = pw_arc(x, a2_beg, y1, r);
    double res = 0.0;
    if (has_a1) res = prim;
    if (has_c) {
        const double cs = (z0 == 1) ? y0 : (z0 == 0 ? -yh : yh);
        const double ce = (z1 == 1) ? y1 : (z1 == 2 ? yh : -yh);
        const double ch = pw_chord(x, cs, ce);
        res = has_a1 ? res + ch : ch;
    }
    if (has_a2) res = res + (has_a1 ? sec : prim);
    return res;
}

// One of the four Intarea terms (pixwt.py:229-235) of the unit pixel centred (x, y):
// side 0..3 in the reference's order of addition.
BP_HD double pixwt_side(double xc, double yc, double r, double x, double y, int side) {
    const double x0 = (x - 0.5) - xc, y0 = (y - 0.5) - yc;
    const double x1 = (x + 0.5) - xc, y1 = (y + 0.5) - yc;
    const double X = (side == 0) ? x1 : (side == 1) ? y1 : (side == 2) ? -x0 : -y0;
    const double A = (side == 0) ? y0 : (side == 1) ? -x1 : (side == 2) ? -y1 : x0;
    const double B = (side == 0) ? y1 : (side == 1) ? -x0 : (side == 2) ? -y0 : x1;
    return pw_oneside(X, A, B, r);
}

// Geometric class of the unit pixel whose centre is (x1, y1) = |offset from the circle centre|:
// 1 = every corner inside the circle (weight 1), 0 = no point of the pixel inside (weight 0),
// 2 = the circle crosses the pixel (needs pixwt).  The 1e-9 px^2 guard band sends anything
// within rounding distance of tangency to pixwt, so classes 0 and 1 are exact statements.
BP_HD int pixel_class(double x1, double y1, double r) {
    const double r2 = r * r;
    const double fx = x1 + 0.5, fy = y1 + 0.5;
    if (fx * fx + fy * fy <= r2 - 1e-9) return 1;
    const double nx = x1 > 0.5 ? x1 - 0.5 : 0.0, ny = y1 > 0.5 ? y1 - 0.5 : 0.0;
    if (nx * nx + ny * ny >= r2 + 1e-9) return 0;
    return 2;
}

// pixwt.py:8-48 + Intarea pixwt.py:215-235: area of the unit pixel centred (x,y) inside the
// circle of radius r centred (xc,yc).
BP_HD double pixwt(double xc, double yc, double r, double x, double y) {
    const double x0 = (x - 0.5) - xc, y0 = (y - 0.5) - yc;
    const double x1 = (x + 0.5) - xc, y1 = (y + 0.5) - yc;
    return pw_oneside(x1, y0, y1, r) + pw_oneside(y1, -x1, -x0, r) +
           pw_oneside(-x0, -y1, -y0, r) + pw_oneside(-y0, x0, x1, r);
}

// -------------------------------------------------------------- PSF model
// daoerf.py:76-90
BP_HD double gauss_cdf(double t) { return 0.5 + 0.5 * erf(t * (1. / sqrt(2.))); }

// daoerf.py:49-74.  a = (height, x0, y0, sigx, sigy).  Returns f; pder[0..4] when non-null.
BP_HD double daoerf(double x, double y, const double* a, double* pder) {
    const double norm = 2.506628275;  // the reference's truncated sqrt(2 pi), daoerf.py:49
    const double u2 = (x - a[1] + 0.5) / a[3], u1 = (x - a[1] - 0.5) / a[3];
    const double v2 = (y - a[2] + 0.5) / a[4], v1 = (y - a[2] - 0.5) / a[4];
    const double fx = norm * a[3] * (gauss_cdf(u2) - gauss_cdf(u1));
    const double fy = norm * a[4] * (gauss_cdf(v2) - gauss_cdf(v1));
    if (pder) {
        const double up = exp(-0.5 * (u2 * u2)), um = exp(-0.5 * (u1 * u1));
        const double vp = exp(-0.5 * (v2 * v2)), vm = exp(-0.5 * (v1 * v1));
        pder[0] = fx * fy;
        pder[1] = a[0] * fy * (-up + um);
        pder[2] = a[0] * fx * (-vp + vm);
        pder[3] = a[0] * fy * (fx / a[3] + u1 * um - u2 * up);
        pder[4] = a[0] * fx * (fy / a[4] + v1 * vm - v2 * vp);
    }
    return a[0] * fx * fy;
}

// Only what the fit needs from daoerf: f, df/dx0, df/dy0 (pder[1], pder[2]).
BP_HD double daoerf_xy(double x, double y, const double* a, double* p1, double* p2) {
    const double norm = 2.506628275;
    const double u2 = (x - a[1] + 0.5) / a[3], u1 = (x - a[1] - 0.5) / a[3];
    const double v2 = (y - a[2] + 0.5) / a[4], v1 = (y - a[2] - 0.5) / a[4];
    const double fx = norm * a[3] * (gauss_cdf(u2) - gauss_cdf(u1));
    const double fy = norm * a[4] * (gauss_cdf(v2) - gauss_cdf(v1));
    const double up = exp(-0.5 * (u2 * u2)), um = exp(-0.5 * (u1 * u1));
    const double vp = exp(-0.5 * (v2 * v2)), vm = exp(-0.5 * (v1 * v1));
    *p1 = a[0] * fy * (-up + um);
    *p2 = a[0] * fx * (-vp + vm);
    return a[0] * fx * fy;
}

// daoerf is separable: fx and (-up + um) depend only on the pixel's column, fy and (-vp + vm) only on
// its row (daoerf.py:49-72).  One axis factor, for offset d from the star along that axis, centre c
// (a[1] or a[2]) and width s (a[3] or a[4]):  f = norm*s*(cdf(u2)-cdf(u1)),  e = -exp(-u2^2/2) + exp(-u1^2/2).
// A fitting box has ~12 columns + ~12 rows but ~80 pixels, so the kernels evaluate 4 erf/exp pairs per
// column and row instead of per pixel; daoerf_from_axes() then reproduces daoerf_xy() bit for bit.
BP_HD void daoerf_axis(double d, double c, double s, double* f, double* e) {
    const double norm = 2.506628275;
    const double u2 = (d - c + 0.5) / s, u1 = (d - c - 0.5) / s;
    *f = norm * s * (gauss_cdf(u2) - gauss_cdf(u1));
    const double up = exp(-0.5 * (u2 * u2)), um = exp(-0.5 * (u1 * u1));
    *e = -up + um;
}
BP_HD double daoerf_from_axes(double a0, double fx, double ex, double fy, double ey, double* p1, double* p2) {
    *p1 = a0 * fy * ex;
    *p2 = a0 * fx * ey;
    return a0 * fx * fy;
}

// One row of rinter.py:161-184: cubic through the flat neighbours k-1..k+2 of the table.
// Flat indices past the end read the last element (rinter.py:117-135 pads the table with it).  Negative flat indices
// (points above the first table row) wrap like numpy's: `wrap` is the length of the array the reference indexes
// (the table, or its padded copy); with wrap = 0 they read NaN.
BP_HD double rinter_at(const double* p, long plen, long idx, long wrap) {
    if (idx < 0) idx += wrap;
    if (idx < 0) return NAN;
    return p[idx < plen ? idx : plen - 1];
}
BP_HD void rinter_row(const double* p, long plen, long k, double xd, double* val, double* der, long wrap = 0) {
    const double pm = rinter_at(p, plen, k - 1, wrap);
    const double p0 = rinter_at(p, plen, k, wrap);
    const double p1 = rinter_at(p, plen, k + 1, wrap);
    const double p2 = rinter_at(p, plen, k + 2, wrap);
    const double c1 = 0.5 * (p1 - pm);
    const double c2 = 2. * p1 + pm - 0.5 * (5. * p0 + p2);
    const double c3 = 0.5 * (3. * (p0 - p1) + p2 - pm);
    *val = xd * (xd * (xd * c3 + c2) + c1) + p0;
    *der = xd * (xd * c3 * 3. + 2. * c2) + c1;
}

// rinter.py:98-195 for one point of an [nr, nc] table (row-major, x = column).
BP_HD double rinter(const double* p, int nr, int nc, double x, double y, double* dfdx, double* dfdy, long wrap = 0) {
    const long plen = (long)nr * nc;
    const long i = (long)x, j = (long)y;  // rinter.py:98 (truncation)
    const double xd = x - (double)i, yd = y - (double)j;
    const long km = (long)nc * (j - 1) + i;  // rinter.py:101-104
    double ym, y0, y1, y2, gm, g0, g1, g2;
    rinter_row(p, plen, km, xd, &ym, &gm, wrap);
    rinter_row(p, plen, km + nc, xd, &y0, &g0, wrap);
    rinter_row(p, plen, km + 2L * nc, xd, &y1, &g1, wrap);
    rinter_row(p, plen, km + 3L * nc, xd, &y2, &g2, wrap);
    {  // rinter.py:185-188: the y-cubic of the four row derivatives
        const double d1 = 0.5 * (g1 - gm);
        const double d2 = 2. * g1 + gm - 0.5 * (5. * g0 + g2);
        const double d3 = 0.5 * (3. * (g0 - g1) + g2 - gm);
        *dfdx = yd * (yd * (yd * d3 + d2) + d1) + g0;
    }
    const double d1 = 0.5 * (y1 - ym);  // rinter.py:190-195
    const double d2 = 2. * y1 + ym - 0.5 * (5. * y0 + y2);
    const double d3 = 0.5 * (3. * (y0 - y1) + y2 - ym);
    *dfdy = yd * (yd * d3 * 3. + 2. * d2) + d1;
    return yd * (yd * (yd * d3 + d2) + d1) + y0;
}

// dao_value.py:95-107 for one point that already passed the edge guard.
BP_HD double psf_value(double xx, double yy, const double* gauss, const double* psf, int npsf,
                       double* dvdx, double* dvdy) {
    const double half = (double)(npsf - 1) / 2.;
    double p1, p2, dfdx, dfdy;
    const double e = daoerf_xy(xx, yy, gauss, &p1, &p2);
    const double v = rinter(psf, npsf, npsf, 2. * xx + half, 2. * yy + half, &dfdx, &dfdy);
    *dvdx = 2. * dfdx - p1;
    *dvdy = 2. * dfdy - p2;
    return e + v;
}

// 1 / sqrt(s) for s > 0: rsqrt() on the device (PK_RSQRT), 1.0 / sqrt(s) on the host and with the knob off
BP_HD double bp_rsqrt(double s) {
#if defined(__CUDA_ARCH__) && PK_RSQRT
    return rsqrt(s);
#else
    return 1.0 / sqrt(s);
#endif
}

// ------------------------------------------------------------ 3x3 inverse
// The general case of inv3 (getrf + getri), on a flat copy of the matrix.  Normal matrices of a fit are symmetric positive
// definite, so this is the cold path: out of line on the device, where its ~400 unrolled instructions would otherwise sit
// in the middle of the fit loop (the loop's code footprint is what the fit warps stall on).
BP_HD_COLD int inv3_lu(double* c9) {
    double (*c)[3] = reinterpret_cast<double (*)[3]>(c9);
    // LU with partial pivoting, then inv(U) and the getri back-substitution.
    double a[3][3];
    int piv[3];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) a[i][j] = c[i][j];
    for (int j = 0; j < 3; j++) {
        int p = j;
        double best = fabs(a[j][j]);
        for (int i = j + 1; i < 3; i++)
            if (fabs(a[i][j]) > best) { best = fabs(a[i][j]); p = i; }
        piv[j] = p;
        if (a[p][j] == 0.0) return 1;
        if (p != j)
            for (int k = 0; k < 3; k++) { const double t = a[j][k]; a[j][k] = a[p][k]; a[p][k] = t; }
        for (int i = j + 1; i < 3; i++) a[i][j] = a[i][j] / a[j][j];
        for (int i = j + 1; i < 3; i++)
            for (int k = j + 1; k < 3; k++) a[i][k] = a[i][k] - a[i][j] * a[j][k];
    }
    for (int j = 0; j < 3; j++) {  // inv(U) in place (trti2)
        a[j][j] = 1.0 / a[j][j];
        const double ajj = -a[j][j];
        for (int jj = 0; jj < j; jj++) {
            const double t = a[jj][j];
            for (int i = 0; i < jj; i++) a[i][j] = a[i][j] + t * a[i][jj];
            a[jj][j] = t * a[jj][jj];
        }
        for (int i = 0; i < j; i++) a[i][j] = a[i][j] * ajj;
    }
    for (int j = 2; j >= 0; j--) {  // inv(A) L = inv(U) (getri)
        double w[3] = {0., 0., 0.};
        for (int i = j + 1; i < 3; i++) { w[i] = a[i][j]; a[i][j] = 0.0; }
        for (int k = j + 1; k < 3; k++)
            for (int i = 0; i < 3; i++) a[i][j] = a[i][j] - a[i][k] * w[k];
    }
    for (int j = 1; j >= 0; j--) {
        const int p = piv[j];
        if (p != j)
            for (int i = 0; i < 3; i++) { const double t = a[i][j]; a[i][j] = a[i][p]; a[i][p] = t; }
    }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) c[i][j] = a[i][j];
    return 0;
}

// What scipy.linalg.inv does to the normal matrix: Cholesky (potrf + potri) when the matrix is
// symmetric and positive definite, LU with partial pivoting (getrf + getri) otherwise.
// Returns 0 on success, 1 when singular (scipy raises LinAlgError -> the fit returns NaN).
BP_HD int inv3(double c[3][3]) {
    bool sym = (c[0][1] == c[1][0]) && (c[0][2] == c[2][0]) && (c[1][2] == c[2][1]);
    if (sym) {
        double l10, l20, l21;
        double s = c[0][0];
        if (s > 0.0) {
#if PK_RECIP
            // three reciprocals instead of nine divisions (every quotient below has l00, l11 or l22 as its divisor); the
            // results move by an ulp or two, far inside the 1e-9 the fits are held to.  On the device the reciprocal of
            // the pivot's root is one rsqrt (the root itself is never needed).
            const double m00 = bp_rsqrt(s);
            l10 = c[1][0] * m00;
            l20 = c[2][0] * m00;
#else
            const double l00 = sqrt(s);
            l10 = c[1][0] / l00;
            l20 = c[2][0] / l00;
#endif
            s = c[1][1] - l10 * l10;
            if (s > 0.0) {
#if PK_RECIP
                const double m11 = bp_rsqrt(s);
                l21 = (c[2][1] - l20 * l10) * m11;
#else
                const double l11 = sqrt(s);
                l21 = (c[2][1] - l20 * l10) / l11;
#endif
                s = c[2][2] - l20 * l20;
                s = s - l21 * l21;
                if (s > 0.0) {
                    // M = inv(L), lower triangular; inv(A) = M^T M
#if PK_RECIP
                    const double m22 = bp_rsqrt(s);
                    const double m10 = -(l10 * m00) * m11;
                    const double m21 = -(l21 * m11) * m22;
                    double t = 0.0 + l20 * m00;
                    t = t + l21 * m10;
                    const double m20 = -t * m22;
#else
                    const double l22 = sqrt(s);
                    const double m00 = 1.0 / l00, m11 = 1.0 / l11, m22 = 1.0 / l22;
                    const double m10 = -(l10 * m00) / l11;
                    const double m21 = -(l21 * m11) / l22;
                    double t = 0.0 + l20 * m00;
                    t = t + l21 * m10;
                    const double m20 = -t / l22;
#endif
                    double a00 = 0.0 + m00 * m00; a00 = a00 + m10 * m10; a00 = a00 + m20 * m20;
                    double a01 = 0.0 + m10 * m11; a01 = a01 + m20 * m21;
                    const double a02 = 0.0 + m20 * m22;
                    double a11 = 0.0 + m11 * m11; a11 = a11 + m21 * m21;
                    const double a12 = 0.0 + m21 * m22;
                    const double a22 = 0.0 + m22 * m22;
                    c[0][0] = a00; c[0][1] = a01; c[0][2] = a02;
                    c[1][0] = a01; c[1][1] = a11; c[1][2] = a12;
                    c[2][0] = a02; c[2][1] = a12; c[2][2] = a22;
                    return 0;
                }
            }
        }
    }
    double t[9];
    for (int i = 0; i < 9; i++) t[i] = c[i / 3][i % 3];
    const int st = inv3_lu(t);
    for (int i = 0; i < 9; i++) c[i / 3][i % 3] = t[i];
    return st;
}

// ------------------------------------------------------------------- mmm
// Sorted-vector searches.  `at(i)` widens element i to double.
template <typename Get>
BP_HD int lower_bound_ge(const Get& at, int n, double v) {  // first i with at(i) >= v (n if none)
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (at(mid) >= v) hi = mid; else lo = mid + 1;
    }
    return lo;
}
template <typename Get>
BP_HD int upper_bound_gt(const Get& at, int n, double v) {  // first i with at(i) > v (n if none)
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (at(mid) > v) hi = mid; else lo = mid + 1;
    }
    return lo;
}

// mmm.py:199-215: where the one-index-at-a-time walk of the LOWER edge stops.  `minimm` is the
// index just below the accepted set (may be -1).  The walk stops at the first index i (stepping from minimm)
// with i == -1, i == nlast, or at(i) <= cut1 <= at(i+1); in a sorted vector those i form [lb-1, ub-1]
// (SURVEY App. A), where lb = first(cut1, false) and ub = first(cut1, true):
//   first(v, strict) = index of the first element > v (strict) or >= v (not strict), nsky if none.
// The walk is split in two so that the searches of both edges can be issued together (one thread per star waits
// for memory, not for arithmetic): *_pre reads the two keys next to the edge and decides whether the reference
// takes its "done" shortcut (sum / sumsq untouched) and in which direction the edge moves; *_post turns the
// search result e = first(cut, ...) - 1 into the new edge.
struct MmmEdge {
    bool tst;   // mmm.py:199 tst_min / mmm.py:224 tst_max: the edge grows outward
    bool done;  // the reference leaves this edge where it is
};
template <typename Get>
BP_HD MmmEdge mmm_min_pre(const Get& at, int minimm, double cut1) {
    const double a = at(minimm + 1), b = at(minimm > 0 ? minimm : 0);
    MmmEdge E;
    E.tst = a >= cut1;
    E.done = E.tst && ((minimm == -1) || (b < cut1));
    return E;
}
BP_HD int mmm_min_post(int nsky, int minimm, bool tst_min, int e) {  // e = first(cut1, tst_min) - 1
    if (tst_min) {  // grow downward: first i <= minimm-1 inside the bracket set, else -1
        const int r = e < (minimm - 1) ? e : (minimm - 1);
        return r < -1 ? -1 : r;
    }
    const int r = e > (minimm + 1) ? e : (minimm + 1);  // shrink upward, else nlast
    return r > nsky - 1 ? nsky - 1 : r;
}
// mmm.py:224-241: the UPPER edge.  `maximm` is the last accepted index.
template <typename Get>
BP_HD MmmEdge mmm_max_pre(const Get& at, int nsky, int maximm, double cut2) {
    const int nlast = nsky - 1;
    const double a = at(maximm), b = at((maximm + 1 < nlast) ? maximm + 1 : nlast);
    MmmEdge E;
    E.tst = a <= cut2;
    E.done = E.tst && ((maximm == nlast) || (b > cut2));
    return E;
}
BP_HD int mmm_max_post(int nsky, int maximm, bool tst_max, int e) {  // e = first(cut2, !tst_max) - 1
    if (tst_max) {  // grow upward
        const int r = e > (maximm + 1) ? e : (maximm + 1);
        return r > nsky - 1 ? nsky - 1 : r;
    }
    const int r = e < (maximm - 1) ? e : (maximm - 1);  // shrink downward
    return r < -1 ? -1 : r;
}

// mmm.py:106-314 on an already sorted, NaN-free vector of nsky elements.
//   at(i)                         element i widened to double
//   delta_sums(lo, hi, mid, &a, &q)  a = sum(at(i)-mid), q = sum((at(i)-mid)^2) over [lo, hi)
//   plain_sum(lo, hi)             sum(at(i)) over [lo, hi)
//   first2(v1, strict1, need1, v2, strict2, need2, &r1, &r2)   two rank queries at once: r = first index with
//                                 at(i) > v (strict) / at(i) >= v, nsky if none; a query with need = false is skipped
//   sums3(lo0, hi0, need0, lo1, hi1, need1, lo2, hi2, mid, out5)   the three range sums of one clipping pass at once:
//                                 out5 = delta_sums over [lo0, hi0) and over [lo1, hi1) (skipped when need = false),
//                                 then plain_sum over [lo2, hi2)
// The functors let a whole warp share the reductions on the device (every lane must receive the same
// bits) and a plain loop stand in on the host (tests/hostcheck).  Failure modes keep the reference's
// in-band codes (sigma = -1).
// log10 of a pixel count (mmm.py:186).  Every lane of the mmm warp evaluates it with the same argument once per
// clipping iteration (~100 fp64 instructions each time, 13 % of stage 2), so the device reads it from a table that
// log10_table_kernel fills with the very same libm call: identical bits, one L1 hit.
constexpr int MMM_LOG10_TAB = 8192;
#if defined(__CUDACC__)
__device__ double g_log10_tab[MMM_LOG10_TAB + 1];
__global__ void log10_table_kernel() {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= MMM_LOG10_TAB; i += gridDim.x * blockDim.x) g_log10_tab[i] = log10((double)i);
}
#endif
BP_HD double mmm_log10n(int n) {
#if defined(__CUDA_ARCH__)
    if (n >= 0 && n <= MMM_LOG10_TAB) return g_log10_tab[n];
#endif
    return log10((double)n);
}

// ------------------------------------------------ indexed sorted vector
// (double)x for a float without the F2F.F64.F32 conversion: a normal float's sign, exponent (re-biased by 1023 - 127) and
// mantissa are moved into place by integer instructions.  Exact; zero, subnormal, inf and NaN take F2F.  Measured on
// B200: in the stage-1 index (32 conversions per thread feeding dependent fp64 adds) it took 4.8 ms per 1 M stars off
// the kernel; in the indexed mmm's 16-key group loads it cost 1.9 ms (the branch per key), so only the index uses it.
BP_HD double bp_widen(float x) {
#if defined(__CUDA_ARCH__)
    const unsigned b = __float_as_uint(x), e = (b >> 23) & 0xffu;
    if (e == 0u || e == 0xffu) return (double)x;
    const unsigned hi = (b & 0x80000000u) | (((b >> 3) & 0x0fffffffu) + 0x38000000u);
    return __hiloint2double((int)hi, (int)(b << 29));
#else
    return (double)x;
#endif
}
BP_HD double bp_widen(double x) { return x; }

// read-only loads: __ldg on the device, plain on the host (tests/hostcheck)
template <typename T>
BP_HD T bp_ld(const T* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}
BP_HD void bp_ld2(const double* p, double* x, double* y) {  // p is 16-byte aligned
#if defined(__CUDA_ARCH__)
    const double2 v = __ldg(reinterpret_cast<const double2*>(p));
    *x = v.x; *y = v.y;
#else
    *x = p[0]; *y = p[1];
#endif
}

// Sorted vector + its index (aper_kernel.cu sort_index builds it), read by ONE thread.  Keys k[0..n) in a 64-byte
// aligned row; L2[j] = k[16 j] and L1[i] = k[256 i] (a 16-ary search tree laid out so that every node is one aligned
// 64-byte group); PT[2g], PT[2g+1] = sums over keys [0, 16 g) of d = key - skymid and d^2.  Everything a thread needs
// per step is a handful of INDEPENDENT aligned 16-key groups, so the loads of one step are in flight together.
template <typename T>
struct IndexedView {
    const T* __restrict__ k;
    const T* __restrict__ L1;
    const T* __restrict__ L2;
    const double* __restrict__ PT;
    int n;
    double skymid;
    BP_HDM double operator()(int i) const { return (double)bp_ld(k + i); }
    BP_HDM static void load16(const T* p, double (&v)[16]) {
#if defined(__CUDA_ARCH__)
        if constexpr (sizeof(T) == 4) {
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const float4 f = __ldg(reinterpret_cast<const float4*>(p) + q);
                v[4 * q] = (double)f.x; v[4 * q + 1] = (double)f.y; v[4 * q + 2] = (double)f.z; v[4 * q + 3] = (double)f.w;
            }
        } else {
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const double2 f = __ldg(reinterpret_cast<const double2*>(p) + q);
                v[2 * q] = f.x; v[2 * q + 1] = f.y;
            }
        }
#else
        for (int q = 0; q < 16; q++) v[q] = (double)p[q];
#endif
    }
    // sums of d and d^2 over keys [0, i): the prefix at the group boundary below i, then the group's first i % 16 keys
    BP_HDM void prefix(int i, double* s1, double* s2) const {
        const int g = i >> 4, c = i & 15;
        double a, q;
        bp_ld2(PT + 2 * g, &a, &q);
        if (c) {
            double v[16];
            load16(k + 16 * g, v);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
            for (int j = 0; j < 15; j++)
                if (j < c) { const double d = v[j] - skymid; a = a + d; q = q + d * d; }
        }
        *s1 = a; *s2 = q;
    }
    // The three range sums of one clipping pass (mmm_sorted's sums3).  Only what is needed is loaded: measured on B200, issuing
    // all six prefix evaluations unconditionally (so that their loads overlap) was slower, 6.1 -> 8.3 ms per 1 M stars -- the
    // kernel is bound by the bytes it pulls through L2, and most passes near convergence move one edge or none.
    BP_HDM void sums3(int lo0, int hi0, bool need0, int lo1, int hi1, bool need1, int lo2, int hi2, double* o5) const {
        if (need0) delta_sums(lo0, hi0, &o5[0], &o5[1]);
        if (need1) delta_sums(lo1, hi1, &o5[2], &o5[3]);
        o5[4] = plain_sum(lo2, hi2);
    }
    BP_HDM void delta_sums(int lo, int hi, double* s1, double* s2) const {
        double a1, q1, a0, q0;
        prefix(hi, &a1, &q1);
        prefix(lo, &a0, &q0);
        *s1 = a1 - a0; *s2 = q1 - q0;
    }
    BP_HDM double plain_sum(int lo, int hi) const {
        if (hi <= lo) return 0.0;
        double a1, q1, a0, q0;
        prefix(hi, &a1, &q1);
        prefix(lo, &a0, &q0);
        return (a1 - a0) + (double)(hi - lo) * skymid;
    }
    // number of entries t = t0 .. 15 of the aligned group at p that exist (t < limit) and FAIL the predicate
    BP_HDM static int count_fail(const T* p, int t0, int limit, double v, bool strict) {
        double x[16];
        load16(p, x);
        int c = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int t = 0; t < 16; t++) c += (t >= t0 && t < limit && !(strict ? x[t] > v : x[t] >= v)) ? 1 : 0;
        return c;
    }
    // Two rank queries in lockstep (first index with key > v / key >= v, n if none): the failing entries of a sorted
    // node are a prefix, so their COUNT is the child to descend into -- three levels, three aligned groups per query.
    BP_HDM void first2(double v1, bool st1, bool need1, double v2, bool st2, bool need2, int* r1, int* r2) const {
        const int n1 = (n + 255) >> 8, n2 = (n + 15) >> 4;
        int c1a = 0, c1b = 0;
        {
            double x[16];
            load16(L1, x);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
            for (int t = 0; t < 16; t++) {
                c1a += (t < n1 && !(st1 ? x[t] > v1 : x[t] >= v1)) ? 1 : 0;
                c1b += (t < n1 && !(st2 ? x[t] > v2 : x[t] >= v2)) ? 1 : 0;
            }
        }
        const bool go1 = need1 && c1a > 0, go2 = need2 && c1b > 0;  // c1 == 0: key[0] passes, the answer is 0
        const int g1 = 16 * (c1a - 1), g2 = 16 * (c1b - 1);         // L2 group whose entry 0 is the last failing L1 key
        int j1 = 0, j2 = 0;
        if (go1) j1 = g1 + count_fail(L2 + g1, 1, n2 - g1, v1, st1);
        if (go2) j2 = g2 + count_fail(L2 + g2, 1, n2 - g2, v2, st2);
        int f1 = 0, f2 = 0;                                         // k[16 j] is the last failing L2 key
        if (go1) f1 = 16 * j1 + 1 + count_fail(k + 16 * j1, 1, n - 16 * j1, v1, st1);
        if (go2) f2 = 16 * j2 + 1 + count_fail(k + 16 * j2, 1, n - 16 * j2, v2, st2);
        *r1 = f1; *r2 = f2;
    }
};

struct MmmResult {
    double skymod, sigma, skew;
    int iters;
};
BP_HD double mmm_nan() { return (double)NAN; }

template <typename Get, typename DSums, typename PSum, typename First2, typename Sums3>
BP_HD MmmResult mmm_sorted(const Get& at, const DSums& delta_sums, const PSum& plain_sum, const First2& first2, const Sums3& sums3,
                           int nsky, double readnoise, int minsky, int mxiter) {
    MmmResult R;
    R.iters = 0;
    R.skew = 0.0;
    if (nsky < minsky) {  // mmm.py:109-112
        R.skymod = mmm_nan(); R.sigma = -1.0;
        return R;
    }
    const int nlast = nsky - 1;
    const double skymid = 0.5 * at((nsky - 1) / 2) + 0.5 * at(nsky / 2);  // mmm.py:121
    double cut1, cut2;
    {
        const double a = skymid - at(0), b = at(nlast) - skymid;  // mmm.py:123-127
        const double c = a < b ? a : b;
        cut2 = skymid + c;
        cut1 = skymid - c;
    }
    int gmin = 0, gmax = 0;
    first2(cut1, false, true, cut2, true, true, &gmin, &gmax);
    gmax = gmax - 1;
    if (gmin > gmax) {  // mmm.py:134-138
        R.skymod = 0.0; R.sigma = -1.0;
        return R;
    }
    double sum, sumsq;
    delta_sums(gmin, gmax + 1, skymid, &sum, &sumsq);  // mmm.py:141-142
    int maximm = gmax, minimm = gmin - 1;      // mmm.py:144-145
    double skymed = 0.5 * at((int)((minimm + maximm + 1) / 2.0)) + 0.5 * at((int)((minimm + maximm) / 2.0 + 1));
    double skymn = sum / (maximm - minimm);
    double sigma = sqrt(sumsq / (maximm - minimm) - skymn * skymn);
    skymn = skymn + skymid;
    double skymod = (skymed < skymn) ? 3. * skymed - 2. * skymn : skymn;  // mmm.py:157-159

    int niter = 0;
    double clamp = 1.0, old = 0.0;
    bool redo = true;
    while (redo) {  // mmm.py:168-301
        niter++;
        if (niter > mxiter || (maximm - minimm) < minsky) {  // mmm.py:170-182
            R.skymod = skymod; R.sigma = -1.0; R.iters = niter;
            return R;
        }
        double r = mmm_log10n(maximm - minimm);  // mmm.py:186-187
        const double r2 = (-0.1042 * r + 1.1695) * r + 0.8895;
        r = 2. > r2 ? 2. : r2;
        const double cut = r * sigma + 0.5 * fabs(skymn - skymod);  // mmm.py:191-193
        cut1 = skymod - cut;
        cut2 = skymod + cut;
        redo = false;
        // mmm.py:199-248: both edges.  The upper walk reads only the vector, maximm and cut2, so it
        // can be taken before the lower edge's sums are applied; the sums then share one call site.
        const MmmEdge E0 = mmm_min_pre(at, minimm, cut1), E1 = mmm_max_pre(at, nsky, maximm, cut2);
        int e0 = 0, e1 = 0;
        first2(cut1, E0.tst, !E0.done, cut2, !E1.tst, !E1.done, &e0, &e1);
        const int newmin = E0.done ? minimm : mmm_min_post(nsky, minimm, E0.tst, e0 - 1);
        const int newmax = E1.done ? maximm : mmm_max_post(nsky, maximm, E1.tst, e1 - 1);
        const bool moved = !E0.done, grew = E0.tst, moved1 = !E1.done, grew1 = E1.tst;
        // lower edge: grown range (newmin, minimm], shrunk range (minimm, newmin]
        // upper edge: grown range (maximm, newmax], shrunk range (newmax, maximm]
        const int from0 = grew ? newmin : minimm, to0 = grew ? minimm : newmin;
        const int from1 = grew1 ? maximm : newmax, to1 = grew1 ? newmax : maximm;
        // robust median: mean of the central 20 % of the NEW window (mmm.py:271-289; np.round = half-to-even).  Its index
        // range depends on the new edges only, so its sum is asked for together with the edge sums (one thread per star
        // then waits for memory once, not three times).
        const int ns = newmax - newmin;  // mmm.py:253-257
        const double center = (newmin + 1 + newmax) / 2.;
        const double side = rint(0.2 * (newmax - newmin)) / 2. + 0.25;
        double j = rint(center - side), k = rint(center + side);
        if (readnoise != 0.0 && ns >= minsky) {  // mmm.py:280-287
            const double L = rint(center - 0.25), M = rint(center + 0.25), Rn = 0.25 * readnoise;
            while (j > 0 && k < ns - 1 && ((at((int)L) - at((int)j)) < Rn || (at((int)k) - at((int)M)) < Rn)) {
                j -= 1; k += 1;
            }
        }
        int ji = (int)j, ki = (int)k + 1;
        if (ji < 0) ji = 0;
        if (ki > nsky) ki = nsky;
        if (ns < minsky) { ji = 0; ki = 0; }  // the pass returns before the median is used
        double o5[5] = {0., 0., 0., 0., 0.};
        sums3(from0 + 1, to0 + 1, moved, from1 + 1, to1 + 1, moved1, ji, ki, skymid, o5);
        if (moved) {
            if (grew) { sum = sum + o5[0]; sumsq = sumsq + o5[1]; }
            else { sum = sum - o5[0]; sumsq = sumsq - o5[1]; }
            redo = true;
        }
        if (moved1) {
            if (grew1) { sum = sum + o5[2]; sumsq = sumsq + o5[3]; }
            else { sum = sum - o5[2]; sumsq = sumsq - o5[3]; }
            redo = true;
        }
        minimm = newmin;
        maximm = newmax;
        if (ns < minsky) {
            R.skymod = skymod; R.sigma = -1.0; R.iters = niter;
            return R;
        }
        skymn = sum / ns;  // mmm.py:259-263
        double var = sumsq / ns - skymn * skymn;
        if (var < 0) var = 0;
        sigma = sqrt(var);
        skymn = skymn + skymid;
        skymed = o5[4] / (k - j + 1);
        const double dmod = (skymed < skymn) ? 3. * skymed - 2. * skymn - skymod : skymn - skymod;  // mmm.py:294-296
        if (dmod * old < 0) clamp = 0.5 * clamp;  // mmm.py:299-301
        skymod = skymod + clamp * dmod;
        old = dmod;
    }
    R.skew = (skymn - skymod) / (sigma > 1. ? sigma : 1.);  // mmm.py:306
    R.skymod = skymod; R.sigma = sigma; R.iters = niter;
    return R;
}

}  // namespace bphot
