// hostcheck.cpp -- TEST INFRASTRUCTURE.  Compiles the kernels' shared scalar header
// (pythonphot_b200/csrc/bphot_math.cuh) with g++ so that its arithmetic and the mmm control code can
// be checked on a machine without a GPU.  Never loaded by the product package.
#include <string.h>

#include <algorithm>
#include <vector>

#include "../../pythonphot_b200/csrc/bphot_math.cuh"

extern "C" {

void hc_pixwt(double xc, double yc, double r, const double* x, const double* y, int n, double* out) {
    for (int i = 0; i < n; i++) out[i] = bphot::pixwt(xc, yc, r, x[i], y[i]);
}

// out rows (SoA, 12 x n) as bphot_psfmodel_device
void hc_psfmodel(const double* xx, const double* yy, int n, const double* g, const double* psf, int npsf, double* out) {
    const double half = (double)(npsf - 1) / 2.;
    for (int i = 0; i < n; i++) {
        double pder[5], dfdx = 0, dfdy = 0, dvx = 0, dvy = 0, z = 0, val = 0;
        const double f = bphot::daoerf(xx[i], yy[i], g, pder);
        const double tx = 2. * xx[i] + half, ty = 2. * yy[i] + half;
        if (tx >= 1. && tx <= npsf - 2. && ty >= 1. && ty <= npsf - 2.) {
            z = bphot::rinter(psf, npsf, npsf, tx, ty, &dfdx, &dfdy);
            val = bphot::psf_value(xx[i], yy[i], g, psf, npsf, &dvx, &dvy);
        }
        out[0 * n + i] = val; out[1 * n + i] = dvx; out[2 * n + i] = dvy; out[3 * n + i] = f;
        for (int k = 0; k < 5; k++) out[(4 + k) * n + i] = pder[k];
        out[9 * n + i] = z; out[10 * n + i] = dfdx; out[11 * n + i] = dfdy;
    }
}

// 1 when daoerf_axis/daoerf_from_axes reproduce daoerf_xy bit for bit on every point
int hc_daoerf_separable(const double* xx, const double* yy, int n, const double* g) {
    for (int i = 0; i < n; i++) {
        double p1, p2, q1, q2, fx, ex, fy, ey;
        const double f = bphot::daoerf_xy(xx[i], yy[i], g, &p1, &p2);
        bphot::daoerf_axis(xx[i], g[1], g[3], &fx, &ex);
        bphot::daoerf_axis(yy[i], g[2], g[4], &fy, &ey);
        const double h = bphot::daoerf_from_axes(g[0], fx, ex, fy, ey, &q1, &q2);
        if (memcmp(&f, &h, 8) || memcmp(&p1, &q1, 8) || memcmp(&p2, &q2, 8)) return 0;
    }
    return 1;
}

int hc_inv3(double* c9) {
    double c[3][3];
    for (int i = 0; i < 9; i++) c[i / 3][i % 3] = c9[i];
    const int st = bphot::inv3(c);
    for (int i = 0; i < 9; i++) c9[i] = c[i / 3][i % 3];
    return st;
}

// mmm on an unsorted vector: drop NaNs, sort, run the shared control code with serial reductions.
int hc_mmm(const double* v, int n, double readnoise, int minsky, int mxiter, double* out3) {
    std::vector<double> s;
    s.reserve(n);
    for (int i = 0; i < n; i++)
        if (v[i] == v[i]) s.push_back(v[i]);
    std::sort(s.begin(), s.end());
    const double* p = s.data();
    auto at = [p](int i) -> double { return p[i]; };
    auto dsum = [&](int lo, int hi, double mid, double* a, double* q) {
        double sa = 0, sq = 0;
        for (int i = lo; i < hi; i++) { const double d = p[i] - mid; sa = sa + d; sq = sq + d * d; }
        *a = sa; *q = sq;
    };
    auto psum = [&](int lo, int hi) -> double {
        double sa = 0;
        for (int i = lo; i < hi; i++) sa = sa + p[i];
        return sa;
    };
    const int ns = (int)s.size();
    auto first = [&](double v, bool strict) -> int {
        return strict ? bphot::upper_bound_gt(at, ns, v) : bphot::lower_bound_ge(at, ns, v);
    };
    bphot::MmmResult R = bphot::mmm_sorted(at, dsum, psum, first, ns, readnoise, minsky, mxiter);
    out3[0] = R.skymod; out3[1] = R.sigma; out3[2] = R.skew;
    return R.iters;
}
}
