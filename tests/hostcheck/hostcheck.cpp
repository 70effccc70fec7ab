// hostcheck.cpp -- TEST INFRASTRUCTURE.  Compiles the kernels' shared scalar header
// (pythonphot_b200/csrc/bphot_math.cuh) with g++ so that its arithmetic and the mmm control code can
// be checked on a machine without a GPU.  Never loaded by the product package.
#include <string.h>

#include <algorithm>
#include <vector>

#include "../../pythonphot_b200/csrc/bphot_math.cuh"

// mmm through the block index of the sorted vector (bphot_math.cuh IndexedView, built like aper_kernel.cu sort_index
// builds it on the device: per block of K keys the first key and the exclusive prefix sums of d and d^2).
// T = float when f32 != 0 (keys are rounded to fp32 first, like an fp32 image).
template <typename T>
static int mmm_indexed(const double* v, int n, int K, double readnoise, int minsky, int mxiter, double* out3) {
    std::vector<T> s;
    s.reserve(n);
    for (int i = 0; i < n; i++)
        if (v[i] == v[i]) s.push_back((T)v[i]);
    std::sort(s.begin(), s.end());
    const int ns = (int)s.size();
    int gs = 0;
    while ((1 << gs) < K) gs++;
    const int nb = (ns + K - 1) / K;
    std::vector<double> PT(2 * (nb + 2), 0.0);
    std::vector<T> bmin(nb + 1);
    bphot::IndexedView<T> V;
    V.k = s.data(); V.PT = PT.data(); V.bmin = bmin.data(); V.gs = gs; V.n = ns;
    V.skymid = ns > 0 ? 0.5 * (double)s[(ns - 1) / 2] + 0.5 * (double)s[ns / 2] : 0.0;
    double a = 0.0, q = 0.0;
    for (int b = 0; b <= nb; b++) {
        PT[2 * b] = a; PT[2 * b + 1] = q;
        if (b < nb) bmin[b] = s[b * K];
        double ba = 0.0, bq = 0.0;
        for (int j = b * K; j < (b + 1) * K && j < ns; j++) { const double d = (double)s[j] - V.skymid; ba = ba + d; bq = bq + d * d; }
        a = a + ba; q = q + bq;
    }
    bphot::MmmResult R = bphot::mmm_sorted(
        V, [&](int lo, int hi, double, double* s1, double* s2) { V.delta_sums(lo, hi, s1, s2); },
        [&](int lo, int hi) { return V.plain_sum(lo, hi); }, [&](double x, bool strict) { return V.first(x, strict); }, ns, readnoise,
        minsky, mxiter);
    out3[0] = R.skymod; out3[1] = R.sigma; out3[2] = R.skew;
    return R.iters;
}

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
int hc_mmm_indexed(const double* v, int n, int K, int f32, double readnoise, int minsky, int mxiter, double* out3) {
    return f32 ? mmm_indexed<float>(v, n, K, readnoise, minsky, mxiter, out3) : mmm_indexed<double>(v, n, K, readnoise, minsky, mxiter, out3);
}
// IndexedView::first against the plain binary searches, on every distinct key and its neighbours
int hc_indexed_first_ok(const double* v, int n, int K) {
    std::vector<double> s(v, v + n);
    std::sort(s.begin(), s.end());
    int gs = 0;
    while ((1 << gs) < K) gs++;
    const int nb = (n + K - 1) / K;
    std::vector<double> PT(2 * (nb + 2), 0.0), bmin(nb + 1);
    for (int b = 0; b < nb; b++) bmin[b] = s[b * K];
    bphot::IndexedView<double> V;
    V.k = s.data(); V.PT = PT.data(); V.bmin = bmin.data(); V.gs = gs; V.n = n; V.skymid = 0.0;
    const double* p = s.data();
    auto at = [p](int i) -> double { return p[i]; };
    for (int i = 0; i < n; i++)
        for (int d = -1; d <= 1; d++) {
            const double x = d == 0 ? s[i] : nextafter(s[i], d < 0 ? -1e300 : 1e300);
            if (V.first(x, true) != bphot::upper_bound_gt(at, n, x)) return 0;
            if (V.first(x, false) != bphot::lower_bound_ge(at, n, x)) return 0;
        }
    return 1;
}
}
