// hostcheck.cpp -- TEST INFRASTRUCTURE.  Compiles the kernels' shared scalar header
// (pythonphot_b200/csrc/bphot_math.cuh) with g++ so that its arithmetic and the mmm control code can
// be checked on a machine without a GPU.  Never loaded by the product package.
#include <string.h>

#include <algorithm>
#include <vector>

#include "../../pythonphot_b200/csrc/bphot_math.cuh"

// The index of a sorted vector as aper_kernel.cu sort_index lays it out (bphot_math.cuh IndexedView): 16-ary search tree
// L1 / L2 and exclusive prefix sums of d, d^2 at every 16th key.  Rows are padded to whole groups of 16.
template <typename T>
struct HostIndex {
    std::vector<T> k, L1, L2;
    std::vector<double> PT;
    bphot::IndexedView<T> V;
    explicit HostIndex(std::vector<T> sorted) : k(std::move(sorted)) {
        const int ns = (int)k.size();
        const int ng = (ns + 15) / 16;
        k.resize((size_t)(ng + 1) * 16, (T)0);
        L1.assign(16, (T)0); L2.assign(4096 / 16 + 16, (T)0); PT.assign(2 * (size_t)(ng + 2), 0.0);
        if ((int)L2.size() < ng + 16) L2.resize(ng + 16, (T)0);
        V.skymid = ns > 0 ? 0.5 * (double)k[(ns - 1) / 2] + 0.5 * (double)k[ns / 2] : 0.0;
        double a = 0.0, q = 0.0;
        for (int g = 0; g <= ng; g++) {
            PT[2 * g] = a; PT[2 * g + 1] = q;
            if (g < ng) L2[g] = k[16 * g];
            for (int j = 16 * g; j < 16 * g + 16 && j < ns; j++) { const double d = (double)k[j] - V.skymid; a = a + d; q = q + d * d; }
        }
        for (int i = 0; i < 16 && 256 * i < ns; i++) L1[i] = k[256 * i];
        V.k = k.data(); V.L1 = L1.data(); V.L2 = L2.data(); V.PT = PT.data(); V.n = ns;
    }
};

// mmm through that index, as aper_mmm_kernel runs it (one thread per star).  T = float: keys rounded to fp32 first,
// like an fp32 image.
template <typename T>
static int mmm_indexed(const double* v, int n, double readnoise, int minsky, int mxiter, double* out3) {
    std::vector<T> s;
    s.reserve(n);
    for (int i = 0; i < n; i++)
        if (v[i] == v[i]) s.push_back((T)v[i]);
    std::sort(s.begin(), s.end());
    const int ns = (int)s.size();
    if (ns > 4096) return -1;
    HostIndex<T> H(s);
    const bphot::IndexedView<T>& V = H.V;
    bphot::MmmResult R = bphot::mmm_sorted(
        V, [&](int lo, int hi, double, double* s1, double* s2) { V.delta_sums(lo, hi, s1, s2); },
        [&](int lo, int hi) { return V.plain_sum(lo, hi); },
        [&](double v1, bool st1, bool need1, double v2, bool st2, bool need2, int* r1, int* r2) { V.first2(v1, st1, need1, v2, st2, need2, r1, r2); },
        [&](int lo0, int hi0, bool need0, int lo1, int hi1, bool need1, int lo2, int hi2, double, double* o5) { V.sums3(lo0, hi0, need0, lo1, hi1, need1, lo2, hi2, o5); },
        ns, readnoise, minsky, mxiter);
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
    auto first2 = [&](double v1, bool st1, bool need1, double v2, bool st2, bool need2, int* r1, int* r2) {
        if (need1) *r1 = first(v1, st1);
        if (need2) *r2 = first(v2, st2);
    };
    auto sums3 = [&](int lo0, int hi0, bool need0, int lo1, int hi1, bool need1, int lo2, int hi2, double mid, double* o5) {
        if (need0) dsum(lo0, hi0, mid, &o5[0], &o5[1]);
        if (need1) dsum(lo1, hi1, mid, &o5[2], &o5[3]);
        o5[4] = psum(lo2, hi2);
    };
    bphot::MmmResult R = bphot::mmm_sorted(at, dsum, psum, first2, sums3, ns, readnoise, minsky, mxiter);
    out3[0] = R.skymod; out3[1] = R.sigma; out3[2] = R.skew;
    return R.iters;
}
int hc_mmm_indexed(const double* v, int n, int f32, double readnoise, int minsky, int mxiter, double* out3) {
    return f32 ? mmm_indexed<float>(v, n, readnoise, minsky, mxiter, out3) : mmm_indexed<double>(v, n, readnoise, minsky, mxiter, out3);
}
// IndexedView::first2 against the plain binary searches, on every key and its floating-point neighbours
int hc_indexed_first_ok(const double* v, int n) {
    std::vector<double> s(v, v + n);
    std::sort(s.begin(), s.end());
    if (n > 4096) return -1;
    HostIndex<double> H(s);
    const double* p = s.data();
    auto at = [p](int i) -> double { return p[i]; };
    for (int i = 0; i < n; i++)
        for (int d = -1; d <= 1; d++) {
            const double x = d == 0 ? s[i] : nextafter(s[i], d < 0 ? -1e300 : 1e300);
            int r1 = -7, r2 = -7;
            H.V.first2(x, true, true, x, false, true, &r1, &r2);
            if (r1 != bphot::upper_bound_gt(at, n, x) || r2 != bphot::lower_bound_ge(at, n, x)) return 0;
        }
    return 1;
}
}
