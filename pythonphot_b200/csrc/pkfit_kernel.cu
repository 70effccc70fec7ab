// pkfit_kernel.cu -- DAOPHOT PSF peak fit, one warp per star (sm_100a).
//
// Replaces the per-star Python loops of
//   variant 0  PythonPhot/pkfit.py:132-458                (analytic noise model, optional recentring)
//   variant 1  PythonPhot/pkfit_noise.py:131-462          (noise + mask images, recentring)
//   variant 2  PythonPhot/pkfit_norecent_noise.py:132-531 (noise + mask images, forced position)
//   variant 3  PythonPhot/pkfit_norecenter.py:346-657     (analytic noise model, forced position; Python-2 semantics)
// and their callee dao_value.dao_value (dao_value.py:9-121 -> daoerf.py, rinter.py).
//
// Per star, one warp runs the damped Gauss-Newton loop.  Pixels of the fitting box are
// enumerated in row-major order (ballot-ranked, so the pixel lists are ordered exactly like the
// reference's np.where output).  The PSF model is the integrated Gaussian, evaluated once per
// COLUMN and once per ROW of the box (daoerf is separable; bphot_math.cuh daoerf_axis), plus the
// bicubic residual look-up per pixel (table staged in shared memory).  The 3-vector and 3x3
// normal matrix are fp64 butterfly reductions (every lane ends with identical bits, so the
// convergence logic is warp-uniform); the 3x3 solve runs redundantly in registers.
//
// The forced fit (variant 2) never moves x, y, so everything that depends only on the pixel --
// the good-pixel list, model value and derivatives, radial weight, sharpness kernel -- is
// computed ONCE per star; an iteration is then ~10 fp64 operations per pixel plus the reductions.
// The reference recomputes all of it every iteration from the same inputs, so the bits agree.
#include "bphot_device.cuh"
#include "bphot_params.h"

namespace bphot {

// Out-of-line warp reductions for the fit kernels (PK_NOINLINE_SUMS=1): ~45 inlined fp64 butterfly reductions were a
// third of the kernel's 80 KB of code and the fit warps stalled on instruction fetch.
#ifndef PK_U
#define PK_U 1  // 32-pixel chunks of the forced fit's pixel list handled per pass (2 and 3 measured slower: registers)
#endif
#ifndef PK_NOINLINE_SUMS
#define PK_NOINLINE_SUMS 0
#endif
#if PK_NOINLINE_SUMS
__device__ __noinline__ double pk_sum(double v) { return warp_sum(v); }
__device__ __noinline__ double pk_min(double v) { return warp_min(v); }
__device__ __noinline__ double pk_max(double v) { return warp_max(v); }
#else
BP_D double pk_sum(double v) { return warp_sum(v); }
BP_D double pk_min(double v) { return warp_min(v); }
BP_D double pk_max(double v) { return warp_max(v); }
#endif


#ifndef PK_MINB_FORCED
#define PK_MINB_FORCED 5  // resident CTAs per SM the forced variant's register budget is compiled for
#endif
#ifndef PK_MINB_FIT
#define PK_MINB_FIT 5  // resident CTAs per SM the recentring variants are compiled for (config5 fits: 323.8 ms at 1, 297.6 at 4, 287.7 at 5 per 2.5 M stars)
#endif
constexpr int PK_WARPS = 4;
constexpr int PK_THREADS = PK_WARPS * 32;
constexpr int PK_NARR_FIT = 10;     // per-warp scratch arrays of `cap` doubles, recentring variants
#ifndef PK_PSF_GLOBAL
#define PK_PSF_GLOBAL 0  // 1 = the residual look-up table is read from global memory (L1-resident, 13 KB) instead of a per-CTA shared copy
#endif
#ifndef PK_FMA_MODEL
#define PK_FMA_MODEL PK_RECIP  // the look-up table's cubics in fused multiply-adds (rinter_f)
#endif
#ifndef PK_FMA_LOOP
#define PK_FMA_LOOP 0  // forced fit, bit mask: 1 = sky taken off the pixel at set-up and the residual as one fma, 2 = dfs / sigma^2 stored, 4 = s3 / schi / e^8 as fmas
#endif
#ifndef PK_FAST_RCP
#define PK_FAST_RCP 1  // 1 / (1 + e^8) of the residual down-weighting by rcp.approx + two Newton steps (6 instructions, ~1 ulp) instead of a division (~30)
#endif
#ifndef PK_PREFETCH
#define PK_PREFETCH 0  // forced fit: prefetch the fitting box of image, noise and mask at the start of a star; measured 14.7 -> 16.0 ms (the other warps already hide those misses), off
#endif
#ifndef PK_FIXPOINT
#define PK_FIXPOINT 1  // forced fit: 1 = leave the loop when an iteration starts from the same (scale, chiold) bits as the previous one; 2 = also within PK_FIX_TOL twice in a row
#endif
#ifndef PK_FIX_TOL
#define PK_FIX_TOL 1e-14
#endif
#ifndef PK_TSUM
#define PK_TSUM 1  // the 13-14 warp sums of an iteration through warp_sum16 (one transposing reduction) instead of one butterfly each
#endif
#ifndef PK_HOIST
#define PK_HOIST 1  // forced fit: W0/sigsq and dfs*dfs/sigsq (operands fixed at set-up) are divided once per star, not per iteration
#endif
constexpr int PK_NARR_FORCED = PK_HOIST ? 9 : 7;   // forced variant

struct FitOut {
    double errmag, chi, sharp, scale, x, y, chi2;
    int niter;
};

struct FitBox {  // pkfit.py:168-177: fitting box, int truncation then clamp
    long long ixlo, iylo;
    int ixx, iyy, npix;
};
BP_D FitBox fit_box(const PkfitParams& P, double x, double y) {
    FitBox b;
    long long ixlo = (long long)(x - P.radius); if (ixlo < 0) ixlo = 0;
    long long iylo = (long long)(y - P.radius); if (iylo < 0) iylo = 0;
    long long ixhi = (long long)(x + P.radius) + 1; if (ixhi > P.ncol - 1) ixhi = P.ncol - 1;
    long long iyhi = (long long)(y + P.radius) + 1; if (iyhi > P.nrow - 1) iyhi = P.nrow - 1;
    const long long ixx = ixhi - ixlo + 1, iyy = iyhi - iylo + 1;
    const bool ok = ixx > 0 && iyy > 0 && ixx < 32768 && iyy < 32768 && ixx * iyy < (1LL << 30);
    b.ixlo = ixlo; b.iylo = iylo;
    b.ixx = ok ? (int)ixx : 0; b.iyy = ok ? (int)iyy : 0;
    b.npix = b.ixx * b.iyy;
    return b;
}

// Column and row factors of the integrated Gaussian for the whole box (daoerf.py:49-72): lanes
// 0..ixx-1 take the columns, the next iyy lanes the rows.  AX = [FX | EX | FY | EY], each `acap` long.
// Returns, for lane i < ixx + iyy <= 32, the lane's term of rho = dx^2/sigx^2 + dy^2/sigy^2 (pkfit.py:292): column i's
// dx^2/sigx^2 or row (i - ixx)'s dy^2/sigy^2 -- the same operations on the same operands as the per-pixel expression, taken once per
// column and row; pixels fetch the two terms by shuffle.
BP_D double fill_axes(const FitBox& b, double x, double y, const double* gauss, double* AX, int acap, int lane) {
    const int nitem = b.ixx + b.iyy;
    double qv = 0.;
    if (lane < nitem && nitem <= 32) {
        const bool col = lane < b.ixx;
        const int k = col ? lane : lane - b.ixx;
        const double d = col ? (double)(k + b.ixlo) - x : (double)(k + b.iylo) - y;
        const double s = col ? gauss[3] : gauss[4];
        qv = d * d / (s * s);
    }
    for (int i = lane; i < nitem; i += 32) {
        const bool col = i < b.ixx;
        const int k = col ? i : i - b.ixx;
        if (k >= acap) continue;
        const double d = col ? (double)(k + b.ixlo) - x : (double)(k + b.iylo) - y;
        double f, e;
        daoerf_axis(d, col ? gauss[1] : gauss[2], col ? gauss[3] : gauss[4], &f, &e);
        double* dst = AX + (col ? 0 : 2 * acap);
        dst[k] = f;
        dst[acap + k] = e;
    }
    return qv;
}

// rinter.py:98-195 as bphot_math.cuh's rinter(), with the cubics in fused multiply-adds (PK_RECIP: the fits are held to
// 1e-9, the look-up table's unit kernel to its goldens' last bits, so only the fits use this form).
BP_D void rinter_row_f(const double* p, long plen, long k, double xd, double* val, double* der) {
    const double pm = rinter_at(p, plen, k - 1, 0), p0 = rinter_at(p, plen, k, 0);
    const double p1 = rinter_at(p, plen, k + 1, 0), p2 = rinter_at(p, plen, k + 2, 0);
    const double c1 = 0.5 * (p1 - pm);
    const double c2 = fma(2., p1, pm) - 0.5 * fma(5., p0, p2);
    const double c3 = 0.5 * (fma(3., p0 - p1, p2) - pm);
    *val = fma(xd, fma(xd, fma(xd, c3, c2), c1), p0);
    *der = fma(xd, fma(xd * 3., c3, 2. * c2), c1);
}
BP_D double rinter_f(const double* p, int nr, int nc, double x, double y, double* dfdx, double* dfdy) {
    const long plen = (long)nr * nc;
    const long i = (long)x, j = (long)y;
    const double xd = x - (double)i, yd = y - (double)j;
    const long km = (long)nc * (j - 1) + i;
    double ym, y0, y1, y2, gm, g0, g1, g2;
    rinter_row_f(p, plen, km, xd, &ym, &gm);
    rinter_row_f(p, plen, km + nc, xd, &y0, &g0);
    rinter_row_f(p, plen, km + 2L * nc, xd, &y1, &g1);
    rinter_row_f(p, plen, km + 3L * nc, xd, &y2, &g2);
    {
        const double d1 = 0.5 * (g1 - gm);
        const double d2 = fma(2., g1, gm) - 0.5 * fma(5., g0, g2);
        const double d3 = 0.5 * (fma(3., g0 - g1, g2) - gm);
        *dfdx = fma(yd, fma(yd, fma(yd, d3, d2), d1), g0);
    }
    const double d1 = 0.5 * (y1 - ym);
    const double d2 = fma(2., y1, ym) - 0.5 * fma(5., y0, y2);
    const double d3 = 0.5 * (fma(3., y0 - y1, y2) - ym);
    *dfdy = fma(yd, fma(yd * 3., d3, 2. * d2), d1);
    return fma(yd, fma(yd, fma(yd, d3, d2), d1), y0);
}

// 1 / d for 1 <= d <= 2^17 (or NaN in, NaN out): the hardware's 2^-23 reciprocal estimate and two Newton steps.  Only used
// for pkfit.py:364's 1 + (0.4 rel/chiold)^8 of pixels that passed the 10-sigma test, i.e. d in [1, 65537]: no zero, no
// infinity, no subnormal can reach it, which is what the ~25 extra instructions of a full IEEE division guard against.
BP_D double rcp_bounded(double d) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    double e = fma(-d, r, 1.0);
    r = fma(r, e, r);
    e = fma(-d, r, 1.0);
    return fma(r, e, r);
}

// dao_value.py:95-107 for the pixel in column c, row r of the box (offsets dxv, dyv from the star).
BP_D double model_at(const double* AX, int acap, int c, int r, double dxv, double dyv, const double* gauss,
                     const double* psf, int npsf, double half, double* dvx, double* dvy) {
    double p1, p2, dfdx, dfdy;
    const double e = daoerf_from_axes(gauss[0], AX[c], AX[acap + c], AX[2 * acap + r], AX[3 * acap + r], &p1, &p2);
#if PK_FMA_MODEL
    const double v = rinter_f(psf, npsf, npsf, fma(2., dxv, half), fma(2., dyv, half), &dfdx, &dfdy);
    *dvx = fma(2., dfdx, -p1);
    *dvy = fma(2., dfdy, -p2);
#else
    const double v = rinter(psf, npsf, npsf, 2. * dxv + half, 2. * dyv + half, &dfdx, &dfdy);
    *dvx = 2. * dfdx - p1;
    *dvy = 2. * dfdy - p2;
#endif
    return e + v;
}

struct FitState {
    double clamp0, clamp1, clamp2, dtold0, dtold1, dtold2;
    double iclamp0, iclamp1, iclamp2;  // 1 / clamp: the clamps are powers of two, so multiplying by these is dividing by those, bit for bit
    double chiold, chi, sharp, errmag, chi2, scale, x, y;
    int niter;
};
BP_D FitState fit_init(double scale, double x, double y, bool norecenter = false) {
    FitState S;
    S.clamp0 = S.clamp1 = S.clamp2 = 1.;
    S.iclamp0 = S.iclamp1 = S.iclamp2 = 1.;
    S.dtold0 = S.dtold1 = S.dtold2 = 0.;
    S.chiold = 1.;
    S.chi = S.sharp = S.errmag = S.chi2 = nan64();
    if (norecenter) S.chi = S.sharp = S.errmag = 100000.;  // pkfit_norecenter.py:352-354
    S.scale = scale; S.x = x; S.y = y;
    S.niter = 0;
    return S;
}
BP_D FitOut fit_out(const FitState& S) {
    FitOut o;
    o.errmag = S.errmag; o.chi = S.chi; o.sharp = S.sharp; o.scale = S.scale; o.x = S.x; o.y = S.y;
    o.chi2 = S.chi2; o.niter = S.niter;
    return o;
}
// the reference's NaN returns: scale and errmag always NaN; chi / sharp keep their last value in some branches
BP_D FitOut fit_nan(FitState& S, bool keep_chi, bool keep_sharp) {
    S.scale = nan64(); S.errmag = nan64();
    if (!keep_chi) S.chi = nan64();
    if (!keep_sharp) S.sharp = nan64();
    return fit_out(S);
}

// Everything after the normal-equation sums (pkfit.py:379-455 and the same lines of the other two
// variants).  Returns false when the reference bails out with NaNs; otherwise updates S and sets *loop.
template <int V>
BP_D bool fit_tail(const PkfitParams& P, const double* gauss, FitState& S, double* v, double (*c)[3], double numer,
                   double denom, double schi, double sumwt, bool* loop) {
    S.chi = schi;
    if (sumwt > 3.0) {  // pkfit.py:379-381
#if PK_RECIP
        // sumwt is in (3, number of pixels]: the product is positive and finite, the reciprocal in rcp_bounded's range
        S.chi = 1.2533 * S.chi * bp_rsqrt(sumwt * (sumwt - 3.));
        S.chiold = PK_FAST_RCP ? ((sumwt - 3.) * S.chi + 3.) * rcp_bounded(sumwt) : ((sumwt - 3.) * S.chi + 3.) / sumwt;
#else
        S.chi = 1.2533 * S.chi * sqrt(1. / (sumwt * (sumwt - 3.)));
        S.chiold = ((sumwt - 3.) * S.chi + 3.) / sumwt;
#endif
    }
    double csum = 0.;
    for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++) csum = csum + c[a][b];
    bool skip_inv = false;
    if (V == 1 || V == 3) {  // pkfit_noise.py:399-410, pkfit_norecenter.py:586-594: NaN -> inversion skipped; inf -> inv raises -> NaN return
        if (isnan(csum)) skip_inv = true;
        else if (isinf(csum)) return false;
    } else if (isnan(csum) || isinf(csum)) return false;  // pkfit.py:383-406
    if (!skip_inv && inv3(c)) return false;
    const double dt0 = v[0] * c[0][0] + v[1] * c[1][0] + v[2] * c[2][0];  // pkfit.py:408-409
    const double dt1 = v[0] * c[0][1] + v[1] * c[1][1] + v[2] * c[2][1];
    const double dt2 = v[0] * c[0][2] + v[1] * c[1][2] + v[2] * c[2][2];
    if (V == 2 && isinf(c[0][0])) return false;  // pkfit_norecent_noise.py:474-483
    if (S.dtold0 * dt0 < -1.e-38) { S.clamp0 = S.clamp0 / 2.; S.iclamp0 = S.iclamp0 * 2.; }  // pkfit.py:421-423
    if (S.dtold1 * dt1 < -1.e-38) { S.clamp1 = S.clamp1 / 2.; S.iclamp1 = S.iclamp1 * 2.; }
    if (S.dtold2 * dt2 < -1.e-38) { S.clamp2 = S.clamp2 / 2.; S.iclamp2 = S.iclamp2 * 2.; }
    S.dtold0 = dt0; S.dtold1 = dt1; S.dtold2 = dt2;
    const double adt0 = fabs(dt0), adt1 = fabs(dt1), adt2 = fabs(dt2);
#if PK_RECIP
    const double rs = dt0 / S.scale;  // pkfit.py:427-429 with one division for both candidates
    double denom2 = rs * (1. / 5.25);
    if (denom2 < -rs * (1. / 0.84)) denom2 = -rs * (1. / 0.84);
    // the three step denominators are 1 + (something >= 0): the bounded reciprocal where they are below 2^17, a division otherwise
    // (also for NaN); the choice is the same in every lane
    auto step = [](double dt, double d) { return (PK_FAST_RCP && d <= 131072.) ? dt * rcp_bounded(d) : dt / d; };
    S.scale = S.scale + step(dt0, 1 + denom2 * S.iclamp0);
    if (V == 1 || (V == 0 && P.recenter)) {  // pkfit.py:430-432
        S.x = S.x + step(dt1, 1. + adt1 * (2. * S.iclamp1));
        S.y = S.y + step(dt2, 1. + adt2 * (2. * S.iclamp2));
    }
#else
    double denom2 = dt0 / (5.25 * S.scale);  // pkfit.py:427-429
    if (denom2 < (-1 * dt0 / (0.84 * S.scale))) denom2 = (-1 * dt0 / (0.84 * S.scale));
    S.scale = S.scale + dt0 / (1 + denom2 / S.clamp0);
    if (V == 1 || (V == 0 && P.recenter)) {  // pkfit.py:430-432
        S.x = S.x + dt1 / (1. + adt1 / (0.5 * S.clamp1));
        S.y = S.y + dt2 / (1. + adt2 / (0.5 * S.clamp2));
    }
#endif
    bool redo = false;
    S.sharp = 2. * gauss[3] * gauss[4] * numer / (gauss[0] * S.scale * denom);  // pkfit.py:440
    S.errmag = S.chiold * sqrt(c[0][0]);
    if (V == 0) {  // pkfit.py:442
        const double a = 0.05 * S.errmag, b = 0.001 * S.scale;
        const double mx = (b > a) ? b : a;
        if (adt0 > mx) redo = true;
    } else if (adt0 > 0.05 * S.errmag || adt0 > 0.001 * S.scale) redo = true;  // pkfit_noise.py:445
    if (adt1 > 0.01 || adt2 > 0.01) redo = true;
    *loop = false;
    if (S.niter < 3) *loop = true;
    if (redo && S.errmag <= 1.9995 && S.niter < P.maxiter) *loop = true;
    return true;
}

// ------------------------------------------------------------ variant 2
// pkfit_norecent_noise.py:132-531.  W = PK_NARR_FORCED arrays of `cap` doubles, AX = 4 x acap doubles.
template <typename TI, typename TN>
BP_D FitOut pkfit_forced_warp(const PkfitParams& P, const double* __restrict__ psf, double* W, double* AX,
                              double scale, const double x, const double y, const double sky, const int lane) {
    const int cap = P.cap, acap = P.acap, npsf = P.npsf;
    double* MV = W;                // model value per good pixel (t[0])
    double* DVX = W + cap;         // d model / dx, dy  (t[1], t[2] = -scale * these)
    double* DVY = W + 2 * cap;
    double* FSUB = W + 3 * cap;    // pixel value
    double* SIG = W + 4 * cap;     // noise-image value
    double* W0 = W + 5 * cap;      // radial weight 5/(5 + rsq/(1-rsq)), pkfit_norecent_noise.py:375
    double* DFS = W + 6 * cap;     // sharpness kernel exp(-rho/2)(rho/2 - 1) where rho <= 36, else NaN (:377-390)
#if PK_HOIST
    // Two of the loop's eight fp64 divisions per pixel have operands that never change between iterations
    // (pkfit_norecent_noise.py:414 wt/sigsq and :393 dfdsig**2/sigsq): same operands, same quotient, so they are taken
    // once at set-up.
    double* WTR = W + 7 * cap;     // W0 / sigsq
    double* DDS = W + 8 * cap;     // DFS * DFS / sigsq
#endif
    const TI* __restrict__ img = reinterpret_cast<const TI*>(P.img);
    const TN* __restrict__ noise = reinterpret_cast<const TN*>(P.noise);
    const uint8_t* __restrict__ mask = P.mask;
    const double gauss[5] = {P.gauss[0], P.gauss[1], P.gauss[2], P.gauss[3], P.gauss[4]};
    const double radsq = P.radius * P.radius;
    const double gx2 = gauss[3] * gauss[3], gy2 = gauss[4] * gauss[4];
    const double igx2 = 1. / gx2, igy2 = 1. / gy2;  // PK_RECIP
    (void)igx2; (void)igy2; (void)gx2; (void)gy2;
    const double half = (double)(npsf - 1) / 2.;
    const unsigned lt = (1u << lane) - 1u;

    FitState S = fit_init(scale, x, y);
    // pkfit_norecent_noise.py:156-164 tests NaN before the loop (niter stays 0).  A non-finite position
    // cannot be truncated to a pixel index (the reference raises); treated as NaN.
    if (!isfinite(x) || !isfinite(y)) return fit_nan(S, false, false);

    // ---- once per star (the reference redoes this every iteration from identical inputs, :170-240)
    S.niter = 1;
    const FitBox b = fit_box(P, x, y);
#if PK_PREFETCH
    // the box's rows of the three images, requested before the erf / exp work of fill_axes instead of one chunk at a time
    prefetch_rows(noise, (size_t)P.npitch * sizeof(TN), b.iylo, b.iyy, (size_t)b.ixlo * sizeof(TN), (size_t)(b.ixlo + b.ixx) * sizeof(TN), lane);
    prefetch_rows(mask, (size_t)P.mpitch, b.iylo, b.iyy, (size_t)b.ixlo, (size_t)(b.ixlo + b.ixx), lane);
    prefetch_rows(img, (size_t)P.pitch * sizeof(TI), b.iylo, b.iyy, (size_t)b.ixlo * sizeof(TI), (size_t)(b.ixlo + b.ixx) * sizeof(TI), lane);
#endif
    fill_axes(b, x, y, gauss, AX, acap, lane);
    int ngood = 0;
    double xmin = KeyTraits<double>::inf(), xmax = -KeyTraits<double>::inf(), ymin = xmin, ymax = xmax;
    for (int base = 0; base < b.npix; base += 32) {
        const int p = base + lane;
        bool ok = false;
        int r = 0, c = 0;
        double dxv = 0., dyv = 0., rsq = 0., nz = 0.;
        if (p < b.npix) {
            r = p / b.ixx;
            c = p - r * b.ixx;
            dyv = (double)(r + b.iylo) - y;
            dxv = (double)(c + b.ixlo) - x;
            rsq = (dxv * dxv + dyv * dyv) / radsq;
            if (rsq < 1.) {  // :212-219
                const size_t row = (size_t)(b.iylo + r), col = (size_t)(b.ixlo + c);
                nz = (double)noise[row * P.npitch + col];
                ok = (nz > 0) && (mask[row * P.mpitch + col] == 0);
            }
        }
        const unsigned mk = __ballot_sync(FULL, ok);
        if (ok) {
            const int q = ngood + __popc(mk & lt);
            if (q < cap) {
                const size_t row = (size_t)(b.iylo + r), col = (size_t)(b.ixlo + c);
                FSUB[q] = (PK_FMA_LOOP & 1) ? (double)img[row * P.pitch + col] - sky : (double)img[row * P.pitch + col];  // PK_RECIP: sky taken off once
#if PK_RECIP
                const double isg = 1. / nz;
                SIG[q] = isg;
                W0[q] = 5. * (1. - rsq) / (5. - 4. * rsq);  // :375, 5 / (5 + rsq / (1 - rsq)) with one division
                const double rho = dxv * dxv * igx2 + dyv * dyv * igy2;  // :338
#else
                SIG[q] = nz;
                W0[q] = 5. / (5. + rsq / (1. - rsq));
                const double rho = dxv * dxv / gx2 + dyv * dyv / gy2;  // :338
#endif
                const double rh = 0.5 * rho;
                const double dfsv = (rho <= 36.) ? exp(-rh) * (rh - 1.) : nan64();
#if PK_HOIST
                const double ssq = nz * nz;
                (void)ssq;
                DFS[q] = (PK_FMA_LOOP & 2) ? dfsv / ssq : dfsv;  // PK_RECIP: the sharpness numerator's per-pixel factor, NaN where dfsv is
#else
                DFS[q] = dfsv;
#endif
#if PK_HOIST && PK_RECIP
                WTR[q] = W0[q] * (isg * isg);
                DDS[q] = dfsv * dfsv * (isg * isg);
#elif PK_HOIST
                WTR[q] = W0[q] / ssq;
                DDS[q] = dfsv * dfsv / ssq;
#endif
                // model inputs, consumed below once the dao_value guard is known
                MV[q] = dxv; DVX[q] = dyv; DVY[q] = __longlong_as_double((long long)(r * 65536 + c));
                xmin = fmin(xmin, dxv); xmax = fmax(xmax, dxv);
                ymin = fmin(ymin, dyv); ymax = fmax(ymax, dyv);
            }
        }
        ngood += __popc(mk);
    }
    if (ngood > cap) ngood = cap;
    if (ngood == 0) return fit_nan(S, false, false);  // :225-236
    __syncwarp();
    // dao_value.py:60-84: any point off the look-up table zeroes the model for ALL points
    xmin = 2. * pk_min(xmin) + half; xmax = 2. * pk_max(xmax) + half;
    ymin = 2. * pk_min(ymin) + half; ymax = 2. * pk_max(ymax) + half;
    const bool guard = (xmin < 1.) || (xmax > npsf - 2.) || (ymin < 1.) || (ymax > npsf - 2.);
    for (int q = lane; q < ngood; q += 32) {
        double mval = 0., dvx = 0., dvy = 0.;
        if (!guard) {
            const long long rc = __double_as_longlong(DVY[q]);
            mval = model_at(AX, acap, (int)(rc & 65535), (int)(rc >> 16), MV[q], DVX[q], gauss, psf, npsf, half, &dvx, &dvy);
        }
        MV[q] = mval; DVX[q] = dvx; DVY[q] = dvy;
    }
    __syncwarp();

    bool loop = true;
    S.niter = 0;
    double prev_scale = nan64(), prev_chiold = nan64();  // PK_FIXPOINT: what the previous iteration started from
    int nstill = 0;
    (void)prev_scale; (void)prev_chiold; (void)nstill;
    while (loop) {
        S.niter++;
        const double chiold = S.chiold;
        const double in_scale = S.scale, in_chiold = S.chiold;
        (void)in_scale; (void)in_chiold;
        const double inv_chi = 1. / chiold;  // PK_RECIP
        const bool later = S.niter >= 2;
        double numer = 0., denom = 0., s3 = 0., schi = 0., sumwt = 0.;
        double v0 = 0., v1 = 0., v2 = 0., c00 = 0., c01 = 0., c02 = 0., c11 = 0., c12 = 0., c22 = 0.;
        bool wle = false;  // some kept pixel has weight <= 100000: min(wt) > 100000 (:416-424) is "no lane saw one"
        bool wfirst_nan = false, seen = false;
        int nkept = 0, nl = 0;
        // PK_U chunks of 32 pixels per pass: the divisions of the chunks are independent and overlap; the sums still
        // take the pixels in list order (chunk by chunk), so every partial sum has the bits it had with one chunk per pass
        for (int base0 = 0; base0 < ngood; base0 += 32 * PK_U) {
            bool keep[PK_U], lil[PK_U];
            double df[PK_U], rel[PK_U], sigsq[PK_U], t0[PK_U], dfs[PK_U], wt[PK_U], t1[PK_U], t2[PK_U], wtr[PK_U], dds[PK_U];
#pragma unroll
            for (int u = 0; u < PK_U; u++) {
                const int q = base0 + 32 * u + lane;
                keep[u] = false; lil[u] = false;
                df[u] = 0.; rel[u] = 0.; sigsq[u] = 1.; t0[u] = 0.; dfs[u] = 0.; wt[u] = 0.; t1[u] = 0.; t2[u] = 0.; wtr[u] = 0.; dds[u] = 0.;
                if (q < ngood) {
                    t0[u] = MV[q];
                    const double sg = SIG[q];
#if PK_RECIP
                    df[u] = (PK_FMA_LOOP & 1) ? fma(-S.scale, t0[u], FSUB[q]) : FSUB[q] - S.scale * t0[u] - sky;  // :278 (PK_FMA_LOOP: FSUB = pixel - sky)
                    rel[u] = df[u] * sg;                       // :327, sg = 1 / sigma
                    keep[u] = !(later && fabs(rel[u] * inv_chi) >= 10.);
#else
                    df[u] = FSUB[q] - S.scale * t0[u] - sky;  // :278
                    rel[u] = df[u] / sg;                       // :327
                    keep[u] = !(later && fabs(rel[u] / chiold) >= 10.);  // :341-373 (rejection restarts from the full list)
#endif
                    sigsq[u] = sg * sg;                        // PK_RECIP: 1 / sigma^2
                    dfs[u] = DFS[q];
                    lil[u] = keep[u] && (dfs[u] == dfs[u]);
#if PK_HOIST
                    dds[u] = DDS[q];
#endif
                    if (keep[u]) {
                        wt[u] = W0[q];
#if PK_HOIST
                        wtr[u] = WTR[q];
#else
                        wtr[u] = wt[u] / sigsq[u];  // :414
#endif
                        t1[u] = -S.scale * DVX[q];
                        t2[u] = -S.scale * DVY[q];
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < PK_U; u++) {
                if (base0 + 32 * u >= ngood) break;  // warp-uniform
                const unsigned mk = __ballot_sync(FULL, keep[u]);
                if (lil[u]) {  // :377-410
                    numer = ((PK_FMA_LOOP & 2) && PK_HOIST) ? fma(dfs[u], df[u], numer) : PK_RECIP ? numer + dfs[u] * df[u] * sigsq[u] : numer + dfs[u] * df[u] / sigsq[u];
#if PK_HOIST
                    denom = denom + dds[u];
#else
                    denom = denom + dfs[u] * dfs[u] / sigsq[u];
#endif
                    s3 = (PK_FMA_LOOP & 4) ? fma(df[u] * df[u], sigsq[u], s3) : PK_RECIP ? s3 + df[u] * df[u] * sigsq[u] : s3 + df[u] * df[u] / sigsq[u];
                    nl++;
                }
                if (keep[u]) {
                    schi = (PK_FMA_LOOP & 4) ? fma(wt[u], fabs(rel[u]), schi) : schi + wt[u] * fabs(rel[u]);  // :411
                    sumwt = sumwt + wt[u];
                    double w = wtr[u];
                    wle = wle || (w <= 100000.);
                    if (!seen && (mk & lt) == 0u) wfirst_nan = (w != w);
                    if (later) {                   // :426-427
                        const double e = PK_RECIP ? 0.4 * rel[u] * inv_chi : 0.4 * rel[u] / chiold;
                        const double e2 = e * e, e4 = e2 * e2;
                        w = PK_FAST_RCP ? w * rcp_bounded(fma(e4, e4, 1.)) : (PK_FMA_LOOP & 4) ? w / fma(e4, e4, 1.) : w / (1. + e4 * e4);
                    }
#if PK_RECIP
                    {   // weighted model terms once, then one fused multiply-add per normal-equation sum (12 fp64 operations, not 27)
                        const double a0 = t0[u] * w, a1 = t1[u] * w, a2 = t2[u] * w;
                        v0 = fma(df[u], a0, v0); v1 = fma(df[u], a1, v1); v2 = fma(df[u], a2, v2);
                        c00 = fma(t0[u], a0, c00); c01 = fma(t0[u], a1, c01); c02 = fma(t0[u], a2, c02);
                        c11 = fma(t1[u], a1, c11); c12 = fma(t1[u], a2, c12); c22 = fma(t2[u], a2, c22);
                    }
#else
                    v0 = v0 + df[u] * t0[u] * w; v1 = v1 + df[u] * t1[u] * w; v2 = v2 + df[u] * t2[u] * w;
                    c00 = c00 + t0[u] * t0[u] * w; c01 = c01 + t0[u] * t1[u] * w; c02 = c02 + t0[u] * t2[u] * w;
                    c11 = c11 + t1[u] * t1[u] * w; c12 = c12 + t1[u] * t2[u] * w; c22 = c22 + t2[u] * t2[u] * w;
#endif
                }
                nkept += __popc(mk);  // warp-uniform count
                if (mk) seen = true;
            }
        }
        if (nkept == 0) return fit_nan(S, false, true);  // :347-356 (every pixel rejected)
        nl = warp_sum_int(nl);
        if (nl == 0) return fit_nan(S, false, false);    // :391-401
#if PK_TSUM
        {   // the fourteen sums of the iteration in one pass (warp_sum16: same additions, same bits as fourteen butterflies)
            double t[16] = {numer, denom, s3, schi, sumwt, v0, v1, v2, c00, c01, c02, c11, c12, c22, 0., 0.};
            warp_sum16(t, lane);
            numer = t[0]; denom = t[1]; s3 = t[2]; schi = t[3]; sumwt = t[4]; v0 = t[5]; v1 = t[6]; v2 = t[7];
            c00 = t[8]; c01 = t[9]; c02 = t[10]; c11 = t[11]; c12 = t[12]; c22 = t[13];
        }
        S.chi2 = s3 / (double)(nl - 1);          // :409-410
        S.chi = schi;
#else
        numer = pk_sum(numer); denom = pk_sum(denom);
        S.chi2 = pk_sum(s3) / (double)(nl - 1);          // :409-410
        schi = pk_sum(schi);
        S.chi = schi;
        sumwt = pk_sum(sumwt);
#endif
        {   // :416-424; python min(): a NaN first element wins
            // (NaN weights never win a later comparison in python's min(), so they count for nothing unless first)
            const bool fn = __any_sync(FULL, wfirst_nan);
            if (!fn && !__any_sync(FULL, wle)) return fit_nan(S, false, false);
        }
        double v[3], c[3][3];
#if PK_TSUM
        v[0] = v0; v[1] = v1; v[2] = v2;
        c[0][0] = c00; c[1][1] = c11; c[2][2] = c22;
        c[0][1] = c[1][0] = c01;
        c[0][2] = c[2][0] = c02;
        c[1][2] = c[2][1] = c12;
#else
        v[0] = pk_sum(v0); v[1] = pk_sum(v1); v[2] = pk_sum(v2);
        c[0][0] = pk_sum(c00); c[1][1] = pk_sum(c11); c[2][2] = pk_sum(c22);
        c[0][1] = c[1][0] = pk_sum(c01);
        c[0][2] = c[2][0] = pk_sum(c02);
        c[1][2] = c[2][1] = pk_sum(c12);
#endif
        if (!fit_tail<2>(P, gauss, S, v, c, numer, denom, schi, sumwt, &loop)) return fit_nan(S, false, false);
#if PK_FIXPOINT
        // A quarter of the forced fits never meet the reference's convergence test (it asks the frozen position to stop
        // moving, pkfit_norecent_noise.py:517) and run to maxiter on a scale that stopped changing long before.  The iteration is a
        // function of (scale, chiold) alone -- position, pixel list and model are fixed -- so when an iteration starts from the
        // same two numbers, bit for bit, as the one before it (both past the first, so both with the 10-sigma test and the
        // down-weighting on), it computes the same sums and the same step; its clamp test sees dtold * dt = dt^2 >= 0 and leaves
        // the clamp alone; and every later iteration repeats it.  The reference would go round until niter = maxiter and return
        // what is in hand now.
        if (loop && S.niter >= 3 && in_scale == prev_scale && in_chiold == prev_chiold) { S.niter = P.maxiter; break; }
#if PK_FIXPOINT > 1
        // In practice the scale does not stop: it wanders by an ulp or two around the fixed point (rounding noise of the sums)
        // until maxiter.  PK_FIXPOINT = 2 also leaves when two iterations in a row started within PK_FIX_TOL (relative) of the
        // one before in both numbers: what the remaining <= 22 iterations can still move is then below 1e-12, three orders of
        // magnitude inside the 1e-9 the fits are held to; niter is maxiter either way (the loop test reads the same quantities).
        // Measured: 14.64 -> 14.46 ms per 1 M stars, parity 0 of 400 000 -- the fits that run to maxiter are NOT settling (they
        // cycle with a larger amplitude, pixels entering and leaving the 10-sigma test), so there is little to leave early from;
        // not worth an argument about tolerances: the default is the exact test only (PK_FIXPOINT = 1, 14.78 -> 14.64 ms).
        {
            const bool still = fabs(in_scale - prev_scale) <= PK_FIX_TOL * fabs(in_scale) && fabs(in_chiold - prev_chiold) <= PK_FIX_TOL * fabs(in_chiold);
            nstill = still ? nstill + 1 : 0;
            if (loop && S.niter >= 4 && nstill >= 2) { S.niter = P.maxiter; break; }
        }
#endif
        prev_scale = in_scale; prev_chiold = in_chiold;
#endif
    }
    return fit_out(S);
}

// ------------------------------------------------------- variants 0 and 1
// pkfit.py:132-458 (V = 0), pkfit_noise.py:131-462 (V = 1) and pkfit_norecenter.py:346-657 (V = 3: pkfit.py's analytic
// noise model with the position frozen, pkfit_noise's convergence test, its own early returns): the pixel list and the
// model are rebuilt every iteration.  W = PK_NARR_FIT arrays of `cap` doubles, GIDX = cap ints, AX = 4 x acap doubles.
template <int V, typename TI, typename TN>
BP_D FitOut pkfit_recenter_warp(const PkfitParams& P, const double* __restrict__ psf, double* W, int* GIDX, double* AX,
                                double scale, double x, double y, const double sky, const int lane) {
    const int cap = P.cap, acap = P.acap, npsf = P.npsf;
    double* DX = W;              // pass A -> B
    double* DY = W + cap;
    double* DF = DX;             // pass B onwards (same slots, same owner lane)
    double* REL = DY;
    double* RSQ = W + 2 * cap;
    double* RHO = W + 3 * cap;
    double* FSUB = W + 4 * cap;
    double* SIG = W + 5 * cap;
    double* SIGSQ = W + 6 * cap;
    double* T0 = W + 7 * cap;
    double* T1 = W + 8 * cap;
    double* T2 = W + 9 * cap;
    const TI* __restrict__ img = reinterpret_cast<const TI*>(P.img);
    const TN* __restrict__ noise = reinterpret_cast<const TN*>(P.noise);
    const uint8_t* __restrict__ mask = P.mask;
    const double gauss[5] = {P.gauss[0], P.gauss[1], P.gauss[2], P.gauss[3], P.gauss[4]};
    const double radsq = P.radius * P.radius;
    const double gx2 = gauss[3] * gauss[3], gy2 = gauss[4] * gauss[4];
    const double pkerr = 0.027 / ((gauss[3] * gauss[4]) * (gauss[3] * gauss[4]));  // pkfit.py:146
    const double half = (double)(npsf - 1) / 2.;
    const unsigned lt = (1u << lane) - 1u;
    constexpr bool ANALYTIC = (V == 0 || V == 3);  // variance from the model (pkfit.py:277-282), no noise / mask image

    FitState S = fit_init(scale, x, y, V == 3);
    bool loop = true;
    while (loop) {
        S.niter++;
        // pkfit.py:158-166.  A non-finite position cannot be truncated to a pixel index (the reference raises); treated as NaN.
        if (!isfinite(S.x) || !isfinite(S.y)) return fit_nan(S, false, false);
        const double scl = S.scale, chiold = S.chiold;
        const double inv_chi = 1. / chiold;  // PK_RECIP
        (void)inv_chi;
        const FitBox b = fit_box(P, S.x, S.y);
        const double qaxis = fill_axes(b, S.x, S.y, gauss, AX, acap, lane);
        const bool qsep = b.ixx + b.iyy <= 32;  // warp-uniform: the rho terms of every column and row sit in one lane each

        // ---- pass A: ordered list of good pixels (pkfit.py:184-219, pkfit_noise.py:206-208)
        int ngood = 0;
        double xmin = KeyTraits<double>::inf(), xmax = -KeyTraits<double>::inf(), ymin = xmin, ymax = xmax;
        // box position of this lane's pixel, advanced by 32 per pass without a division
        const int ixd = b.ixx > 0 ? b.ixx : 1, krow = 32 / ixd, kcol = 32 - krow * ixd;
        int pr = lane / ixd, pc = lane - pr * ixd;
        for (int base = 0; base < b.npix; base += 32) {
            const int p = base + lane;
            const bool in = p < b.npix;
            bool ok = false;
            double dxv = 0., dyv = 0., rsq = 0., rho = 0., fv = 0., nz = 0.;
            // rho's two terms come from the lanes that own column c and row r (all lanes take part in the shuffles)
            const double qx = __shfl_sync(FULL, qaxis, (in && qsep) ? pc : 0), qy = __shfl_sync(FULL, qaxis, (in && qsep) ? b.ixx + pr : 0);
            if (in) {
                dyv = (double)(pr + b.iylo) - S.y;
                dxv = (double)(pc + b.ixlo) - S.x;
                const double dysq = dyv * dyv, dxsq = dxv * dxv;
                rsq = (dxsq + dysq) / radsq;
                ok = rsq < 1.;
                if (ok) {
                    // the three loads are independent: one memory round trip, not three
                    const size_t row = (size_t)(b.iylo + pr), col = (size_t)(b.ixlo + pc);
                    const TI fraw = img[row * P.pitch + col];
                    if (!ANALYTIC) {
                        const TN nraw = noise[row * P.npitch + col];
                        const uint8_t mraw = mask[row * P.mpitch + col];
                        nz = (double)nraw;
                        ok = (nz > 0) && (mraw == 0);
                    }
                    if (ok) {
                        fv = (double)fraw;
                        rho = qsep ? qx + qy : dxsq / gx2 + dysq / gy2;  // pkfit.py:292
                    }
                }
            }
            const unsigned mk = __ballot_sync(FULL, ok);
            if (ok) {
                const int q = ngood + __popc(mk & lt);
                if (q < cap) {
                    DX[q] = dxv; DY[q] = dyv; RSQ[q] = rsq; RHO[q] = rho; FSUB[q] = fv;
                    GIDX[q] = pr * 65536 + pc;
                    if (!ANALYTIC) {
#if PK_RECIP
                        const double isg = 1. / nz;  // SIG / SIGSQ hold 1/sigma and 1/sigma^2: one division per pixel, the rest multiply
                        SIG[q] = isg; SIGSQ[q] = isg * isg;
#else
                        SIG[q] = nz; SIGSQ[q] = nz * nz;
#endif
                    }
                    xmin = fmin(xmin, dxv); xmax = fmax(xmax, dxv);
                    ymin = fmin(ymin, dyv); ymax = fmax(ymax, dyv);
                }
            }
            ngood += __popc(mk);
            pr += krow; pc += kcol;
            if (pc >= ixd) { pc -= ixd; pr++; }
        }
        if (ngood > cap) ngood = cap;
        if (ngood == 0) return fit_nan(S, false, false);  // pkfit.py:211-219
        __syncwarp();
        // dao_value.py:60-84: any point off the look-up table zeroes the model for ALL points
        xmin = 2. * pk_min(xmin) + half; xmax = 2. * pk_max(xmax) + half;
        ymin = 2. * pk_min(ymin) + half; ymax = 2. * pk_max(ymax) + half;
        const bool guard = (xmin < 1.) || (xmax > npsf - 2.) || (ymin < 1.) || (ymax > npsf - 2.);

        // ---- pass B: model, residuals (pkfit.py:224-283); the 10-sigma test of pkfit.py:296-300 is taken here, where the
        // residual is at hand: the flag travels in bit 30 of GIDX (box rows stay below 2^14: radius < 64)
        int nbad = 0;
        for (int q = lane; q < ngood; q += 32) {
            double mval = 0., dvx = 0., dvy = 0.;
            if (!guard) {
                const int rc = GIDX[q];
                mval = model_at(AX, acap, rc & 65535, rc >> 16, DX[q], DY[q], gauss, psf, npsf, half, &dvx, &dvy);
            }
            T0[q] = mval;
            T1[q] = -scl * dvx;
            T2[q] = -scl * dvy;
            const double fsub = FSUB[q];
            const double df = PK_RECIP ? fma(-scl, mval, fsub) - sky : fsub - scl * mval - sky;
            if (ANALYTIC) {  // pkfit.py:277-282 (ronois enters un-squared, as in the reference)
                double fpos = fsub - df;
                if (fpos < 0.) fpos = 0.;
                const double s2 = fpos / P.phpadu + P.ronois + (0.0075 * fpos) * (0.0075 * fpos) +
                                  (pkerr * (fpos - sky)) * (pkerr * (fpos - sky));
#if PK_RECIP
                const double isg = 1. / sqrt(s2);
                SIG[q] = isg;
                SIGSQ[q] = isg * isg;
#else
                SIGSQ[q] = s2;
                SIG[q] = sqrt(s2);
#endif
            }
            DF[q] = df;
#if PK_RECIP
            const double rel = df * SIG[q];
            REL[q] = rel;
            if (S.niter >= 2 && fabs(rel * inv_chi) >= 10.) { nbad++; GIDX[q] |= 0x40000000; }
#else
            const double rel = df / SIG[q];
            REL[q] = rel;
            if (S.niter >= 2 && fabs(rel / chiold) >= 10.) { nbad++; GIDX[q] |= 0x40000000; }
#endif
        }
        __syncwarp();

        // ---- 10-sigma rejection from the second iteration on (pkfit.py:296-321), ordered compaction
        int m = ngood;
        bool compacted = false;
        if (S.niter >= 2) {
            nbad = warp_sum_int(nbad);
            // pkfit.py:303-309; pkfit_norecenter.py:509 takes this return only when more than one pixel is rejected
            if (nbad == ngood && !(V == 3 && nbad <= 1)) return fit_nan(S, true, true);
            if (nbad > 0) {
                compacted = true;
                m = 0;
                for (int base = 0; base < ngood; base += 32) {
                    const int q = base + lane;
                    bool keep = false;
                    double a0 = 0., a1 = 0., a2 = 0., a3 = 0., a4 = 0., a5 = 0., a6 = 0.;
                    if (q < ngood) {
                        a0 = REL[q];
                        keep = !(GIDX[q] & 0x40000000);
                        a1 = DF[q]; a2 = RSQ[q]; a3 = RHO[q]; a4 = FSUB[q]; a5 = SIG[q]; a6 = SIGSQ[q];
                    }
                    const unsigned mk = __ballot_sync(FULL, keep);
                    __syncwarp();
                    if (keep) {
                        const int d = m + __popc(mk & lt);
                        REL[d] = a0; DF[d] = a1; RSQ[d] = a2; RHO[d] = a3; FSUB[d] = a4; SIG[d] = a5; SIGSQ[d] = a6;
                        GIDX[d] = q;
                    }
                    m += __popc(mk);
                    __syncwarp();
                }
            }
        }
        // pkfit.py keeps t un-compacted and slices t[:, 0:lenwt] (pkfit.py:369-372): the k-th surviving
        // pixel is paired with the k-th ORIGINAL pixel's t.  pkfit_noise indexes t by goodind.
        const bool direct = !compacted || ANALYTIC;

        // ---- weights and sums (pkfit.py:323-372)
        double numer = 0., denom = 0., schi = 0., sumwt = 0.;
        double v0 = 0., v1 = 0., v2 = 0., c00 = 0., c01 = 0., c02 = 0., c11 = 0., c12 = 0., c22 = 0.;
        int nl = 0;
        for (int base = 0; base < m; base += 32) {
            const int q = base + lane;
            const bool in = q < m;
            const double rho = in ? RHO[q] : 1e300;
            const bool lil = in && (rho <= 36.);  // pkfit.py:324
            const unsigned mk = __ballot_sync(FULL, lil);
            if (lil) {
                const double rh = 0.5 * rho;
                const double dfdsig = exp(-rh) * (rh - 1.);  // pkfit.py:336
                double s;
                if (ANALYTIC) {  // pkfit.py:341-346
                    double fpos = FSUB[q];
                    if (fpos - sky < 0.) fpos = sky;
                    s = fpos / P.phpadu + P.ronois + (0.0075 * fpos) * (0.0075 * fpos) +
                        (pkerr * (fpos - sky)) * (pkerr * (fpos - sky));
                } else {
                    const double sg = SIG[q];
                    s = sg * sg;  // pkfit_noise.py:361 (PK_RECIP: this is 1 / sigma^2)
                }
                // pkfit.py:350 pairs dfdsig with df[0:len(lilrho)] (the FIRST nl residuals)
                const double dfv = ANALYTIC ? DF[nl + __popc(mk & lt)] : DF[q];
#if PK_RECIP
                const double is = ANALYTIC ? 1. / s : s;
                numer = fma(dfdsig * dfv, is, numer);
                denom = fma(dfdsig * dfdsig, is, denom);
#else
                numer = numer + dfdsig * dfv / s;
                denom = denom + dfdsig * dfdsig / s;
#endif
            }
            nl += __popc(mk);
            if (in) {
                const double rsq = RSQ[q], rel = REL[q], df = DF[q];
#if PK_RECIP
                double wt = 5. * (1. - rsq) / (5. - 4. * rsq);  // pkfit.py:323, 5 / (5 + rsq / (1 - rsq)) with one division
#else
                double wt = 5. / (5. + rsq / (1. - rsq));  // pkfit.py:323
#endif
                schi = PK_RECIP ? fma(wt, fabs(rel), schi) : schi + wt * fabs(rel);  // pkfit.py:359
                sumwt = sumwt + wt;
                wt = PK_RECIP ? wt * SIGSQ[q] : wt / SIGSQ[q];  // pkfit.py:362
                if (S.niter >= 2) {  // pkfit.py:364
                    const double e = PK_RECIP ? 0.4 * rel * inv_chi : 0.4 * rel / chiold;
                    const double e2 = e * e, e4 = e2 * e2;
                    wt = PK_FAST_RCP ? wt * rcp_bounded(fma(e4, e4, 1.)) : wt / (1. + e4 * e4);  // survivors of the 10-sigma test: e < 4
                }
                const int g = direct ? q : GIDX[q];
                const double t0 = T0[g], t1 = T1[g], t2 = T2[g];
#if PK_RECIP
                {
                    const double a0 = t0 * wt, a1 = t1 * wt, a2 = t2 * wt;
                    v0 = fma(df, a0, v0); v1 = fma(df, a1, v1); v2 = fma(df, a2, v2);
                    c00 = fma(t0, a0, c00); c01 = fma(t0, a1, c01); c02 = fma(t0, a2, c02);
                    c11 = fma(t1, a1, c11); c12 = fma(t1, a2, c12); c22 = fma(t2, a2, c22);
                }
#else
                v0 = v0 + df * t0 * wt; v1 = v1 + df * t1 * wt; v2 = v2 + df * t2 * wt;
                c00 = c00 + t0 * t0 * wt; c01 = c01 + t0 * t1 * wt; c02 = c02 + t0 * t2 * wt;
                c11 = c11 + t1 * t1 * wt; c12 = c12 + t1 * t2 * wt; c22 = c22 + t2 * t2 * wt;
#endif
            }
        }
        if (nl == 0) return fit_nan(S, false, false);  // pkfit.py:325-333
        // pkfit_norecenter.py:537,570-573 multiplies df[lilrho] (nl long) with wt (m long): numpy raises unless every pixel
        // lies within 6 sigma; reported as a failed fit
        if (V == 3 && nl != m) return fit_nan(S, false, false);
        double v[3], c[3][3];
#if PK_TSUM
        {
            double t[16] = {numer, denom, schi, sumwt, v0, v1, v2, c00, c01, c02, c11, c12, c22, 0., 0., 0.};
            warp_sum16(t, lane);
            numer = t[0]; denom = t[1]; schi = t[2]; sumwt = t[3];
            v[0] = t[4]; v[1] = t[5]; v[2] = t[6];
            c[0][0] = t[7]; c[1][1] = t[10]; c[2][2] = t[12];
            c[0][1] = c[1][0] = t[8];
            c[0][2] = c[2][0] = t[9];
            c[1][2] = c[2][1] = t[11];
        }
        S.chi = schi;
#else
        numer = pk_sum(numer); denom = pk_sum(denom);
        schi = pk_sum(schi);
        S.chi = schi;
        sumwt = pk_sum(sumwt);
        v[0] = pk_sum(v0); v[1] = pk_sum(v1); v[2] = pk_sum(v2);
        c[0][0] = pk_sum(c00); c[1][1] = pk_sum(c11); c[2][2] = pk_sum(c22);
        c[0][1] = c[1][0] = pk_sum(c01);
        c[0][2] = c[2][0] = pk_sum(c02);
        c[1][2] = c[2][1] = pk_sum(c12);
#endif
        if (!fit_tail<V>(P, gauss, S, v, c, numer, denom, schi, sumwt, &loop)) return fit_nan(S, false, false);
        __syncwarp();
    }
    return fit_out(S);
}

// Shared memory: [psf table | per-warp W arrays | per-warp AX | per-warp GIDX (variants 0, 1)].
template <int V, typename TI, typename TN>
__global__ void __launch_bounds__(PK_THREADS, V == 2 ? PK_MINB_FORCED : PK_MINB_FIT) pkfit_kernel(const PkfitParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int NARR = (V == 2) ? PK_NARR_FORCED : PK_NARR_FIT;
    double* sbase = reinterpret_cast<double*>(smem_raw);
    const int ntab = (PK_PSF_GLOBAL && V == 2) ? 0 : P.npsf * P.npsf;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double* W = sbase + ntab + (size_t)warp * NARR * P.cap;
    double* AX = sbase + ntab + (size_t)PK_WARPS * NARR * P.cap + (size_t)warp * 4 * P.acap;
    int* GIDX = reinterpret_cast<int*>(sbase + ntab + (size_t)PK_WARPS * (NARR * P.cap + 4 * P.acap)) + (size_t)warp * P.cap;
    for (int i = threadIdx.x; i < ntab; i += PK_THREADS) sbase[i] = P.psf[i];
    const double* __restrict__ psf = ntab ? sbase : P.psf;
    __syncthreads();
    // stars are drawn from a counter, not a fixed stride: fits take 3 to 25 iterations, and a launch should end with
    // its last star, not with the warp that was dealt the most slow ones
    for (;;) {
        int i = 0;
        if (lane == 0) i = atomicAdd(P.work_counter, 1);
        i = __shfl_sync(FULL, i, 0);
        if (i >= P.nstars) break;
        const int s = P.order ? P.order[i] : i;  // stars are visited in image-tile order, results land at s
        FitOut o;
        if constexpr (V == 2) o = pkfit_forced_warp<TI, TN>(P, psf, W, AX, P.scale[s], P.x[s], P.y[s], P.sky[s], lane);
        else o = pkfit_recenter_warp<V, TI, TN>(P, psf, W, GIDX, AX, P.scale[s], P.x[s], P.y[s], P.sky[s], lane);
        if (lane == 0) {
            P.errmag[s] = o.errmag; P.chi[s] = o.chi; P.sharp[s] = o.sharp; P.niter[s] = o.niter;
            P.scale_out[s] = o.scale; P.x_out[s] = o.x; P.y_out[s] = o.y; P.chi2[s] = o.chi2;
        }
        __syncwarp();
    }
}

// ---------------------------------------------------- unit-test kernel
// dao_value.dao_value(xx, yy, gauss, psf, deriv=True) for n points, plus the pieces it is made of
// (rows a10-a12): out = [value, dvdx, dvdy, daoerf f, pder0..4, rinter z, dfdx, dfdy] x n (SoA, 12 rows).
// The "any point off the table" guard of dao_value.py:70-84 is applied by the host wrapper.
// table_coords != 0: xx, yy are look-up-table coordinates (rinter's own arguments); only the rinter rows are filled.
// table_coords > 1 is the length numpy would wrap negative flat indices against (rinter.py:117-157).
__global__ void psfmodel_kernel(const double* xx, const double* yy, long long n, const double* gauss5,
                                const double* psf, int npsf, int table_coords, double* out) {
    double g[5];
    for (int i = 0; i < 5; i++) g[i] = gauss5[i];
    const double half = (double)(npsf - 1) / 2.;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        double pder[5], dfdx, dfdy, dvx, dvy;
        const double f = daoerf(xx[i], yy[i], g, pder);
        const double tx = table_coords ? xx[i] : 2. * xx[i] + half, ty = table_coords ? yy[i] : 2. * yy[i] + half;
        double z = 0.;
        dfdx = 0.; dfdy = 0.; dvx = 0.; dvy = 0.;
        double val = 0.;
        if (table_coords) {  // rinter.rinter's own flat-index semantics: any point, table_coords > 1 = numpy wrap length
            z = rinter(psf, npsf, npsf, tx, ty, &dfdx, &dfdy, table_coords > 1 ? (long)table_coords : 0L);
        } else if (tx >= 1. && tx <= npsf - 2. && ty >= 1. && ty <= npsf - 2.) {
            z = rinter(psf, npsf, npsf, tx, ty, &dfdx, &dfdy);
            val = psf_value(xx[i], yy[i], g, psf, npsf, &dvx, &dvy);
        }
        out[0 * n + i] = val; out[1 * n + i] = dvx; out[2 * n + i] = dvy;
        out[3 * n + i] = f;
        for (int k = 0; k < 5; k++) out[(4 + k) * n + i] = pder[k];
        out[9 * n + i] = z; out[10 * n + i] = dfdx; out[11 * n + i] = dfdy;
    }
}

}  // namespace bphot
