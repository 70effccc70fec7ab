// pkfit_kernel.cu -- DAOPHOT PSF peak fit, one warp per star (sm_100a).
//
// Replaces the per-star Python loops of
//   variant 0  PythonPhot/pkfit.py:132-458                (analytic noise model, optional recentring)
//   variant 1  PythonPhot/pkfit_noise.py:131-462          (noise + mask images, recentring)
//   variant 2  PythonPhot/pkfit_norecent_noise.py:132-531 (noise + mask images, forced position)
// and their callee dao_value.dao_value (dao_value.py:9-121 -> daoerf.py, rinter.py).
//
// Per star, one warp runs the damped Gauss-Newton loop.  Pixels of the fitting box are
// enumerated in row-major order (ballot-ranked, so the pixel lists are ordered exactly like the
// reference's np.where output); each lane evaluates the PSF model (integrated Gaussian +
// bicubic residual look-up, table staged in shared memory) for its pixels; the 3-vector and
// 3x3 normal matrix are fp64 butterfly reductions (every lane ends with identical bits, so
// the convergence logic is warp-uniform); the 3x3 solve runs redundantly in registers.
#include "bphot_device.cuh"
#include "bphot_params.h"

namespace bphot {

constexpr int PK_WARPS = 4;
constexpr int PK_THREADS = PK_WARPS * 32;
constexpr int PK_NARR = 10;  // per-warp scratch arrays of `cap` doubles

struct FitOut {
    double errmag, chi, sharp, scale, x, y, chi2;
    int niter;
};

template <typename TI, typename TN>
BP_D FitOut pkfit_warp(const PkfitParams& P, const double* __restrict__ psf, double* W, int* GIDX,
                       double scale, double x, double y, const double sky, const int lane) {
    const int cap = P.cap, variant = P.variant, npsf = P.npsf;
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
    const double radius = P.radius, radsq = radius * radius;
    const double gx2 = gauss[3] * gauss[3], gy2 = gauss[4] * gauss[4];
    const double pkerr = 0.027 / ((gauss[3] * gauss[4]) * (gauss[3] * gauss[4]));  // pkfit.py:146
    const double half = (double)(npsf - 1) / 2.;
    const unsigned lt = (1u << lane) - 1u;

    double clamp0 = 1., clamp1 = 1., clamp2 = 1., dtold0 = 0., dtold1 = 0., dtold2 = 0.;
    double chiold = 1., chi = nan64(), sharp = nan64(), errmag = nan64(), chi2 = nan64();
    int niter = 0;

#define PK_RETURN()                                                                              \
    do {                                                                                         \
        FitOut o_;                                                                               \
        o_.errmag = errmag; o_.chi = chi; o_.sharp = sharp; o_.scale = scale; o_.x = x; o_.y = y; \
        o_.chi2 = chi2; o_.niter = niter;                                                        \
        return o_;                                                                               \
    } while (0)
#define PK_RETURN_NAN(keep_chi, keep_sharp)      \
    do {                                         \
        scale = nan64(); errmag = nan64();       \
        if (!(keep_chi)) chi = nan64();          \
        if (!(keep_sharp)) sharp = nan64();      \
        PK_RETURN();                             \
    } while (0)

    // pkfit_norecent_noise.py:156-164 tests NaN before the loop (niter stays 0); the others inside it.
    // A non-finite position cannot be truncated to a pixel index (the reference raises); treated as NaN.
    const bool badpos = !isfinite(x) || !isfinite(y);
    if (variant == 2 && badpos) PK_RETURN_NAN(0, 0);

    bool loop = true;
    while (loop) {
        niter++;
        if (badpos) PK_RETURN_NAN(0, 0);  // pkfit.py:158-166
        // pkfit.py:168-177: fitting box, int truncation then clamp
        long long ixlo = (long long)(x - radius); if (ixlo < 0) ixlo = 0;
        long long iylo = (long long)(y - radius); if (iylo < 0) iylo = 0;
        long long ixhi = (long long)(x + radius) + 1; if (ixhi > P.ncol - 1) ixhi = P.ncol - 1;
        long long iyhi = (long long)(y + radius) + 1; if (iyhi > P.nrow - 1) iyhi = P.nrow - 1;
        const long long ixx = ixhi - ixlo + 1, iyy = iyhi - iylo + 1;
        const int npix = (ixx > 0 && iyy > 0 && ixx * iyy < (1LL << 30)) ? (int)(ixx * iyy) : 0;

        // ---- pass A: ordered list of good pixels (pkfit.py:184-219, pkfit_noise.py:206-208)
        int ngood = 0;
        double xmin = KeyTraits<double>::inf(), xmax = -KeyTraits<double>::inf(), ymin = xmin, ymax = xmax;
        for (int base = 0; base < npix; base += 32) {
            const int p = base + lane;
            bool ok = false;
            double dxv = 0., dyv = 0., rsq = 0., rho = 0., fv = 0., nz = 0.;
            if (p < npix) {
                const int r = p / (int)ixx, c = p - r * (int)ixx;
                dyv = (double)(r + iylo) - y;
                dxv = (double)(c + ixlo) - x;
                const double dysq = dyv * dyv, dxsq = dxv * dxv;
                rsq = (dxsq + dysq) / radsq;
                ok = rsq < 1.;
                if (ok) {
                    const size_t row = (size_t)(iylo + r), col = (size_t)(ixlo + c);
                    if (variant != 0) {
                        nz = (double)noise[row * P.npitch + col];
                        ok = (nz > 0) && (mask[row * P.mpitch + col] == 0);
                    }
                    if (ok) {
                        fv = (double)img[row * P.pitch + col];
                        rho = dxsq / gx2 + dysq / gy2;  // pkfit.py:292
                    }
                }
            }
            const unsigned mk = __ballot_sync(FULL, ok);
            if (ok) {
                const int q = ngood + __popc(mk & lt);
                if (q < cap) {
                    DX[q] = dxv; DY[q] = dyv; RSQ[q] = rsq; RHO[q] = rho; FSUB[q] = fv;
                    if (variant != 0) { SIG[q] = nz; SIGSQ[q] = nz * nz; }
                    xmin = fmin(xmin, dxv); xmax = fmax(xmax, dxv);
                    ymin = fmin(ymin, dyv); ymax = fmax(ymax, dyv);
                }
            }
            ngood += __popc(mk);
        }
        if (ngood > cap) ngood = cap;
        if (ngood == 0) PK_RETURN_NAN(0, 0);  // pkfit.py:211-219
        __syncwarp();
        // dao_value.py:60-84: any point off the look-up table zeroes the model for ALL points
        xmin = 2. * warp_min(xmin) + half; xmax = 2. * warp_max(xmax) + half;
        ymin = 2. * warp_min(ymin) + half; ymax = 2. * warp_max(ymax) + half;
        const bool guard = (xmin < 1.) || (xmax > npsf - 2.) || (ymin < 1.) || (ymax > npsf - 2.);

        // ---- pass B: model, residuals (pkfit.py:224-283)
        for (int q = lane; q < ngood; q += 32) {
            double mval = 0., dvx = 0., dvy = 0.;
            if (!guard) mval = psf_value(DX[q], DY[q], gauss, psf, npsf, &dvx, &dvy);
            T0[q] = mval;
            T1[q] = -scale * dvx;
            T2[q] = -scale * dvy;
            const double fsub = FSUB[q];
            const double df = fsub - scale * mval - sky;
            if (variant == 0) {  // pkfit.py:277-282 (ronois enters un-squared, as in the reference)
                double fpos = fsub - df;
                if (fpos < 0.) fpos = 0.;
                const double s2 = fpos / P.phpadu + P.ronois + (0.0075 * fpos) * (0.0075 * fpos) +
                                  (pkerr * (fpos - sky)) * (pkerr * (fpos - sky));
                SIGSQ[q] = s2;
                SIG[q] = sqrt(s2);
            }
            DF[q] = df;
            REL[q] = df / SIG[q];
        }
        __syncwarp();

        // ---- 10-sigma rejection from the second iteration on (pkfit.py:296-321), ordered compaction
        int m = ngood;
        bool compacted = false;
        if (niter >= 2) {
            int nbad = 0;
            for (int q = lane; q < ngood; q += 32)
                if (fabs(REL[q] / chiold) >= 10.) nbad++;
            nbad = warp_sum_int(nbad);
            if (nbad == ngood) {
                if (variant == 2) PK_RETURN_NAN(0, 1);  // pkfit_norecent_noise.py:347-356
                PK_RETURN_NAN(1, 1);                    // pkfit.py:303-309
            }
            if (nbad > 0) {
                compacted = true;
                m = 0;
                for (int base = 0; base < ngood; base += 32) {
                    const int q = base + lane;
                    bool keep = false;
                    double a0 = 0., a1 = 0., a2 = 0., a3 = 0., a4 = 0., a5 = 0., a6 = 0.;
                    if (q < ngood) {
                        a0 = REL[q];
                        keep = !(fabs(a0 / chiold) >= 10.);
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
        // pixel is paired with the k-th ORIGINAL pixel's t.  The noise variants index t by goodind.
        const bool direct = !compacted || variant == 0;

        // ---- weights and sums (pkfit.py:323-372)
        double numer = 0., denom = 0., s3 = 0., schi = 0., sumwt = 0.;
        double v0 = 0., v1 = 0., v2 = 0., c00 = 0., c01 = 0., c02 = 0., c11 = 0., c12 = 0., c22 = 0.;
        double wmin = KeyTraits<double>::inf();
        bool wfirst_nan = false;
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
                if (variant == 0) {  // pkfit.py:341-346
                    double fpos = FSUB[q];
                    if (fpos - sky < 0.) fpos = sky;
                    s = fpos / P.phpadu + P.ronois + (0.0075 * fpos) * (0.0075 * fpos) +
                        (pkerr * (fpos - sky)) * (pkerr * (fpos - sky));
                } else {
                    const double sg = SIG[q];
                    s = sg * sg;  // pkfit_noise.py:361
                }
                // pkfit.py:350 pairs dfdsig with df[0:len(lilrho)] (the FIRST nl residuals)
                const double dfv = (variant == 0) ? DF[nl + __popc(mk & lt)] : DF[q];
                numer = numer + dfdsig * dfv / s;
                denom = denom + dfdsig * dfdsig / s;
                s3 = s3 + dfv * dfv / s;
            }
            nl += __popc(mk);
            if (in) {
                const double rsq = RSQ[q], rel = REL[q], df = DF[q];
                double wt = 5. / (5. + rsq / (1. - rsq));  // pkfit.py:323
                schi = schi + wt * fabs(rel);               // pkfit.py:359
                sumwt = sumwt + wt;
                wt = wt / SIGSQ[q];                          // pkfit.py:362
                if (q == 0) wfirst_nan = (wt != wt);
                wmin = fmin(wmin, wt);
                if (niter >= 2) {  // pkfit.py:364
                    const double e = 0.4 * rel / chiold;
                    const double e2 = e * e, e4 = e2 * e2;
                    wt = wt / (1. + e4 * e4);
                }
                const int g = direct ? q : GIDX[q];
                const double t0 = T0[g], t1 = T1[g], t2 = T2[g];
                v0 = v0 + df * t0 * wt; v1 = v1 + df * t1 * wt; v2 = v2 + df * t2 * wt;
                c00 = c00 + t0 * t0 * wt; c01 = c01 + t0 * t1 * wt; c02 = c02 + t0 * t2 * wt;
                c11 = c11 + t1 * t1 * wt; c12 = c12 + t1 * t2 * wt; c22 = c22 + t2 * t2 * wt;
            }
        }
        if (nl == 0) PK_RETURN_NAN(0, 0);  // pkfit.py:325-333
        numer = warp_sum(numer); denom = warp_sum(denom);
        if (variant == 2) chi2 = warp_sum(s3) / (double)(nl - 1);  // pkfit_norecent_noise.py:409-410
        chi = warp_sum(schi);
        sumwt = warp_sum(sumwt);
        if (variant == 2) {  // pkfit_norecent_noise.py:416-424; python min(): a NaN first element wins
            wmin = warp_min(wmin);
            const bool fn = __any_sync(FULL, wfirst_nan);
            if (!fn && wmin > 100000) PK_RETURN_NAN(0, 0);
        }
        double v[3], c[3][3];
        v[0] = warp_sum(v0); v[1] = warp_sum(v1); v[2] = warp_sum(v2);
        c[0][0] = warp_sum(c00); c[1][1] = warp_sum(c11); c[2][2] = warp_sum(c22);
        c[0][1] = c[1][0] = warp_sum(c01);
        c[0][2] = c[2][0] = warp_sum(c02);
        c[1][2] = c[2][1] = warp_sum(c12);
        if (sumwt > 3.0) {  // pkfit.py:379-381
            chi = 1.2533 * chi * sqrt(1. / (sumwt * (sumwt - 3.)));
            chiold = ((sumwt - 3.) * chi + 3.) / sumwt;
        }
        double csum = 0.;
        for (int a = 0; a < 3; a++)
            for (int b = 0; b < 3; b++) csum = csum + c[a][b];
        bool skip_inv = false;
        if (variant == 1) {  // pkfit_noise.py:399-410: NaN -> inversion skipped; inf -> inv raises -> NaN return
            if (isnan(csum)) skip_inv = true;
            else if (isinf(csum)) PK_RETURN_NAN(0, 0);
        } else if (isnan(csum) || isinf(csum)) PK_RETURN_NAN(0, 0);  // pkfit.py:383-406
        if (!skip_inv && inv3(c)) PK_RETURN_NAN(0, 0);
        double dt0 = v[0] * c[0][0] + v[1] * c[1][0] + v[2] * c[2][0];  // pkfit.py:408-409
        double dt1 = v[0] * c[0][1] + v[1] * c[1][1] + v[2] * c[2][1];
        double dt2 = v[0] * c[0][2] + v[1] * c[1][2] + v[2] * c[2][2];
        if (variant == 2 && isinf(c[0][0])) PK_RETURN_NAN(0, 0);  // pkfit_norecent_noise.py:474-483
        if (dtold0 * dt0 < -1.e-38) clamp0 = clamp0 / 2.;  // pkfit.py:421-423
        if (dtold1 * dt1 < -1.e-38) clamp1 = clamp1 / 2.;
        if (dtold2 * dt2 < -1.e-38) clamp2 = clamp2 / 2.;
        dtold0 = dt0; dtold1 = dt1; dtold2 = dt2;
        const double adt0 = fabs(dt0), adt1 = fabs(dt1), adt2 = fabs(dt2);
        double denom2 = dt0 / (5.25 * scale);  // pkfit.py:427-429
        if (denom2 < (-1 * dt0 / (0.84 * scale))) denom2 = (-1 * dt0 / (0.84 * scale));
        scale = scale + dt0 / (1 + denom2 / clamp0);
        if (variant == 1 || (variant == 0 && P.recenter)) {  // pkfit.py:430-432
            x = x + dt1 / (1. + adt1 / (0.5 * clamp1));
            y = y + dt2 / (1. + adt2 / (0.5 * clamp2));
        }
        bool redo = false;
        sharp = 2. * gauss[3] * gauss[4] * numer / (gauss[0] * scale * denom);  // pkfit.py:440
        errmag = chiold * sqrt(c[0][0]);
        if (variant == 0) {  // pkfit.py:442
            const double a = 0.05 * errmag, b = 0.001 * scale;
            const double mx = (b > a) ? b : a;
            if (adt0 > mx) redo = true;
        } else if (adt0 > 0.05 * errmag || adt0 > 0.001 * scale) redo = true;  // pkfit_noise.py:445
        if (adt1 > 0.01 || adt2 > 0.01) redo = true;
        if (niter >= 3) loop = false;
        if (redo && errmag <= 1.9995 && niter < P.maxiter) loop = true;
        __syncwarp();
    }
    PK_RETURN();
#undef PK_RETURN
#undef PK_RETURN_NAN
}

template <typename TI, typename TN>
__global__ void __launch_bounds__(PK_THREADS) pkfit_kernel(const PkfitParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* psf = reinterpret_cast<double*>(smem_raw);
    const int ntab = P.npsf * P.npsf;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double* W = psf + ntab + (size_t)warp * PK_NARR * P.cap;
    int* GIDX = reinterpret_cast<int*>(psf + ntab + (size_t)PK_WARPS * PK_NARR * P.cap) + (size_t)warp * P.cap;
    for (int i = threadIdx.x; i < ntab; i += PK_THREADS) psf[i] = P.psf[i];
    __syncthreads();
    const int wstride = gridDim.x * PK_WARPS;
    for (int s = blockIdx.x * PK_WARPS + warp; s < P.nstars; s += wstride) {
        FitOut o = pkfit_warp<TI, TN>(P, psf, W, GIDX, P.scale[s], P.x[s], P.y[s], P.sky[s], lane);
        if (lane == 0) {
            P.errmag[s] = o.errmag; P.chi[s] = o.chi; P.sharp[s] = o.sharp; P.niter[s] = o.niter;
            P.scale_out[s] = o.scale; P.x_out[s] = o.x; P.y_out[s] = o.y; P.chi2[s] = o.chi2;
        }
        __syncwarp();
    }
}

template __global__ void pkfit_kernel<float, float>(const PkfitParams);
template __global__ void pkfit_kernel<float, double>(const PkfitParams);
template __global__ void pkfit_kernel<double, float>(const PkfitParams);
template __global__ void pkfit_kernel<double, double>(const PkfitParams);

// ---------------------------------------------------- unit-test kernel
// dao_value.dao_value(xx, yy, gauss, psf, deriv=True) for n points, plus the pieces it is made of
// (rows a10-a12): out = [value, dvdx, dvdy, daoerf f, pder0..4, rinter z, dfdx, dfdy] x n (SoA, 12 rows).
// The "any point off the table" guard of dao_value.py:70-84 is applied by the host wrapper.
// table_coords != 0: xx, yy are look-up-table coordinates (rinter's own arguments); only the rinter rows are filled.
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
        if (tx >= 1. && tx <= npsf - 2. && ty >= 1. && ty <= npsf - 2.) {
            z = rinter(psf, npsf, npsf, tx, ty, &dfdx, &dfdy);
            if (!table_coords) val = psf_value(xx[i], yy[i], g, psf, npsf, &dvx, &dvy);
        }
        out[0 * n + i] = val; out[1 * n + i] = dvx; out[2 * n + i] = dvy;
        out[3 * n + i] = f;
        for (int k = 0; k < 5; k++) out[(4 + k) * n + i] = pder[k];
        out[9 * n + i] = z; out[10 * n + i] = dfdx; out[11 * n + i] = dfdy;
    }
}

}  // namespace bphot
