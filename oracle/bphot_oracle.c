/*
 * bphot_oracle.c -- CPU restatement of the PythonPhot hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity checker for the CUDA
 * kernels in pythonphot_b200/csrc and the "port" CPU baseline timed by
 * bench.py.  The product package never links, imports or calls it.
 *
 * It restates, in plain scalar C and in the reference's own order of
 * operations (sort + index walk for mmm, row-major pixel lists for pkfit),
 * the algorithms of djones1040/PythonPhot @ 7992eb6:
 *
 *   bo_pixwt      PythonPhot/pixwt.py:8-235     (vector path, as aper calls it)
 *   bo_mmm        PythonPhot/mmm.py:106-314
 *   bo_aper       PythonPhot/aper.py:134-433
 *   bo_daoerf     PythonPhot/daoerf.py:49-90
 *   bo_rinter     PythonPhot/rinter.py:98-195
 *   bo_dao_value  PythonPhot/dao_value.py:60-107
 *   bo_pkfit      PythonPhot/pkfit.py:132-458, pkfit_noise.py:131-462,
 *                 pkfit_norecent_noise.py:132-531
 *
 * Parity status: PINNED against the unmodified reference executed in the
 * authoring container (oracle/refloader.py) on seeded synthetic inputs; the
 * vectors are committed under tests/golden/ (generator: tests/golden/gen_golden.py).
 * The reference has no tests or golden vectors of its own apart from the
 * pixwt docstring value 0.6502 (pixwt.py:25-30), which tests/ also checks.
 *
 * Numeric contract: the reference is called with fp64 image / noise arrays
 * (SURVEY.md F8).  Image and noise may be passed here as fp32 or fp64; fp32
 * is widened on load, which is bit-equivalent to image.astype(float64).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define BO_F32 0
#define BO_F64 1

static inline double ld(const void *p, int dt, size_t i) {
    return dt == BO_F32 ? (double)((const float *)p)[i] : ((const double *)p)[i];
}

/* numpy's pairwise float64 add-reduce (numpy/_core/src/umath/loops_utils.h.src,
 * PW_BLOCKSIZE 128, 8 accumulators).  Used wherever the reference calls np.sum
 * so that the oracle reproduces the reference's rounding, not merely 1e-9. */
static double pw_sum(const double *a, long n) {
    if (n < 8) {
        double r = 0.0;
        for (long i = 0; i < n; i++) r += a[i];
        return r;
    } else if (n <= 128) {
        double r[8];
        long i;
        for (i = 0; i < 8; i++) r[i] = a[i];
        for (i = 8; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; j++) r[j] += a[i + j];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; i++) res += a[i];
        return res;
    } else {
        long n2 = n / 2;
        n2 -= n2 % 8;
        return pw_sum(a, n2) + pw_sum(a + n2, n - n2);
    }
}

/* ------------------------------------------------------------------ pixwt */
/* pixwt.py:50-62 */
static double bo_arc(double x, double y0, double y1, double r) {
    return 0.5 * r * r * (atan(y1 / x) - atan(y0 / x));
}
/* pixwt.py:64-72 */
static double bo_chord(double x, double y0, double y1) { return 0.5 * x * (y1 - y0); }

/* pixwt.py:112-213 (array branch, evaluated per element; IEEE division by a
 * zero x yields +-inf and atan(+-inf) = +-pi/2, exactly as numpy does) */
static double bo_oneside(double x, double y0, double y1, double r) {
    if (fabs(x) >= r) return bo_arc(x, y0, y1, r);
    double yh = sqrt(r * r - x * x);
    if (y0 <= -yh) {
        if (y1 <= -yh) return bo_arc(x, y0, y1, r);
        else if (y1 <= yh) return bo_arc(x, y0, -yh, r) + bo_chord(x, -yh, y1);
        else return bo_arc(x, y0, -yh, r) + bo_chord(x, -yh, yh) + bo_arc(x, yh, y1, r);
    } else if (y0 < yh) {
        if (y1 <= -yh) return bo_chord(x, y0, -yh) + bo_arc(x, -yh, y1, r);
        else if (y1 <= yh) return bo_chord(x, y0, y1);
        else return bo_chord(x, y0, yh) + bo_arc(x, yh, y1, r);
    } else {
        if (y1 <= -yh) return bo_arc(x, y0, yh, r) + bo_chord(x, yh, -yh) + bo_arc(x, -yh, y1, r);
        else if (y1 <= yh) return bo_arc(x, y0, yh, r) + bo_chord(x, yh, y1);
        else return bo_arc(x, y0, y1, r);
    }
}

/* pixwt.py:8-48 + Intarea pixwt.py:215-235 */
double bo_pixwt(double xc, double yc, double r, double x, double y) {
    double x0 = (x - 0.5) - xc, y0 = (y - 0.5) - yc;
    double x1 = (x + 0.5) - xc, y1 = (y + 0.5) - yc;
    return bo_oneside(x1, y0, y1, r) + bo_oneside(y1, -x1, -x0, r) +
           bo_oneside(-x0, -y1, -y0, r) + bo_oneside(-y0, x0, x1, r);
}

void bo_pixwt_vec(double xc, double yc, double r, const double *x, const double *y, int n, double *out) {
    for (int i = 0; i < n; i++) out[i] = bo_pixwt(xc, yc, r, x[i], y[i]);
}

/* -------------------------------------------------------------------- mmm */
static int cmp_dbl(const void *a, const void *b) {
    double x = *(const double *)a, y = *(const double *)b;
    return (x > y) - (x < y);
}

/* mmm.py:7-314.  `work` must hold n doubles.  Returns the number of
 * rejection-loop iterations performed; info[0] = maximm-minimm at exit,
 * info[1] = status (0 ok, 1 too few input, 2 empty first window, 3 too many
 * iterations, 4 too few valid, 5 rejection left too few). */
int bo_mmm(const double *v, int n_in, double readnoise, int minsky, int mxiter,
           double *skymod_o, double *sigma_o, double *skew_o, int *info, double *work) {
    double *sky = work;
    int nsky = 0;
    for (int i = 0; i < n_in; i++)
        if (v[i] == v[i]) sky[nsky++] = v[i]; /* mmm.py:106 */
    info[0] = 0; info[1] = 0;
    if (nsky < minsky) { /* mmm.py:109-112 */
        *skymod_o = NAN; *sigma_o = -1.0; *skew_o = 0.0; info[1] = 1;
        return 0;
    }
    int nlast = nsky - 1;
    qsort(sky, nsky, sizeof(double), cmp_dbl); /* mmm.py:119 */
    double skymid = 0.5 * sky[(nsky - 1) / 2] + 0.5 * sky[nsky / 2]; /* mmm.py:121 */
    double a = skymid - sky[0], b = sky[nsky - 1] - skymid;
    double cut1 = a < b ? a : b; /* mmm.py:123 */
    double cut2 = skymid + cut1;
    cut1 = skymid - cut1;
    int gmin = -1, gmax = -1; /* mmm.py:131 */
    for (int i = 0; i < nsky; i++)
        if (sky[i] <= cut2 && sky[i] >= cut1) { if (gmin < 0) gmin = i; gmax = i; }
    if (gmin < 0) { /* mmm.py:134-138 */
        *skymod_o = 0.0; *sigma_o = -1.0; *skew_o = 0.0; info[1] = 2;
        return 0;
    }
    /* the selection is contiguous in a sorted vector */
    double *tmp = (double *)malloc(sizeof(double) * (size_t)nsky);
    int ng = gmax - gmin + 1;
    for (int i = 0; i < ng; i++) tmp[i] = sky[gmin + i] - skymid;
    double sum = pw_sum(tmp, ng); /* mmm.py:141 */
    for (int i = 0; i < ng; i++) tmp[i] = tmp[i] * tmp[i];
    double sumsq = pw_sum(tmp, ng); /* mmm.py:142 */
    int maximm = gmax, minimm = gmin - 1; /* mmm.py:144-145 */
    double skymed = 0.5 * sky[(int)((minimm + maximm + 1) / 2.0)] +
                    0.5 * sky[(int)((minimm + maximm) / 2.0 + 1)]; /* mmm.py:149 */
    double skymn = sum / (maximm - minimm);
    double sigma = sqrt(sumsq / (maximm - minimm) - skymn * skymn);
    skymn = skymn + skymid;
    double skymod = (skymed < skymn) ? 3. * skymed - 2. * skymn : skymn; /* mmm.py:157-159 */

    int niter = 0;
    double clamp = 1, old = 0;
    int redo = 1;
    while (redo) { /* mmm.py:168 */
        niter++;
        if (niter > mxiter) { /* mmm.py:170-175 */
            *skymod_o = skymod; *sigma_o = -1.0; *skew_o = 0.0; info[1] = 3;
            info[0] = maximm - minimm; free(tmp); return niter;
        }
        if (maximm - minimm < minsky) { /* mmm.py:177-182 */
            *skymod_o = skymod; *sigma_o = -1.0; *skew_o = 0.0; info[1] = 4;
            info[0] = maximm - minimm; free(tmp); return niter;
        }
        double r = log10((double)(maximm - minimm)); /* mmm.py:186-187 */
        double r2 = (-0.1042 * r + 1.1695) * r + 0.8895;
        r = 2. > r2 ? 2. : r2;
        if (r2 != r2) r = r2; /* np.max propagates NaN */
        double cut = r * sigma + 0.5 * fabs(skymn - skymod); /* mmm.py:191 */
        cut1 = skymod - cut; cut2 = skymod + cut;

        redo = 0;
        int newmin = minimm; /* mmm.py:199-222 */
        int tst_min = sky[newmin + 1] >= cut1;
        int done = (newmin == -1) && tst_min;
        if (!done) {
            int skyind = newmin > 0 ? newmin : 0;
            if (sky[skyind] < cut1 && tst_min) done = 1;
        }
        if (!done) {
            int istep = 1 - 2 * tst_min;
            while (!done) {
                newmin += istep;
                if (newmin == -1 || newmin == nlast) done = 1;
                if (!done)
                    if (sky[newmin] <= cut1 && sky[newmin + 1] >= cut1) done = 1;
            }
            int lo, hi; /* python slice [lo:hi) */
            if (tst_min) { lo = newmin + 1; hi = minimm + 1; } else { lo = minimm + 1; hi = newmin + 1; }
            int m = hi - lo; if (m < 0) m = 0;
            for (int i = 0; i < m; i++) tmp[i] = sky[lo + i] - skymid;
            sum = sum - istep * pw_sum(tmp, m);
            for (int i = 0; i < m; i++) tmp[i] = tmp[i] * tmp[i];
            sumsq = sumsq - istep * pw_sum(tmp, m);
            redo = 1;
            minimm = newmin;
        }
        int newmax = maximm; /* mmm.py:224-248 */
        int tst_max = sky[maximm] <= cut2;
        done = (maximm == nlast) && tst_max;
        if (!done) {
            int skyind = (maximm + 1 < nlast) ? maximm + 1 : nlast;
            if (tst_max && sky[skyind] > cut2) done = 1;
        }
        if (!done) {
            int istep = -1 + 2 * tst_max;
            while (!done) {
                newmax += istep;
                if (newmax == nlast || newmax == -1) done = 1;
                if (!done)
                    if (sky[newmax] <= cut2 && sky[newmax + 1] >= cut2) done = 1;
            }
            int lo, hi;
            if (tst_max) { lo = maximm + 1; hi = newmax + 1; } else { lo = newmax + 1; hi = maximm + 1; }
            int m = hi - lo; if (m < 0) m = 0;
            for (int i = 0; i < m; i++) tmp[i] = sky[lo + i] - skymid;
            sum = sum + istep * pw_sum(tmp, m);
            for (int i = 0; i < m; i++) tmp[i] = tmp[i] * tmp[i];
            sumsq = sumsq + istep * pw_sum(tmp, m);
            redo = 1;
            maximm = newmax;
        }
        int ns = maximm - minimm; /* mmm.py:253-257 */
        if (ns < minsky) {
            *skymod_o = skymod; *sigma_o = -1.0; *skew_o = 0.0; info[1] = 5;
            info[0] = ns; free(tmp); return niter;
        }
        skymn = sum / ns; /* mmm.py:259-263 */
        double var = sumsq / ns - skymn * skymn;
        if (var < 0) var = 0;
        sigma = sqrt(var);
        skymn = skymn + skymid;

        double center = (minimm + 1 + maximm) / 2.; /* mmm.py:271-274 */
        double side = rint(0.2 * (maximm - minimm)) / 2. + 0.25;
        double j = rint(center - side), k = rint(center + side);
        if (readnoise != 0.0) { /* mmm.py:280-287 */
            double L = rint(center - 0.25), M = rint(center + 0.25), R = 0.25 * readnoise;
            while (j > 0 && k < ns - 1 &&
                   ((sky[(int)L] - sky[(int)j]) < R || (sky[(int)k] - sky[(int)M]) < R)) { j -= 1; k += 1; }
        }
        int ji = (int)j, ki = (int)k + 1; /* mmm.py:289 */
        if (ji < 0) ji = 0;
        if (ki > nsky) ki = nsky;
        skymed = pw_sum(sky + ji, ki > ji ? ki - ji : 0) / (k - j + 1);

        double dmod = (skymed < skymn) ? 3. * skymed - 2. * skymn - skymod : skymn - skymod; /* mmm.py:294-296 */
        if (dmod * old < 0) clamp = 0.5 * clamp; /* mmm.py:299-301 */
        skymod = skymod + clamp * dmod;
        old = dmod;
    }
    *skew_o = (skymn - skymod) / (sigma > 1. ? sigma : 1.); /* mmm.py:306 (python max([1.,sigma])) */
    *skymod_o = skymod; *sigma_o = sigma;
    info[0] = maximm - minimm;
    free(tmp);
    return niter;
}

/* ------------------------------------------------------------------- aper */
/* aper.py:26-433 with skyalgorithm='mmm' or setskyval.  Outputs are [nstars*naper]
 * (mag, magerr, flux, fluxerr, badflag) and [nstars] (sky, skyerr, nsky, mmm_iter).
 * compat_py2 selects the Python-2 integer row for yy at aper.py:248.
 * fragile[nstars*naper] (may be NULL): 1 when the reference's own badflag / flux for that aperture hinges on
 * rounding noise -- a non-finite or out-of-range pixel sits where Pixwt returns |w| < 1e-9 (outside the
 * circle but inside the bigrad box), so whether it is "in" the aperture (aper.py:329-330, w > 0) depends on
 * the last bits of eight atan() values.  fragile bits: 1 = such a pixel exists, 2 = a non-finite one, 4 = a finite
 * out-of-range one.  alt5[nstars*naper*5] (may be NULL) = (flux, fluxerr, mag, magerr, badflag) of the aperture with every
 * fragile pixel left OUT: together with "a non-finite one got in" (flux sum zeroed, flag 1) and "a finite one got in"
 * (same values, flag 1) these are all the outcomes the reference's arithmetic admits, so the tests check a fragile
 * aperture against that set instead of skipping it. */
int bo_aper(const void *img, int img_dt, int nrow, int ncol,
            const double *xc, const double *yc, int nstars,
            double phpadu, const double *apr, int naper, double zeropoint,
            double skyrad_in, double skyrad_out, double badlo, double badhi,
            int use_setsky, const double *setsky, int minsky, int exact,
            double readnoise, int compat_py2,
            double *mag, double *magerr, double *flux, double *fluxerr,
            double *sky, double *skyerr, double *badflag, int *nsky_o, int *mmm_iter_o,
            unsigned char *fragile, double *alt5, int nthreads) {
    if (use_setsky) { skyrad_in = 0.; skyrad_out = apr[0]; for (int k = 1; k < naper; k++) if (apr[k] > skyrad_out) skyrad_out = apr[k]; skyrad_out += 1; } /* aper.py:156-160 */
    double rinsq = skyrad_in * skyrad_in, routsq = skyrad_out * skyrad_out;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel
    {
        size_t cap = 0;
        double *skybuf = NULL, *work = NULL, *wv = NULL, *v2 = NULL, *w2 = NULL;
        unsigned char *fr = NULL;
#pragma omp for schedule(dynamic, 16)
        for (int i = 0; i < nstars; i++) {
            double *o_mag = mag + (size_t)i * naper, *o_magerr = magerr + (size_t)i * naper;
            double *o_flux = flux + (size_t)i * naper, *o_fluxerr = fluxerr + (size_t)i * naper;
            double *o_bad = badflag + (size_t)i * naper;
            for (int k = 0; k < naper; k++) { o_mag[k] = o_magerr[k] = o_flux[k] = o_fluxerr[k] = NAN; o_bad[k] = 1; if (fragile) fragile[(size_t)i * naper + k] = 0; }
            double skymod = 0., skysig = 0., skyskw = 0., skyvar = 0., sigsq = 0.;
            sky[i] = 0.; skyerr[i] = 0.; nsky_o[i] = 0; mmm_iter_o[i] = 0;
            double x = xc[i], y = yc[i];
            /* aper.py:209-210; NaN positions make the reference crash, we flag them bad */
            if (!(x >= 0.5) || !(x <= ncol - 1.5) || !(y >= 0.5) || !(y <= nrow - 1.5)) continue;
            long lx = (long)(x - skyrad_out), ly = (long)(y - skyrad_out); /* aper.py:188-199 */
            long ux = (long)(x + skyrad_out), uy = (long)(y + skyrad_out);
            if (lx < 0) lx = 0;
            if (ux > ncol - 1) ux = ncol - 1;
            if (ly < 0) ly = 0;
            if (uy > nrow - 1) uy = nrow - 1;
            long nx = ux - lx + 1, ny = uy - ly + 1;
            double dx = x - lx, dy = y - ly; /* aper.py:201-202 */
            double edge = dx - 0.5; /* aper.py:206-208 */
            if (nx + 0.5 - dx < edge) edge = nx + 0.5 - dx;
            if (dy - 0.5 < edge) edge = dy - 0.5;
            if (ny + 0.5 - dy < edge) edge = ny + 0.5 - dy;
            size_t npix = (size_t)nx * ny;
            if (npix > cap) {
                cap = npix; free(skybuf); free(work); free(wv); free(fr); free(v2); free(w2);
                fr = (unsigned char *)malloc(cap);
                v2 = (double *)malloc(cap * sizeof(double));
                w2 = (double *)malloc(cap * sizeof(double));
                skybuf = (double *)malloc(cap * sizeof(double));
                work = (double *)malloc(cap * sizeof(double));
                wv = (double *)malloc(cap * sizeof(double));
            }
            int nsky = 0;
            if (!use_setsky) {
                /* aper.py:240-267 */
                for (long r = 0; r < ny; r++) {
                    double dysq = ((double)r - dy) * ((double)r - dy);
                    for (long c = 0; c < nx; c++) {
                        double dxs = ((double)c - dx) * ((double)c - dx);
                        double rsq = dxs + dysq;
                        if (rsq >= rinsq && rsq <= routsq) {
                            double val = ld(img, img_dt, (size_t)(ly + r) * ncol + (lx + c));
                            if (badlo != badhi && (val < badlo || val > badhi)) continue;
                            skybuf[nsky++] = val;
                        }
                    }
                }
                nsky_o[i] = nsky;
                if (nsky < minsky) continue; /* aper.py:269-272 */
                int info[2];
                mmm_iter_o[i] = bo_mmm(skybuf, nsky, readnoise, minsky, 50, &skymod, &skysig, &skyskw, info, work); /* aper.py:288 */
                skyvar = skysig * skysig; /* aper.py:290-291 */
                sigsq = skyvar / nsky;
                if (skysig < 0.0) { sky[i] = skymod; skyerr[i] = skysig; continue; } /* aper.py:293-296,411-412 */
                if (skysig > 999.99) skysig = 999; /* aper.py:298-300 */
                if (skyskw < -99) skyskw = -99;
                if (skyskw > 999.9) skyskw = 999.9;
            } else { /* aper.py:302-308 */
                skymod = setsky[0]; skysig = setsky[1];
                skyvar = skysig * skysig; sigsq = skyvar / setsky[2];
                nsky_o[i] = (int)setsky[2];
            }
            sky[i] = skymod; skyerr[i] = skysig;
            for (int k = 0; k < naper; k++) { /* aper.py:310-380 */
                double ap = apr[k], area = M_PI * ap * ap;
                if (!(edge >= ap)) continue; /* aper.py:311,350-353 */
                int nw = 0, anynonfinite = 0, anybad = 0, nfr = 0;
                /* collect (value, weight) in row-major order */
                if (exact) {
                    double bigrad = ap + 0.5, smallrad = ap / sqrt(2.) - 0.5; /* aper.py:179-181 */
                    for (long r = 0; r < ny; r++)
                        for (long c = 0; c < nx; c++) {
                            double xxp = (double)c;
                            double yyp = compat_py2 ? (double)r : (double)(r * nx + c) / (double)nx; /* aper.py:246-248 */
                            double x1 = fabs(xxp - dx), y1 = fabs(yyp - dy);
                            double w;
                            if (x1 > bigrad || y1 > bigrad) continue; /* aper.py:319-320 */
                            if (x1 < smallrad && y1 < smallrad) w = 1.0; /* aper.py:316-318 */
                            else {
                                int isfr = 0;
                                w = bo_pixwt(dx, dy, ap, xxp, yyp); /* aper.py:328-329 */
                                if (fragile && fabs(w) < 1e-9) {
                                    double pv = ld(img, img_dt, (size_t)(ly + r) * ncol + (lx + c));
                                    if (!isfinite(pv)) { fragile[(size_t)i * naper + k] |= 3; isfr = 1; }
                                    else if (badlo < badhi && (pv <= badlo || pv >= badhi)) { fragile[(size_t)i * naper + k] |= 5; isfr = 1; }
                                }
                                if (w < 0.0) w = 0.0;
                                if (!(w > 0.0)) continue; /* aper.py:330 */
                                skybuf[nw] = ld(img, img_dt, (size_t)(ly + r) * ncol + (lx + c));
                                wv[nw] = w; fr[nw] = (unsigned char)isfr; nfr += isfr; nw++;
                                continue;
                            }
                            skybuf[nw] = ld(img, img_dt, (size_t)(ly + r) * ncol + (lx + c));
                            wv[nw] = w; fr[nw] = 0; nw++;
                        }
                } else { /* aper.py:336-349 */
                    int nfull = 0; int np_ = 0;
                    for (long r = 0; r < ny; r++) {
                        double dysq = ((double)r - dy) * ((double)r - dy);
                        for (long c = 0; c < nx; c++) {
                            double dxs = ((double)c - dx) * ((double)c - dx);
                            double rr = sqrt(dxs + dysq) - 0.5;
                            if (!(rr < ap)) continue;
                            double f = ap - rr;
                            if (f > 1) f = 1;
                            if (f < 0) f = 0;
                            skybuf[nw] = ld(img, img_dt, (size_t)(ly + r) * ncol + (lx + c));
                            wv[nw] = f; nw++;
                            if (f == 1.0) nfull++; else work[np_++] = f;
                        }
                    }
                    double factor = (area - nfull) / pw_sum(work, np_);
                    for (int q = 0; q < nw; q++) if (wv[q] != 1.0) wv[q] = wv[q] * factor;
                }
                if (alt5 && exact && fragile && fragile[(size_t)i * naper + k]) {
                    /* the same aperture with the fragile pixels left out (aper.py:358-403 on the shorter lists) */
                    double *o5 = alt5 + ((size_t)i * naper + k) * 5;
                    int m = 0, nf2 = 0, nb2 = 0;
                    for (int q = 0; q < nw; q++) if (!fr[q]) { v2[m] = skybuf[q]; w2[m] = wv[q]; m++; }
                    for (int q = 0; q < m; q++) if (!isfinite(v2[q])) { nf2 = 1; v2[q] = 0; }
                    if (nf2) for (int q = 0; q < m; q++) w2[q] = 0;
                    if (badlo < badhi) for (int q = 0; q < m; q++) if (v2[q] <= badlo || v2[q] >= badhi) { nb2 = 1; v2[q] = 0; }
                    for (int q = 0; q < m; q++) v2[q] = v2[q] * w2[q];
                    double f2 = pw_sum(v2, m) - skymod * area;
                    double er2 = sqrt(area * skyvar + fabs(f2) / phpadu + sigsq * (area * area));
                    o5[0] = f2; o5[1] = er2; o5[2] = NAN; o5[3] = NAN; o5[4] = (nf2 || nb2) ? 1 : 0;
                    if (f2 > 0.0) { o5[3] = 1.0857 * er2 / f2; o5[2] = zeropoint - 2.5 * log10(f2); }
                }
                o_bad[k] = 0; /* aper.py:358 */
                for (int q = 0; q < nw; q++) if (!isfinite(skybuf[q])) { anynonfinite = 1; skybuf[q] = 0; } /* aper.py:359-366 */
                if (anynonfinite) { o_bad[k] = 1; for (int q = 0; q < nw; q++) wv[q] = 0; }
                if (badlo < badhi) { /* aper.py:367-378 */
                    for (int q = 0; q < nw; q++) if (skybuf[q] <= badlo || skybuf[q] >= badhi) { anybad = 1; skybuf[q] = 0; }
                    if (anybad) o_bad[k] = 1;
                }
                for (int q = 0; q < nw; q++) work[q] = skybuf[q] * wv[q];
                double apflux = pw_sum(work, nw); /* aper.py:380 */
                apflux = apflux - skymod * area; /* aper.py:391 */
                double e1 = area * skyvar, e2 = fabs(apflux) / phpadu, e3 = sigsq * (area * area); /* aper.py:394-396 */
                double err = sqrt(e1 + e2 + e3);
                o_flux[k] = apflux; o_fluxerr[k] = err;
                if (apflux > 0.0) { /* aper.py:399-403 */
                    o_magerr[k] = 1.0857 * err / apflux;
                    o_mag[k] = zeropoint - 2.5 * log10(apflux);
                }
            }
        }
        free(skybuf); free(work); free(wv); free(fr); free(v2); free(w2);
    }
    return 0;
}

/* ------------------------------------------------------------ PSF model */
/* daoerf.py:76-90 */
static double bo_cdf(double x) { return 0.5 + 0.5 * erf(x * (1. / sqrt(2.))); }

/* daoerf.py:49-74 ; pder[5] */
double bo_daoerf(double x, double y, const double *a, double *pder) {
    const double norm = 2.506628275;
    double u2 = (x - a[1] + 0.5) / a[3], u1 = (x - a[1] - 0.5) / a[3];
    double v2 = (y - a[2] + 0.5) / a[4], v1 = (y - a[2] - 0.5) / a[4];
    double fx = norm * a[3] * (bo_cdf(u2) - bo_cdf(u1));
    double fy = norm * a[4] * (bo_cdf(v2) - bo_cdf(v1));
    double f = a[0] * fx * fy;
    if (pder) {
        pder[0] = fx * fy;
        double uplus = exp(-0.5 * (u2 * u2)), uminus = exp(-0.5 * (u1 * u1));
        pder[1] = a[0] * fy * (-uplus + uminus);
        double vplus = exp(-0.5 * (v2 * v2)), vminus = exp(-0.5 * (v1 * v1));
        pder[2] = a[0] * fx * (-vplus + vminus);
        pder[3] = a[0] * fy * (fx / a[3] + u1 * uminus - u2 * uplus);
        pder[4] = a[0] * fx * (fy / a[4] + v1 * vminus - v2 * vplus);
    }
    return f;
}

/* rinter.py:98-195 for one point.  p is the flattened table of length plen with
 * row length nc; indices past the end read the last element (rinter.py:117-135). */
static inline double P(const double *p, long plen, long idx) { return p[idx < plen ? idx : plen - 1]; }
static void cubic_row(const double *p, long plen, long k, double xd, double *val, double *der) {
    double p_1 = P(p, plen, k - 1), p0 = P(p, plen, k), p1 = P(p, plen, k + 1), p2 = P(p, plen, k + 2);
    double c1 = 0.5 * (p1 - p_1);                       /* rinter.py:164 */
    double c2 = 2. * p1 + p_1 - 0.5 * (5. * p0 + p2);   /* rinter.py:165 */
    double c3 = 0.5 * (3. * (p0 - p1) + p2 - p_1);      /* rinter.py:166 */
    *val = xd * (xd * (xd * c3 + c2) + c1) + p0;        /* rinter.py:170-177 */
    *der = xd * (xd * c3 * 3. + 2. * c2) + c1;          /* rinter.py:181-184 */
}
double bo_rinter(const double *p, int nr, int nc, double x, double y, double *dfdx, double *dfdy) {
    long plen = (long)nr * nc;
    long i = (long)x, j = (long)y; /* rinter.py:98 */
    double xd = x - i, yd = y - j;
    long x_1 = (long)nc * (j - 1) + i, x0 = x_1 + nc, x1 = x0 + nc, x2 = x1 + nc; /* rinter.py:101-104 */
    double y_1, y0, y1, y2, dy_1, dy0, dy1, dy2;
    cubic_row(p, plen, x_1, xd, &y_1, &dy_1);
    cubic_row(p, plen, x0, xd, &y0, &dy0);
    cubic_row(p, plen, x1, xd, &y1, &dy1);
    cubic_row(p, plen, x2, xd, &y2, &dy2);
    if (dfdx) { /* rinter.py:185-188 */
        double d1 = 0.5 * (dy1 - dy_1);
        double d2 = 2. * dy1 + dy_1 - 0.5 * (5. * dy0 + dy2);
        double d3 = 0.5 * (3. * (dy0 - dy1) + dy2 - dy_1);
        *dfdx = yd * (yd * (yd * d3 + d2) + d1) + dy0;
    }
    double d1 = 0.5 * (y1 - y_1); /* rinter.py:190-193 */
    double d2 = 2. * y1 + y_1 - 0.5 * (5. * y0 + y2);
    double d3 = 0.5 * (3. * (y0 - y1) + y2 - y_1);
    if (dfdy) *dfdy = yd * (yd * d3 * 3. + 2. * d2) + d1; /* rinter.py:195 */
    return yd * (yd * (yd * d3 + d2) + d1) + y0;
}

/* dao_value.py:60-107 for n points; returns 1 if the edge guard fired (all outputs 0). */
int bo_dao_value(const double *xx, const double *yy, int n, const double *gauss,
                 const double *psf, int npsf, double *value, double *dvdx, double *dvdy) {
    double half = (double)(npsf - 1) / 2.;
    double xmin = INFINITY, xmax = -INFINITY, ymin = INFINITY, ymax = -INFINITY;
    for (int q = 0; q < n; q++) {
        double x = 2. * xx[q] + half, y = 2. * yy[q] + half;
        if (x < xmin) xmin = x;
        if (x > xmax) xmax = x;
        if (y < ymin) ymin = y;
        if (y > ymax) ymax = y;
    }
    if (xmin < 1. || xmax > npsf - 2. || ymin < 1. || ymax > npsf - 2.) { /* dao_value.py:70-77 */
        for (int q = 0; q < n; q++) { value[q] = 0; if (dvdx) { dvdx[q] = 0; dvdy[q] = 0; } }
        return 1;
    }
    for (int q = 0; q < n; q++) {
        double pder[5], dfdx, dfdy;
        double x = 2. * xx[q] + half, y = 2. * yy[q] + half;
        double e = bo_daoerf(xx[q], yy[q], gauss, pder);
        double val = bo_rinter(psf, npsf, npsf, x, y, &dfdx, &dfdy);
        value[q] = e + val;                       /* dao_value.py:100-103 */
        if (dvdx) { dvdx[q] = 2. * dfdx - pder[1]; dvdy[q] = 2. * dfdy - pder[2]; } /* dao_value.py:104-105 */
    }
    return 0;
}

/* --------------------------------------------------------- 3x3 inverse */
/* scipy.linalg.inv as installed beside the reference (scipy 1.18) detects a
 * symmetric positive-definite input and takes the Cholesky route (potrf+potri);
 * otherwise LU with partial pivoting (getrf+getri).  Returns 0 ok, 1 singular. */
static int inv3(double c[3][3]) {
    double L[3][3] = {{0}};
    int spd = (c[0][1] == c[1][0] && c[0][2] == c[2][0] && c[1][2] == c[2][1]);
    if (spd) {
        for (int j = 0; j < 3 && spd; j++) {
            double s = c[j][j];
            for (int k = 0; k < j; k++) s -= L[j][k] * L[j][k];
            if (!(s > 0.0)) { spd = 0; break; }
            L[j][j] = sqrt(s);
            for (int i = j + 1; i < 3; i++) {
                double t = c[i][j];
                for (int k = 0; k < j; k++) t -= L[i][k] * L[j][k];
                L[i][j] = t / L[j][j];
            }
        }
    }
    if (spd) {
        /* M = inv(L) (lower), inv(A) = M^T M */
        double M[3][3] = {{0}};
        for (int j = 0; j < 3; j++) {
            M[j][j] = 1.0 / L[j][j];
            for (int i = j + 1; i < 3; i++) {
                double s = 0;
                for (int k = j; k < i; k++) s += L[i][k] * M[k][j];
                M[i][j] = -s / L[i][i];
            }
        }
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                double s = 0;
                for (int k = (i > j ? i : j); k < 3; k++) s += M[k][i] * M[k][j];
                c[i][j] = s;
            }
        return 0;
    }
    double a[3][3]; int piv[3];
    memcpy(a, c, sizeof(a));
    for (int j = 0; j < 3; j++) {
        int p = j; double best = fabs(a[j][j]);
        for (int i = j + 1; i < 3; i++) if (fabs(a[i][j]) > best) { best = fabs(a[i][j]); p = i; }
        piv[j] = p;
        if (a[p][j] == 0.0) return 1;
        if (p != j) for (int k = 0; k < 3; k++) { double t = a[j][k]; a[j][k] = a[p][k]; a[p][k] = t; }
        for (int i = j + 1; i < 3; i++) a[i][j] /= a[j][j];
        for (int i = j + 1; i < 3; i++)
            for (int k = j + 1; k < 3; k++) a[i][k] -= a[i][j] * a[j][k];
    }
    /* inv(U) in place (dtrti2) */
    for (int j = 0; j < 3; j++) {
        a[j][j] = 1.0 / a[j][j];
        double ajj = -a[j][j];
        for (int jj = 0; jj < j; jj++) {
            double t = a[jj][j];
            for (int i = 0; i < jj; i++) a[i][j] += t * a[i][jj];
            a[jj][j] = t * a[jj][jj];
        }
        for (int i = 0; i < j; i++) a[i][j] *= ajj;
    }
    /* solve inv(A)*L = inv(U) (dgetri) */
    for (int j = 2; j >= 0; j--) {
        double w[3] = {0, 0, 0};
        for (int i = j + 1; i < 3; i++) { w[i] = a[i][j]; a[i][j] = 0.0; }
        for (int k = j + 1; k < 3; k++)
            for (int i = 0; i < 3; i++) a[i][j] -= a[i][k] * w[k];
    }
    for (int j = 1; j >= 0; j--) {
        int p = piv[j];
        if (p != j) for (int i = 0; i < 3; i++) { double t = a[i][j]; a[i][j] = a[i][p]; a[i][p] = t; }
    }
    memcpy(c, a, sizeof(a));
    return 0;
}

void bo_inv3(double *c9, int *status) { double c[3][3]; memcpy(c, c9, sizeof(c)); *status = inv3(c); memcpy(c9, c, sizeof(c)); }

/* ------------------------------------------------------------------ pkfit */
#define BO_PKFIT 0          /* pkfit.py:132 (no noise/mask image, analytic variance) */
#define BO_PKFIT_NOISE 1    /* pkfit_noise.py:131 */
#define BO_PKFIT_NORECENT 2 /* pkfit_norecent_noise.py:132 */
#define BO_PKFIT_NORECENTER 3 /* pkfit_norecenter.py:346 (Python-2 semantics: it raises under Python 3, :435) */

typedef struct { double errmag, chi, sharp, scale, x, y, chi2; int niter; } bo_fit;

static void fit_nan(bo_fit *o, int keep_chi, int keep_sharp) {
    o->scale = NAN; o->errmag = NAN;
    if (!keep_chi) o->chi = NAN;
    if (!keep_sharp) o->sharp = NAN;
}

static void pkfit_one(int variant, const void *f, int f_dt, const void *fnoise, int n_dt,
                      const uint8_t *fmask, int ny, int nx, const double *gauss,
                      const double *psf, int npsf, double ronois, double phpadu,
                      double scale, double x, double y, double sky, double radius,
                      int maxiter, int recenter, bo_fit *o) {
    double pkerr = 0.027 / ((gauss[3] * gauss[4]) * (gauss[3] * gauss[4]));
    double clamp[3] = {1, 1, 1}, dtold[3] = {0, 0, 0};
    int niter = 0;
    double chiold = 1., chi = NAN, sharp = NAN, errmag = NAN, chi2 = NAN;
    /* pkfit_norecenter: analytic variance like pkfit.py, position frozen; errmag = chi = sharp = 100000 until computed (:352-354) */
    const int analytic = (variant == BO_PKFIT || variant == BO_PKFIT_NORECENTER);
    if (variant == BO_PKFIT_NORECENTER) { chi = 100000; sharp = 100000; errmag = 100000; }
    o->chi2 = NAN;
    int cap = (int)((2 * fabs(radius) + 4) * (2 * fabs(radius) + 4)) + 16;
    double *buf = (double *)malloc(sizeof(double) * (size_t)cap * 16);
    double *gdx = buf, *gdy = gdx + cap, *t0 = gdy + cap, *t1 = t0 + cap, *t2 = t1 + cap;
    double *fsub = t2 + cap, *sig = fsub + cap, *sigsq = sig + cap, *df = sigsq + cap;
    double *relerr = df + cap, *rsqv = relerr + cap, *rhosq = rsqv + cap, *wt = rhosq + cap;
    double *tmp = wt + cap, *dvx = tmp + cap, *dvy = dvx + cap;
    int *gidx = (int *)malloc(sizeof(int) * (size_t)cap);

#define RETURN_FIT do { o->errmag = errmag; o->chi = chi; o->sharp = sharp; o->niter = niter; \
        o->scale = scale; o->x = x; o->y = y; o->chi2 = chi2; free(buf); free(gidx); return; } while (0)
#define RETURN_NAN(keep_chi, keep_sharp) do { scale = NAN; errmag = NAN; if (!(keep_chi)) chi = NAN; \
        if (!(keep_sharp)) sharp = NAN; RETURN_FIT; } while (0)

    if (variant == BO_PKFIT_NORECENT && (isnan(x) || isnan(y))) RETURN_NAN(0, 0); /* pkfit_norecent_noise.py:156-164 */

    int loop = 1;
    while (loop) {
        niter++;
        if (variant != BO_PKFIT_NORECENT && (isnan(x) || isnan(y))) RETURN_NAN(0, 0); /* pkfit.py:158-166 */
        long ixlo = (long)(x - radius); if (ixlo < 0) ixlo = 0; /* pkfit.py:168-177 */
        long iylo = (long)(y - radius); if (iylo < 0) iylo = 0;
        long ixhi = (long)(x + radius) + 1; if (ixhi > nx - 1) ixhi = nx - 1;
        long iyhi = (long)(y + radius) + 1; if (iyhi > ny - 1) iyhi = ny - 1;
        long ixx = ixhi - ixlo + 1, iyy = iyhi - iylo + 1;
        double radsq = radius * radius;
        int ngood = 0;
        for (long r = 0; r < iyy; r++) {
            double dyv = (double)(r + iylo) - y, dysq = dyv * dyv;
            for (long c = 0; c < ixx; c++) {
                double dxv = (double)(c + ixlo) - x, dxsq = dxv * dxv;
                double rsq = (dxsq + dysq) / radsq; /* pkfit.py:184 */
                if (!(rsq < 1.)) continue;
                size_t gi = (size_t)(iylo + r) * nx + (ixlo + c);
                if (!analytic) { /* pkfit_noise.py:206-208 */
                    if (!(ld(fnoise, n_dt, gi) > 0)) continue;
                    if (fmask[gi] != 0) continue;
                }
                if (ngood >= cap) continue;
                gdx[ngood] = dxv; gdy[ngood] = dyv; rsqv[ngood] = rsq;
                rhosq[ngood] = dxsq / (gauss[3] * gauss[3]) + dysq / (gauss[4] * gauss[4]); /* pkfit.py:292 */
                fsub[ngood] = ld(f, f_dt, gi);
                if (!analytic) { sig[ngood] = ld(fnoise, n_dt, gi); sigsq[ngood] = sig[ngood] * sig[ngood]; }
                ngood++;
            }
        }
        if (ngood == 0) RETURN_NAN(0, 0); /* pkfit.py:211-219 */
        bo_dao_value(gdx, gdy, ngood, gauss, psf, npsf, t0, dvx, dvy); /* pkfit.py:224 */
        for (int q = 0; q < ngood; q++) { t1[q] = -scale * dvx[q]; t2[q] = -scale * dvy[q]; } /* pkfit.py:243-244 */
        for (int q = 0; q < ngood; q++) df[q] = fsub[q] - scale * t0[q] - sky; /* pkfit.py:262 */
        if (analytic) { /* pkfit.py:277-282 */
            for (int q = 0; q < ngood; q++) {
                double fpos = fsub[q] - df[q];
                if (fpos < 0.) fpos = 0;
                sigsq[q] = fpos / phpadu + ronois + (0.0075 * fpos) * (0.0075 * fpos) +
                           (pkerr * (fpos - sky)) * (pkerr * (fpos - sky));
                sig[q] = sqrt(sigsq[q]);
            }
        }
        for (int q = 0; q < ngood; q++) relerr[q] = df[q] / sig[q]; /* pkfit.py:283 */
        int m = ngood; /* compacted length */
        for (int q = 0; q < ngood; q++) gidx[q] = q;
        if (niter >= 2) { /* pkfit.py:296-321 */
            int nbad = 0;
            for (int q = 0; q < ngood; q++) if (fabs(relerr[q] / chiold) >= 10.) nbad++;
            if (nbad == ngood && !(variant == BO_PKFIT_NORECENTER && nbad <= 1)) { /* pkfit_norecenter.py:509: only if len(badpix) > 1 */
                if (variant == BO_PKFIT_NORECENT) RETURN_NAN(0, 1); /* pkfit_norecent_noise.py:347-356 */
                RETURN_NAN(1, 1);                                   /* pkfit.py:303-309 */
            }
            if (nbad > 0) {
                m = 0;
                for (int q = 0; q < ngood; q++) {
                    if (fabs(relerr[q] / chiold) >= 10.) continue;
                    fsub[m] = fsub[q]; df[m] = df[q]; sigsq[m] = sigsq[q]; sig[m] = sig[q];
                    relerr[m] = relerr[q]; rsqv[m] = rsqv[q]; rhosq[m] = rhosq[q];
                    gidx[m] = q; m++;
                }
            }
        }
        /* pkfit.py keeps t un-compacted and slices t[:,0:lenwt] (pkfit.py:369-372):
         * the k-th surviving pixel is paired with the k-th ORIGINAL pixel's t. */
        if (analytic) for (int q = 0; q < m; q++) gidx[q] = q;
        for (int q = 0; q < m; q++) wt[q] = 5. / (5. + rsqv[q] / (1. - rsqv[q])); /* pkfit.py:323 */
        int nl = 0; double numer, denom;
        {
            double *a1 = tmp, *a2 = dvx, *a3 = dvy; /* scratch */
            for (int q = 0; q < m; q++) {
                if (!(rhosq[q] <= 36.)) continue; /* pkfit.py:324 */
                double rh = 0.5 * rhosq[q];
                double dfdsig = exp(-rh) * (rh - 1.); /* pkfit.py:336 */
                double s;
                if (analytic) { /* pkfit.py:341-346 */
                    double fpos = fsub[q];
                    if (fsub[q] - sky < 0.) fpos = sky;
                    s = fpos / phpadu + ronois + (0.0075 * fpos) * (0.0075 * fpos) +
                        (pkerr * (fpos - sky)) * (pkerr * (fpos - sky));
                } else s = sig[q] * sig[q]; /* pkfit_noise.py:361 */
                /* pkfit.py:350 pairs dfdsig with df[0:len(lilrho)] (first nl entries) */
                double dfv = analytic ? df[nl] : df[q];
                a1[nl] = dfdsig * dfv / s;
                a2[nl] = dfdsig * dfdsig / s;
                a3[nl] = dfv * dfv / s;
                nl++;
            }
            if (nl == 0) RETURN_NAN(0, 0); /* pkfit.py:325-333 */
            /* pkfit_norecenter.py:537,570-573 multiplies df[lilrho] (nl long) with wt (m long): numpy raises unless every pixel is
             * within 6 sigma; reported as a failed fit here */
            if (variant == BO_PKFIT_NORECENTER && nl != m) RETURN_NAN(0, 0);
            numer = pw_sum(a1, nl); denom = pw_sum(a2, nl);
            if (variant == BO_PKFIT_NORECENT) chi2 = pw_sum(a3, nl) / (nl - 1); /* pkfit_norecent_noise.py:409-410 */
        }
        for (int q = 0; q < m; q++) tmp[q] = wt[q] * fabs(relerr[q]);
        chi = pw_sum(tmp, m); /* pkfit.py:359 */
        double sumwt = pw_sum(wt, m);
        double wmin = INFINITY;
        for (int q = 0; q < m; q++) { wt[q] = wt[q] / sigsq[q]; } /* pkfit.py:362 */
        if (variant == BO_PKFIT_NORECENT) { /* pkfit_norecent_noise.py:416-424 (python min(): first element wins NaN ties) */
            wmin = wt[0];
            for (int q = 1; q < m; q++) if (wt[q] < wmin) wmin = wt[q];
            if (wmin > 100000) RETURN_NAN(0, 0);
        }
        if (niter >= 2)
            for (int q = 0; q < m; q++) wt[q] = wt[q] / (1. + pow(0.4 * relerr[q] / chiold, 8)); /* pkfit.py:364 */
        double v[3], c[3][3];
        const double *tt[3] = {t0, t1, t2};
        for (int kk = 0; kk < 3; kk++) { /* pkfit.py:369-372 */
            for (int q = 0; q < m; q++) tmp[q] = df[q] * tt[kk][gidx[q]] * wt[q];
            v[kk] = pw_sum(tmp, m);
            for (int ll = 0; ll < 3; ll++) {
                for (int q = 0; q < m; q++) tmp[q] = tt[kk][gidx[q]] * tt[ll][gidx[q]] * wt[q];
                c[ll][kk] = pw_sum(tmp, m);
            }
        }
        if (sumwt > 3.0) { /* pkfit.py:379-381 */
            chi = 1.2533 * chi * sqrt(1. / (sumwt * (sumwt - 3.)));
            chiold = ((sumwt - 3.) * chi + 3.) / sumwt;
        }
        double csum = 0;
        for (int a_ = 0; a_ < 3; a_++) for (int b_ = 0; b_ < 3; b_++) csum += c[a_][b_];
        int skip_inv = 0;
        if (variant == BO_PKFIT_NOISE || variant == BO_PKFIT_NORECENTER) { /* pkfit_noise.py:399-410, pkfit_norecenter.py:586-594: NaN => inversion skipped; inf => inv raises => NaN return */
            if (isnan(csum)) skip_inv = 1;
            else if (isinf(csum)) RETURN_NAN(0, 0);
        } else if (isnan(csum) || isinf(csum)) RETURN_NAN(0, 0); /* pkfit.py:383-406 */
        if (!skip_inv && inv3(c)) RETURN_NAN(0, 0);
        double dt[3];
        for (int j = 0; j < 3; j++) dt[j] = v[0] * c[0][j] + v[1] * c[1][j] + v[2] * c[2][j]; /* pkfit.py:408-409 */
        if (variant == BO_PKFIT_NORECENT && isinf(c[0][0])) RETURN_NAN(0, 0); /* pkfit_norecent_noise.py:474-483 */
        for (int j = 0; j < 3; j++) if (dtold[j] * dt[j] < -1.e-38) clamp[j] = clamp[j] / 2.; /* pkfit.py:421-423 */
        double adt[3];
        for (int j = 0; j < 3; j++) { dtold[j] = dt[j]; adt[j] = fabs(dt[j]); }
        double denom2 = dt[0] / (5.25 * scale); /* pkfit.py:427-429 */
        if (denom2 < (-1 * dt[0] / (0.84 * scale))) denom2 = (-1 * dt[0] / (0.84 * scale));
        scale = scale + dt[0] / (1 + denom2 / clamp[0]);
        if (variant == BO_PKFIT_NOISE || (variant == BO_PKFIT && recenter)) { /* pkfit.py:430-432 */
            x = x + dt[1] / (1. + adt[1] / (0.5 * clamp[1]));
            y = y + dt[2] / (1. + adt[2] / (0.5 * clamp[2]));
        }
        int redo = 0;
        sharp = 2. * gauss[3] * gauss[4] * numer / (gauss[0] * scale * denom); /* pkfit.py:440 */
        errmag = chiold * sqrt(c[0][0]);
        if (variant == BO_PKFIT) { /* pkfit.py:442 (python max: b if b > a else a); pkfit_norecenter.py:631 has the `or` form */
            double a_ = 0.05 * errmag, b_ = 0.001 * scale;
            double mx = (b_ > a_) ? b_ : a_;
            if (adt[0] > mx) redo = 1;
        } else if (adt[0] > 0.05 * errmag || adt[0] > 0.001 * scale) redo = 1; /* pkfit_noise.py:445 */
        if (adt[1] > 0.01 || adt[2] > 0.01) redo = 1;
        if (niter >= 3) loop = 0;
        if (redo && errmag <= 1.9995 && niter < maxiter) loop = 1;
    }
    RETURN_FIT;
#undef RETURN_FIT
#undef RETURN_NAN
}

/* Batched driver: one reference call per star (README.md:73-75). out arrays [nstars]. */
int bo_pkfit(int variant, const void *f, int f_dt, const void *fnoise, int n_dt,
             const uint8_t *fmask, int ny, int nx, const double *gauss, const double *psf, int npsf,
             double ronois, double phpadu, const double *scale, const double *x, const double *y,
             const double *sky, double radius, int maxiter, int recenter, int nstars,
             double *errmag, double *chi, double *sharp, int *niter, double *scale_o,
             double *x_o, double *y_o, double *chi2, int nthreads) {
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 16)
    for (int i = 0; i < nstars; i++) {
        bo_fit o;
        pkfit_one(variant, f, f_dt, fnoise, n_dt, fmask, ny, nx, gauss, psf, npsf, ronois, phpadu,
                  scale[i], x[i], y[i], sky[i], radius, maxiter, recenter, &o);
        errmag[i] = o.errmag; chi[i] = o.chi; sharp[i] = o.sharp; niter[i] = o.niter;
        scale_o[i] = o.scale; x_o[i] = o.x; y_o[i] = o.y; chi2[i] = o.chi2;
    }
    return 0;
}

int bo_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
