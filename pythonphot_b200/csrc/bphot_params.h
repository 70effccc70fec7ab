// bphot_params.h -- kernel argument blocks (plain data, passed by value to the kernels).
#pragma once
#include <stdint.h>

#define BPHOT_MAX_APER 16

namespace bphot {

struct AperParams {
    // image: [nrow, ncol] row-major, x = column; `pitch` = elements between rows
    const void* img;
    int nrow, ncol;
    long long pitch;
    // star list
    const double* xc;
    const double* yc;
    int nstars;
    // aper.py:26-29 keywords
    double phpadu, zeropoint;
    double apr[BPHOT_MAX_APER];
    int naper;
    double skyrad_in, skyrad_out;  // already [0, max(apr)+1] when use_setsky (aper.py:156-160)
    double badlo, badhi;
    float badlo_t, badhi_t;  // fp32 images: smallest float >= badlo, largest float <= badhi (exact fp32 form of the window test)
    int use_setsky;
    double setsky[3];
    int minsky, exact, compat_py2, mxiter;
    int* work_counter;  // stage 1 / stage 2: zeroed counter the CTAs (warps) draw their next star from, one per launch
    int retry_wide;   // stage 1, second pass: only stars the 96-thread pass marked -3 (annulus larger than 96 x K keys)
    int sky_sigclip;  // 1: sigma-clipped mean sky (aper.py:275-282) instead of mmm
    double readnoise;
    int sort_cap;  // power of two >= the largest possible annulus pixel count
    // stage-1 -> stage-2 hand-over: sorted annulus keys of stars [s0, s0 + scount), `sstride` keys per star
    void* skeys;
    long long sstride;
    int s0, scount;
    int row0;  // stage 1: scratch row of star s0 (a stage-2 launch covers several stage-1 launches)
    // index of each sorted row for the thread-per-star mmm (aper_kernel.cu sort_index): null = not built
    void* sidx;
    long long idx_stride;  // bytes per star
    int span_rows;  // rows of the per-CTA column-span table (>= stamp height)
    const int* order;  // visiting order of the stars (image-tile sorted), or null
    // outputs
    double* mag;      // [nstars, naper]
    double* magerr;   // [nstars, naper]
    double* flux;     // [nstars, naper]
    double* fluxerr;  // [nstars, naper]
    double* sky;      // [nstars]
    double* skyerr;   // [nstars]
    uint8_t* badflag; // [nstars, naper]
    int* nsky;        // [nstars]  sky pixels handed to mmm (aper.py:267)
    int* mmm_iter;    // [nstars]  mmm clipping iterations
};

struct PkfitParams {
    int variant;  // 0 pkfit.py, 1 pkfit_noise.py, 2 pkfit_norecent_noise.py
    const void* img;
    const void* noise;
    const uint8_t* mask;
    int nrow, ncol;
    long long pitch, npitch, mpitch;
    double gauss[5];
    const double* psf;  // [npsf, npsf] fp64, device
    int npsf;
    double ronois, phpadu, radius;
    int maxiter, recenter;
    const double* scale;
    const double* x;
    const double* y;
    const double* sky;
    int nstars;
    int* work_counter;  // zeroed counter the warps draw their next star from
    const int* order;  // visiting order of the stars (image-tile sorted), or null
    int cap;   // per-star pixel-list capacity
    int acap;  // columns / rows of the fitting box (capacity of the per-axis Gaussian factors)
    double* errmag;
    double* chi;
    double* sharp;
    int* niter;
    double* scale_out;
    double* x_out;
    double* y_out;
    double* chi2;
};

}  // namespace bphot
