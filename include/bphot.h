/*
 * bphot.h -- C ABI of libbphot.so, the B200 (sm_100a) batched stellar-photometry library.
 *
 * The reference (djones1040/PythonPhot @ 7992eb6) is pure Python and has no FFI of its own; the
 * boundary replaced here is its per-star Python API.  Each entry point
 * below names the reference callable whose per-star loop it executes for a whole star list.
 * Plain pointers and sizes only.  Two flavours per operation:
 *
 *   *_device  every array argument is a DEVICE pointer (e.g. torch.Tensor.data_ptr()); the work
 *             is enqueued on `stream` (a cudaStream_t, NULL = default stream) and the call
 *             returns without synchronising.  Small parameter vectors (apr, setskyval, gauss)
 *             are HOST pointers, copied at call time.
 *   *_host    every array argument is a HOST pointer; the call uploads, runs, downloads and
 *             synchronises on CUDA device `device`.
 *
 * Conventions shared with the reference:
 *   - images are [nrow, ncol] row-major, x = column, y = row; `pitch` = elements between rows;
 *   - per-star failures are reported in band (NaN outputs, sigma = -1, badflag = 1, niter
 *     sentinels) exactly as the reference reports them; the int return value is only about the
 *     call itself: 0 = ok, negative = BPHOT_ERR_*; bphot_last_error() has the message.
 *   - the library never retains or frees caller memory.
 */
#ifndef BPHOT_H
#define BPHOT_H

#ifdef __cplusplus
extern "C" {
#endif

#define BPHOT_VERSION 100 /* 0.1.0 */

#define BPHOT_F32 0
#define BPHOT_F64 1

#define BPHOT_PKFIT 0          /* PythonPhot/pkfit.py:132            pkfit_class.pkfit */
#define BPHOT_PKFIT_NOISE 1    /* PythonPhot/pkfit_noise.py:131      pkfit_class.pkfit_noise */
#define BPHOT_PKFIT_NORECENT 2 /* PythonPhot/pkfit_norecent_noise.py:132 pkfit_class.pkfit_norecent_noise */
#define BPHOT_PKFIT_NORECENTER 3 /* PythonPhot/pkfit_norecenter.py:346  pkfit_class.pkfit_norecenter (its Python-2 semantics) */

#define BPHOT_MAX_APERTURES 16

#define BPHOT_APER_PY2 1       /* bphot_aper_*: `flags` bits */
#define BPHOT_APER_SIGMACLIP 2
#define BPHOT_APER_NARROW_CTAS 256 /* testing aid, see `flags` below */

#define BPHOT_OK 0
#define BPHOT_ERR_ARG (-1)      /* bad argument (NULL pointer, naper out of range, ...) */
#define BPHOT_ERR_CUDA (-2)     /* CUDA runtime error, see bphot_last_error() */
#define BPHOT_ERR_CAPACITY (-3) /* fitting box larger than one CTA's shared memory, output columns too short, no-alloc scratch too small */
#define BPHOT_ERR_NODEVICE (-4) /* no CUDA device: there is no CPU fallback */

int bphot_version(void);
/* the build knobs this library was compiled with ("AP_ASM_SELECT=1 AP_BATCH=65536 ..."), for A/B logs */
const char* bphot_build_flags(void);
const char* bphot_last_error(void);
int bphot_device_count(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
long long bphot_launch_count(void);
/*
 * Measurement aid (bench.py's roofline): while enabled, every launch of the four hot kernels is bracketed by CUDA
 * events on its own stream and the owning call synchronises before it returns.  bphot_profile(1) also clears the
 * accumulators; returns the previous setting.  bphot_kernel_times fills ms[4] / launches[4] with the summed
 * durations and launch counts of 0 aper_stage1 (gather + sort), 1 aper_sums, 2 aper_stage2 (mmm), 3 pkfit.
 */
/*
 * Copies a PAGE-LOCKED host buffer into device memory with a kernel that reads the host pages directly (zero-copy loads
 * over PCIe) instead of a cudaMemcpyAsync: the transfer then does not queue on the copy engine behind image uploads that
 * are already in flight (a star list of 16 MB otherwise waits for 600 MB of exposure).  Returns BPHOT_ERR_ARG when h_src is
 * not page-locked (the caller falls back to cudaMemcpyAsync).
 */
int bphot_fetch_host(void* d_dst, const void* h_src, long long nbytes, void* stream);
/* slots of the per-star pixel list the fit kernels allocate for this fitting radius: a rigorous upper bound on the number
 * of pixel centres strictly inside a disc of that radius, for any centre (tests/test_abi_cpu.py checks it against brute force) */
int bphot_pkfit_capacity(double radius);
int bphot_profile(int enable);
int bphot_kernel_times(double* ms, long long* launches);

/*
 * aper.aper(image, xc, yc, phpadu, apr, zeropoint, skyrad, badpix, setskyval, minsky,
 *           skyalgorithm='mmm', exact, readnoise)            PythonPhot/aper.py:26-433
 * with its callees mmm.mmm (mmm.py:7) and pixwt.Pixwt (pixwt.py:8).
 *   apr[naper]        aperture radii (host)
 *   skyrad_in/out     skyrad[0], skyrad[1]
 *   badlo, badhi      badpix[0], badpix[1] (equal = no bad-pixel test, aper.py:260)
 *   use_setsky        1 = setskyval given; setsky[3] = (sky, sigma, nsky) (aper.py:156-160, 302-308)
 *   minsky            20 when the reference default [] is used (aper.py:148)
 *   exact             1 = pixwt exact overlap (aper.py:313-333), 0 = polygon approximation
 *   flags             bit 0 (BPHOT_APER_PY2): 0 = Python-3 semantics of aper.py:246-248 (yy = row + col/nx), 1 = Python-2
 *                     integer row (what the reference was written for); only matters when exact=1
 *                     bit 1 (BPHOT_APER_SIGMACLIP): skyalgorithm='sigmaclipping' -- sky = mean, skyerr = std of the
 *                     annulus after scipy.stats.sigmaclip(low=4, high=4) (aper.py:275-282) instead of mmm;
 *                     mmm_iter then counts clipping passes
 *                     bit 8 (BPHOT_APER_NARROW_CTAS): always run the sort stage with 96-thread CTAs first, whatever the
 *                     expected annulus count; stars that do not fit are redone by the 128-thread pass.  Results do
 *                     not change; the parity tests use it to reach the retry pass, which default radii never need
 * Outputs (fp64 unless noted): mag, magerr, flux, fluxerr [nstars*naper]; sky, skyerr [nstars];
 * badflag uint8 [nstars*naper]; nsky int32 [nstars] (annulus pixels handed to mmm, aper.py:267);
 * mmm_iter int32 [nstars] (clipping iterations mmm ran).
 */
int bphot_aper_device(const void* d_image, int image_dtype, int nrow, int ncol, long long pitch,
                      const double* d_xc, const double* d_yc, int nstars,
                      double phpadu, const double* apr, int naper, double zeropoint,
                      double skyrad_in, double skyrad_out, double badlo, double badhi,
                      int use_setsky, const double* setsky, int minsky, int exact,
                      double readnoise, int flags,
                      double* d_mag, double* d_magerr, double* d_flux, double* d_fluxerr,
                      double* d_sky, double* d_skyerr, unsigned char* d_badflag,
                      int* d_nsky, int* d_mmm_iter, void* stream);
/*
 * The same call for an image that is STILL BEING UPLOADED: the caller copies the image into d_image in bands of
 * `band_rows` rows, top to bottom, on a copy stream of its own, and records band_events[k] (cudaEvent_t) once rows
 * [0, (k+1)*band_rows) are resident.  Stars are visited in image-tile order, so every batch of the sort stage waits only
 * for the band that covers its lowest stamp row and the upload runs under the kernels instead of in front of them
 * (bench.py's end-to-end leg; pythonphot_b200.aper.aper does this for host arrays).  nbands = 0 is bphot_aper_device.
 */
int bphot_aper_device_banded(const void* d_image, int image_dtype, int nrow, int ncol, long long pitch,
                             const double* d_xc, const double* d_yc, int nstars,
                             double phpadu, const double* apr, int naper, double zeropoint,
                             double skyrad_in, double skyrad_out, double badlo, double badhi,
                             int use_setsky, const double* setsky, int minsky, int exact,
                             double readnoise, int flags,
                             double* d_mag, double* d_magerr, double* d_flux, double* d_fluxerr,
                             double* d_sky, double* d_skyerr, unsigned char* d_badflag,
                             int* d_nsky, int* d_mmm_iter,
                             const void* const* band_events, int nbands, int band_rows, void* stream);
/*
 * bphot_aper_device for callers that may not allocate.  NO counterpart in the reference: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.  Same arguments,
 * same outputs; the call makes no cudaMalloc / cudaMallocAsync / cudaFree and uses no allocating library routine.  Its
 * scratch (work counters, sorted sky-annulus rows, row indexes) lives in ONE static device array of
 * BPHOT_NOALLOC_SCRATCH_BYTES that every GPU loading libbphot.so holds until the process exits.
 *   limit   nstars <= 65536 and 256 + nstars * (row + index bytes) <= BPHOT_NOALLOC_SCRATCH_BYTES; for fp32 pixels and
 *           skyrad=[40,50] that is 18 240 B per star = 3 679 stars per call (setsky: no rows, 65 536 stars).  Beyond it the
 *           call returns BPHOT_ERR_CAPACITY and bphot_last_error() names the largest star count that fits.
 *   sharing every call on a device uses the same array, which keeps the last call's contents (the call clears the counters
 *           it needs at zero itself): two calls that OVERLAP on the GPU (different streams, or concurrent host threads) race.
 *   order   stars are processed in caller order on the caller's stream (no visiting-order sort, no side stream), so results
 *           are those of bphot_aper_device; large star lists are faster through bphot_aper_device.
 */
#define BPHOT_NOALLOC_SCRATCH_BYTES (64u << 20)
int bphot_aper_device_noalloc(const void* d_image, int image_dtype, int nrow, int ncol, long long pitch,
                              const double* d_xc, const double* d_yc, int nstars,
                              double phpadu, const double* apr, int naper, double zeropoint,
                              double skyrad_in, double skyrad_out, double badlo, double badhi,
                              int use_setsky, const double* setsky, int minsky, int exact,
                              double readnoise, int flags,
                              double* d_mag, double* d_magerr, double* d_flux, double* d_fluxerr,
                              double* d_sky, double* d_skyerr, unsigned char* d_badflag,
                              int* d_nsky, int* d_mmm_iter, void* stream);

int bphot_aper_host(const void* image, int image_dtype, int nrow, int ncol,
                    const double* xc, const double* yc, int nstars,
                    double phpadu, const double* apr, int naper, double zeropoint,
                    double skyrad_in, double skyrad_out, double badlo, double badhi,
                    int use_setsky, const double* setsky, int minsky, int exact,
                    double readnoise, int flags,
                    double* mag, double* magerr, double* flux, double* fluxerr,
                    double* sky, double* skyerr, unsigned char* badflag,
                    int* nsky, int* mmm_iter, int device);

/*
 * pkfit_class(image, gauss, psf, ronois, phpadu[, noise, mask]).{pkfit, pkfit_noise,
 * pkfit_norecent_noise}(scale, x, y, sky, radius, maxiter)  for a list of stars.
 *   PythonPhot/pkfit.py:119-458, pkfit_noise.py:118-462, pkfit_norecent_noise.py:119-531,
 *   with dao_value.dao_value (dao_value.py:9), daoerf.daoerf (daoerf.py:9), rinter.rinter (rinter.py:8).
 *   variant           BPHOT_PKFIT / BPHOT_PKFIT_NOISE / BPHOT_PKFIT_NORECENT
 *   noise, mask       ignored (may be NULL) for BPHOT_PKFIT; mask is uint8, nonzero = masked
 *   gauss[5]          (height, x0, y0, sigx, sigy) (host)
 *   psf[npsf*npsf]    residual look-up table, fp64
 *   recenter          pkfit.py's `recenter` keyword (BPHOT_PKFIT only)
 * Outputs [nstars]: errmag, chi, sharp, scale, x, y, chi2 (fp64), niter (int32).  x, y are the
 * fitted (or unchanged) positions (`xyout=True`); chi2 is pkfit_norecent_noise's `returnchi2`
 * value, NaN for the other variants.
 */
int bphot_pkfit_device(int variant, const void* d_image, int image_dtype,
                       const void* d_noise, int noise_dtype, const unsigned char* d_mask,
                       int nrow, int ncol, long long pitch, long long noise_pitch, long long mask_pitch,
                       const double* gauss, const double* d_psf, int npsf,
                       double ronois, double phpadu,
                       const double* d_scale, const double* d_x, const double* d_y, const double* d_sky,
                       double radius, int maxiter, int recenter, int nstars,
                       double* d_errmag, double* d_chi, double* d_sharp, int* d_niter,
                       double* d_scale_out, double* d_x_out, double* d_y_out, double* d_chi2,
                       void* stream);

int bphot_pkfit_host(int variant, const void* image, int image_dtype,
                     const void* noise, int noise_dtype, const unsigned char* mask,
                     int nrow, int ncol,
                     const double* gauss, const double* psf, int npsf,
                     double ronois, double phpadu,
                     const double* scale, const double* x, const double* y, const double* sky,
                     double radius, int maxiter, int recenter, int nstars,
                     double* errmag, double* chi, double* sharp, int* niter,
                     double* scale_out, double* x_out, double* y_out, double* chi2,
                     int device);

/*
 * mmm.mmm(sky_vector, readnoise, minsky, mxiter) for nvec vectors packed back to back
 * (PythonPhot/mmm.py:7-314).  offs[nvec+1] = element offsets; max_len = longest vector.
 * out3[nvec*3] = (skymod, sigma, skew) per vector; iters[nvec] = clipping iterations.
 * Vectors longer than one CTA sorts (32 768 fp32 / 16 384 fp64 elements) are sorted in global memory (the call then reads
 * offs[nvec] back and synchronises the stream once).
 */
int bphot_mmm_device(const void* d_vec, int dtype, const long long* d_offs, int nvec, int max_len,
                     double readnoise, int minsky, int mxiter,
                     double* d_out3, int* d_iters, void* stream);

/* pixwt.Pixwt(xc, yc, r, x[n], y[n]) -> out[n]            PythonPhot/pixwt.py:8-48 */
int bphot_pixwt_device(double xc, double yc, double r, const double* d_x, const double* d_y,
                       long long n, double* d_out, void* stream);

/*
 * Per-point PSF model pieces (dao_value.py:95-107, daoerf.py:49-74, rinter.py:98-195).
 * out[12*n], structure-of-arrays rows: value, dvdx, dvdy, daoerf f, pder0..pder4, rinter z, dfdx, dfdy.
 * Points whose table coordinates fall outside [1, npsf-2] get zeros in the table-dependent rows;
 * dao_value's all-or-nothing guard (dao_value.py:70-84) is the caller's to apply.
 * table_coords != 0: xx, yy are already look-up-table coordinates (rinter.rinter's own x, y); only the
 * rinter rows (z, dfdx, dfdy) are meaningful then, and they follow rinter.py's flat-index arithmetic for ANY
 * point: flat indices past the end read the last table element (rinter.py:117-135); negative ones read NaN
 * when table_coords == 1, or wrap numpy-style against an array of length table_coords when it is > 1
 * (the table, or the padded copy the reference builds: max flat index + 3).
 */
int bphot_psfmodel_device(const double* d_xx, const double* d_yy, long long n,
                          const double* d_gauss, const double* d_psf, int npsf, int table_coords,
                          double* d_out, void* stream);

/*
 * find.find(image, hmin, fwhm, roundlim, sharplim)            PythonPhot/find.py:8-394 (DAOFIND)
 * Convolution with the lowered truncated Gaussian (find.py:106-188), threshold + local-maximum
 * search (find.py:211-233), per-candidate sharpness / roundness / marginal-Gaussian centroid
 * (find.py:259-346).  Stars come out in the reference's order (row-major order of their peak pixel).
 *   max_out           capacity of the five output columns (fp64, device / host)
 *   counts[5]         HOST int array, always filled: [0] stars meeting every criterion (the number of
 *                     valid output rows), [1] local maxima above hmin, [2..4] rejected by sharpness,
 *                     roundness, centroid (the figures find.py prints when verbose)
 * The call synchronises `stream` (the counts decide the output size).  If counts[0] > max_out nothing
 * is written and BPHOT_ERR_CAPACITY is returned; call again with larger columns.  counts[0] == 0 is
 * the case in which the reference returns None.
 */
int bphot_find_device(const void* d_image, int image_dtype, int nrow, int ncol, long long pitch,
                      double hmin, double fwhm, double round_lo, double round_hi,
                      double sharp_lo, double sharp_hi, int max_out,
                      double* d_x, double* d_y, double* d_flux, double* d_sharp, double* d_round,
                      int* counts, void* stream);

int bphot_find_host(const void* image, int image_dtype, int nrow, int ncol,
                    double hmin, double fwhm, double round_lo, double round_hi,
                    double sharp_lo, double sharp_hi, int max_out,
                    double* x, double* y, double* flux, double* sharp, double* roundness,
                    int* counts, int device);

/*
 * cntrd.cntrd(img, x, y, fwhm, extendbox, keepcenter)         PythonPhot/cntrd.py:7-231
 * Derivative-search centroid around approximate positions x[npts], y[npts] (device, fp64).
 * extend_x / extend_y: the `extendbox` keyword (scalar -> both); keepcenter as the keyword.
 * Outputs: xcen, ycen (fp64; -1, -1 where the reference gives up) and status (int32):
 */
#define BPHOT_CNTRD_OK 0
#define BPHOT_CNTRD_EDGE 1             /* 'too near edge of image'            cntrd.py:113-119 */
#define BPHOT_CNTRD_MOVED_EDGE 2       /* 'moved too near edge of image'      cntrd.py:146-152 */
#define BPHOT_CNTRD_X_NOT_DECREASING 3 /* 'Unable to compute X centroid'      cntrd.py:181-186 */
#define BPHOT_CNTRD_X_OUT_OF_RANGE 4   /* 'Computed X centroid out of range'  cntrd.py:188-193 */
#define BPHOT_CNTRD_Y_NOT_DECREASING 5 /* cntrd.py:202-207 */
#define BPHOT_CNTRD_Y_OUT_OF_RANGE 6   /* cntrd.py:209-214 */
#define BPHOT_CNTRD_NO_FINITE_PIXEL 7  /* search box all NaN (the reference raises) */
int bphot_cntrd_device(const void* d_image, int image_dtype, int nrow, int ncol, long long pitch,
                       const double* d_x, const double* d_y, int npts, double fwhm,
                       int extend_x, int extend_y, int keepcenter,
                       double* d_xcen, double* d_ycen, int* d_status, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BPHOT_H */
