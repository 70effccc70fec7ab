// bphot_capi.cu -- the C ABI of libbphot.so (include/bphot.h): argument checks, launch
// configuration, and the host-buffer convenience entry points.  Unity build: the kernels are
// included so that one nvcc invocation produces the library.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <cub/device/device_radix_sort.cuh>

#include "../../include/bphot.h"
#include "aper_kernel.cu"
#include "pkfit_kernel.cu"

namespace {

thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define CU_TRY(expr)                                                                          \
    do {                                                                                      \
        cudaError_t e_ = (expr);                                                              \
        if (e_ != cudaSuccess) return fail(BPHOT_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
    } while (0)

int next_pow2(long long v) {
    int m = 2;
    while (m < v) m <<= 1;
    return m;
}

constexpr int APER_BATCH = 65536;  // stars per stage-1 / stage-2 launch pair (scratch = batch * ~13 KB)

struct DevInfo {
    int sms;
    int smem_optin;
};
int dev_info(DevInfo* d) {
    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    CU_TRY(cudaDeviceGetAttribute(&d->sms, cudaDevAttrMultiProcessorCount, dev));
    CU_TRY(cudaDeviceGetAttribute(&d->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    return 0;
}

// Persistent grid: as many CTAs as fit on the device at once, capped by the work.
template <typename K>
int persistent_grid(K kernel, int threads, size_t smem, long long work_items, int* grid) {
    DevInfo d;
    int rc = dev_info(&d);
    if (rc) return rc;
    if (smem > 48 * 1024) CU_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem));
    if (per_sm < 1) return fail(BPHOT_ERR_CAPACITY, "kernel does not fit on an SM with %zu B of shared memory", smem);
    long long g = (long long)per_sm * d.sms;
    if (g > work_items) g = work_items;
    if (g < 1) g = 1;
    *grid = (int)g;
    return 0;
}

size_t esize(int dtype) { return dtype == BPHOT_F32 ? 4 : 8; }

// Upper bound on the number of pixel centres in the annulus rin <= r <= rout: every pixel whose
// centre is in the annulus has its unit cell inside the annulus widened by sqrt(2)/2 on each side.
long long annulus_bound(double rin, double rout) {
    const double h = 0.70710678118654757;
    const double ro = rout + h, ri = rin > h ? rin - h : 0.0;
    const double area = M_PI * (ro * ro - ri * ri);
    const double box = (2. * rout + 3.) * (2. * rout + 3.);
    return (long long)((area < box ? area : box) + 8.);
}

// ---- visiting order: stars sorted by the 32x32-pixel image tile they sit in (tile rows outermost).
// Neighbouring CTAs then read overlapping stamps, so stamp reads hit L1/L2 instead of DRAM: a random
// star order costs ~27 KB of DRAM reads per star at skyrad=[40,50], tile order about 1 KB.
__global__ void tile_key_kernel(const double* x, const double* y, int n, int nrow, int ncol, unsigned* keys, int* idx) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double xv = x[i], yv = y[i];
        if (!(xv >= 0.)) xv = 0.;  // also NaN
        if (!(yv >= 0.)) yv = 0.;
        if (xv > ncol - 1) xv = ncol - 1;
        if (yv > nrow - 1) yv = nrow - 1;
        keys[i] = ((unsigned)((int)yv >> 5) << 16) | (unsigned)((int)xv >> 5);
        idx[i] = i;
    }
}
struct StarOrder {  // stream-ordered allocations, released by release()
    void* block = nullptr;
    const int* order = nullptr;
    void release(cudaStream_t st) { if (block) cudaFreeAsync(block, st); block = nullptr; }
};
constexpr int ORDER_MIN_STARS = 2048;  // below this the launch is latency-bound anyway
int make_star_order(const double* d_x, const double* d_y, int n, int nrow, int ncol, cudaStream_t st, StarOrder* out) {
    if (n < ORDER_MIN_STARS || nrow > (65535 << 5) || ncol > (65535 << 5)) return 0;
    size_t tmp_bytes = 0;
    CU_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const unsigned*)nullptr, (unsigned*)nullptr, (const int*)nullptr,
                                           (int*)nullptr, n, 0, 32, st));
    const size_t col = ((size_t)n * 4 + 255) & ~(size_t)255;
    tmp_bytes = (tmp_bytes + 255) & ~(size_t)255;
    CU_TRY(cudaMallocAsync(&out->block, 4 * col + tmp_bytes, st));
    char* b = reinterpret_cast<char*>(out->block);
    unsigned* k_in = reinterpret_cast<unsigned*>(b), *k_out = reinterpret_cast<unsigned*>(b + col);
    int* i_in = reinterpret_cast<int*>(b + 2 * col), *i_out = reinterpret_cast<int*>(b + 3 * col);
    int blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    tile_key_kernel<<<blocks, 256, 0, st>>>(d_x, d_y, n, nrow, ncol, k_in, i_in);
    CU_TRY(cub::DeviceRadixSort::SortPairs(b + 4 * col, tmp_bytes, k_in, k_out, i_in, i_out, n, 0, 32, st));
    g_launches += 1;  // ours; the radix sort passes are CUB's
    out->order = i_out;
    return 0;
}

struct DevBuf {  // frees on scope exit
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 1); }
    template <typename T> T* as() { return reinterpret_cast<T*>(p); }
};

}  // namespace

extern "C" {

int bphot_version(void) { return BPHOT_VERSION; }
const char* bphot_last_error(void) { return g_err; }
long long bphot_launch_count(void) { return g_launches.load(); }

int bphot_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int bphot_aper_device(const void* d_image, int image_dtype, int nrow, int ncol, long long pitch,
                      const double* d_xc, const double* d_yc, int nstars,
                      double phpadu, const double* apr, int naper, double zeropoint,
                      double skyrad_in, double skyrad_out, double badlo, double badhi,
                      int use_setsky, const double* setsky, int minsky, int exact,
                      double readnoise, int compat_py2,
                      double* d_mag, double* d_magerr, double* d_flux, double* d_fluxerr,
                      double* d_sky, double* d_skyerr, unsigned char* d_badflag,
                      int* d_nsky, int* d_mmm_iter, void* stream) {
    using namespace bphot;
    if (!d_image || !apr || nstars < 0 || nrow <= 0 || ncol <= 0) return fail(BPHOT_ERR_ARG, "bphot_aper: null image/apr or bad shape");
    if (naper < 1 || naper > BPHOT_MAX_APER) return fail(BPHOT_ERR_ARG, "bphot_aper: naper=%d outside 1..%d", naper, BPHOT_MAX_APER);
    if (image_dtype != BPHOT_F32 && image_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_aper: image dtype %d", image_dtype);
    if (use_setsky && !setsky) return fail(BPHOT_ERR_ARG, "bphot_aper: use_setsky without setsky[3]");
    if (nstars == 0) return BPHOT_OK;
    if (!d_xc || !d_yc || !d_mag || !d_magerr || !d_flux || !d_fluxerr || !d_sky || !d_skyerr || !d_badflag || !d_nsky || !d_mmm_iter)
        return fail(BPHOT_ERR_ARG, "bphot_aper: null star list or output pointer");
    AperParams P;
    memset(&P, 0, sizeof(P));
    P.img = d_image; P.nrow = nrow; P.ncol = ncol; P.pitch = pitch > 0 ? pitch : ncol;
    P.xc = d_xc; P.yc = d_yc; P.nstars = nstars;
    P.phpadu = phpadu; P.zeropoint = zeropoint; P.naper = naper;
    double amax = apr[0];
    for (int k = 0; k < naper; k++) { P.apr[k] = apr[k]; if (apr[k] > amax) amax = apr[k]; }
    P.use_setsky = use_setsky ? 1 : 0;
    if (use_setsky) {  // aper.py:156-160
        skyrad_in = 0.; skyrad_out = amax + 1.;
        P.setsky[0] = setsky[0]; P.setsky[1] = setsky[1]; P.setsky[2] = setsky[2];
    }
    if (!(skyrad_out >= 0.) || !(skyrad_in >= 0.) || !(skyrad_out < 1e4)) return fail(BPHOT_ERR_ARG, "bphot_aper: skyrad out of range");
    P.skyrad_in = skyrad_in; P.skyrad_out = skyrad_out; P.badlo = badlo; P.badhi = badhi;
    P.minsky = minsky; P.exact = exact ? 1 : 0; P.compat_py2 = compat_py2 ? 1 : 0; P.mxiter = 50;
    P.readnoise = readnoise;
    // register-sort variants hold 128*K keys; larger annuli fall back to the shared-memory sort (K = 0)
    const long long bound = use_setsky ? 4 : annulus_bound(skyrad_in, skyrad_out);
    const int need = next_pow2(bound);
    const int K = need <= 1024 ? 8 : need <= 2048 ? 16 : need <= 4096 ? 32 : 0;
    P.sort_cap = K ? AP_THREADS * K : need;
    P.mag = d_mag; P.magerr = d_magerr; P.flux = d_flux; P.fluxerr = d_fluxerr;
    P.sky = d_sky; P.skyerr = d_skyerr; P.badflag = d_badflag; P.nsky = d_nsky; P.mmm_iter = d_mmm_iter;
    P.span_rows = (int)(2. * skyrad_out) + 4;
    const size_t es = esize(image_dtype);
    const size_t key_bytes = ((size_t)(K ? AP_THREADS * (K + 1) : need) * es + 15) & ~(size_t)15;
    const size_t smem = key_bytes + (size_t)P.span_rows * 16;
    DevInfo d;
    int rc = dev_info(&d);
    if (rc) return rc;
    if (smem + 16384 > (size_t)d.smem_optin)
        return fail(BPHOT_ERR_CAPACITY, "bphot_aper: sky annulus [%g,%g] needs %zu B of shared memory per star (limit %d)",
                    skyrad_in, skyrad_out, smem, d.smem_optin);
    // scratch rows of sorted annulus keys (stage 1 -> stage 2), one batch of stars at a time
    P.sstride = (bound + 7) / 8 * 8;
    const int batch = nstars < APER_BATCH ? nstars : APER_BATCH;
    cudaStream_t st = (cudaStream_t)stream;
    void* scratch = nullptr;
    if (!use_setsky) CU_TRY(cudaMallocAsync(&scratch, (size_t)batch * P.sstride * es, st));
    P.skeys = scratch;
    StarOrder ord;
    rc = make_star_order(d_xc, d_yc, nstars, nrow, ncol, st, &ord);
    if (rc) return rc;
    P.order = ord.order;
    int grid1 = 0, grid2 = 0;
#define BP_STAGE1(T, KK)                                                                          \
    do {                                                                                          \
        rc = persistent_grid(aper_stage1_kernel<T, KK>, AP_THREADS, smem, P.scount, &grid1);      \
        if (!rc) aper_stage1_kernel<T, KK><<<grid1, AP_THREADS, smem, st>>>(P);                   \
    } while (0)
#define BP_STAGE1_K(T)                       \
    do {                                     \
        if (K == 8) BP_STAGE1(T, 8);         \
        else if (K == 16) BP_STAGE1(T, 16);  \
        else if (K == 32) BP_STAGE1(T, 32);  \
        else BP_STAGE1(T, 0);                \
    } while (0)
#define BP_STAGE2(T)                                                                                              \
    do {                                                                                                          \
        rc = persistent_grid(aper_stage2_kernel<T>, AP_THREADS, 0, (P.scount + AP_WARPS - 1) / AP_WARPS, &grid2); \
        if (!rc) aper_stage2_kernel<T><<<grid2, AP_THREADS, 0, st>>>(P);                                          \
    } while (0)
    for (int s0 = 0; s0 < nstars && !rc; s0 += batch) {
        P.s0 = s0;
        P.scount = (nstars - s0) < batch ? (nstars - s0) : batch;
        if (image_dtype == BPHOT_F32) BP_STAGE1_K(float);
        else BP_STAGE1_K(double);
        if (rc) break;
        if (image_dtype == BPHOT_F32) BP_STAGE2(float);
        else BP_STAGE2(double);
        g_launches += 2;
    }
#undef BP_STAGE2
#undef BP_STAGE1_K
#undef BP_STAGE1
    if (scratch) cudaFreeAsync(scratch, st);
    ord.release(st);
    if (rc) return rc;
    CU_TRY(cudaGetLastError());
    return BPHOT_OK;
}

int bphot_pkfit_device(int variant, const void* d_image, int image_dtype,
                       const void* d_noise, int noise_dtype, const unsigned char* d_mask,
                       int nrow, int ncol, long long pitch, long long noise_pitch, long long mask_pitch,
                       const double* gauss, const double* d_psf, int npsf,
                       double ronois, double phpadu,
                       const double* d_scale, const double* d_x, const double* d_y, const double* d_sky,
                       double radius, int maxiter, int recenter, int nstars,
                       double* d_errmag, double* d_chi, double* d_sharp, int* d_niter,
                       double* d_scale_out, double* d_x_out, double* d_y_out, double* d_chi2,
                       void* stream) {
    using namespace bphot;
    if (variant < 0 || variant > 2) return fail(BPHOT_ERR_ARG, "bphot_pkfit: variant %d", variant);
    if (!d_image || !gauss || !d_psf || npsf < 4 || nrow <= 0 || ncol <= 0 || nstars < 0) return fail(BPHOT_ERR_ARG, "bphot_pkfit: null image/gauss/psf or bad shape");
    if (variant != BPHOT_PKFIT && (!d_noise || !d_mask)) return fail(BPHOT_ERR_ARG, "bphot_pkfit: noise and mask images are required for variant %d", variant);
    if (image_dtype != BPHOT_F32 && image_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_pkfit: image dtype %d", image_dtype);
    if (variant != BPHOT_PKFIT && noise_dtype != BPHOT_F32 && noise_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_pkfit: noise dtype %d", noise_dtype);
    if (!(radius > 0.) || !(radius < 64.)) return fail(BPHOT_ERR_ARG, "bphot_pkfit: radius %g outside (0, 64)", radius);
    if (nstars == 0) return BPHOT_OK;
    if (!d_scale || !d_x || !d_y || !d_sky || !d_errmag || !d_chi || !d_sharp || !d_niter || !d_scale_out || !d_x_out || !d_y_out || !d_chi2)
        return fail(BPHOT_ERR_ARG, "bphot_pkfit: null star list or output pointer");
    PkfitParams P;
    memset(&P, 0, sizeof(P));
    P.variant = variant; P.img = d_image; P.noise = d_noise; P.mask = d_mask;
    P.nrow = nrow; P.ncol = ncol;
    P.pitch = pitch > 0 ? pitch : ncol; P.npitch = noise_pitch > 0 ? noise_pitch : ncol; P.mpitch = mask_pitch > 0 ? mask_pitch : ncol;
    for (int i = 0; i < 5; i++) P.gauss[i] = gauss[i];
    P.psf = d_psf; P.npsf = npsf; P.ronois = ronois; P.phpadu = phpadu; P.radius = radius;
    P.maxiter = maxiter; P.recenter = recenter ? 1 : 0;
    P.scale = d_scale; P.x = d_x; P.y = d_y; P.sky = d_sky; P.nstars = nstars;
    {   // pixel centres with (dx^2+dy^2) < radius^2: unit cells lie inside a disc of radius + sqrt(2)/2
        const double rr = radius + 0.70710678118654757;
        long long c = (long long)(M_PI * rr * rr + 8.);
        P.cap = (int)((c + 31) / 32 * 32);
    }
    P.errmag = d_errmag; P.chi = d_chi; P.sharp = d_sharp; P.niter = d_niter;
    P.scale_out = d_scale_out; P.x_out = d_x_out; P.y_out = d_y_out; P.chi2 = d_chi2;
    P.acap = 2 * (int)ceil(radius) + 4;  // int(x+r)+1 - int(x-r) + 1 <= 2r + 3
    const int narr = (variant == BPHOT_PKFIT_NORECENT) ? PK_NARR_FORCED : PK_NARR_FIT;
    const size_t smem = (size_t)npsf * npsf * 8 + (size_t)PK_WARPS * ((size_t)P.cap * (narr * 8 + 4) + (size_t)P.acap * 32);
    DevInfo d;
    int rc = dev_info(&d);
    if (rc) return rc;
    if (smem + 1024 > (size_t)d.smem_optin)
        return fail(BPHOT_ERR_CAPACITY, "bphot_pkfit: radius %g / npsf %d need %zu B of shared memory (limit %d)", radius, npsf, smem, d.smem_optin);
    const long long nblocks = ((long long)nstars + PK_WARPS - 1) / PK_WARPS;
    int grid = 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int nd = (variant == BPHOT_PKFIT) ? image_dtype : noise_dtype;
    StarOrder ord;
    rc = make_star_order(d_x, d_y, nstars, nrow, ncol, st, &ord);
    if (rc) return rc;
    P.order = ord.order;
#define BP_LAUNCH(V, TI, TN)                                                              \
    do {                                                                                  \
        rc = persistent_grid(pkfit_kernel<V, TI, TN>, PK_THREADS, smem, nblocks, &grid);  \
        if (rc) { ord.release(st); return rc; }                                           \
        pkfit_kernel<V, TI, TN><<<grid, PK_THREADS, smem, st>>>(P);                       \
    } while (0)
#define BP_LAUNCH_T(V)                                                           \
    do {                                                                         \
        if (image_dtype == BPHOT_F32 && nd == BPHOT_F32) BP_LAUNCH(V, float, float);   \
        else if (image_dtype == BPHOT_F32) BP_LAUNCH(V, float, double);          \
        else if (nd == BPHOT_F32) BP_LAUNCH(V, double, float);                   \
        else BP_LAUNCH(V, double, double);                                       \
    } while (0)
    if (variant == BPHOT_PKFIT) {  // no noise image: TN is unused
        if (image_dtype == BPHOT_F32) BP_LAUNCH(0, float, float);
        else BP_LAUNCH(0, double, double);
    } else if (variant == BPHOT_PKFIT_NOISE) BP_LAUNCH_T(1);
    else BP_LAUNCH_T(2);
#undef BP_LAUNCH_T
#undef BP_LAUNCH
    ord.release(st);
    g_launches++;
    CU_TRY(cudaGetLastError());
    return BPHOT_OK;
}

int bphot_mmm_device(const void* d_vec, int dtype, const long long* d_offs, int nvec, int max_len,
                     double readnoise, int minsky, int mxiter, double* d_out3, int* d_iters, void* stream) {
    using namespace bphot;
    if (!d_vec || !d_offs || !d_out3 || !d_iters || nvec < 0 || max_len < 0) return fail(BPHOT_ERR_ARG, "bphot_mmm: null pointer or negative size");
    if (dtype != BPHOT_F32 && dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_mmm: dtype %d", dtype);
    if (nvec == 0) return BPHOT_OK;
    const int need = next_pow2(max_len);
    const int K = need <= 1024 ? 8 : need <= 2048 ? 16 : need <= 4096 ? 32 : 0;
    const int cap = K ? AP_THREADS * K : need;
    const size_t smem = (size_t)(K ? AP_THREADS * (K + 1) : need) * esize(dtype);
    DevInfo d;
    int rc = dev_info(&d);
    if (rc) return rc;
    if (smem + 1024 > (size_t)d.smem_optin) return fail(BPHOT_ERR_CAPACITY, "bphot_mmm: vector of %d elements exceeds shared memory", max_len);
    int grid = 0;
    cudaStream_t st = (cudaStream_t)stream;
#define BP_LAUNCH_MMM(T, KK)                                                                                         \
    do {                                                                                                             \
        rc = persistent_grid(mmm_kernel<T, KK>, AP_THREADS, smem, nvec, &grid);                                      \
        if (rc) return rc;                                                                                           \
        mmm_kernel<T, KK><<<grid, AP_THREADS, smem, st>>>((const T*)d_vec, d_offs, nvec, readnoise, minsky, mxiter,  \
                                                          cap, d_out3, d_iters);                                     \
    } while (0)
#define BP_LAUNCH_MMM_K(T)                      \
    do {                                        \
        if (K == 8) BP_LAUNCH_MMM(T, 8);        \
        else if (K == 16) BP_LAUNCH_MMM(T, 16); \
        else if (K == 32) BP_LAUNCH_MMM(T, 32); \
        else BP_LAUNCH_MMM(T, 0);               \
    } while (0)
    if (dtype == BPHOT_F32) BP_LAUNCH_MMM_K(float);
    else BP_LAUNCH_MMM_K(double);
#undef BP_LAUNCH_MMM_K
#undef BP_LAUNCH_MMM
    g_launches++;
    CU_TRY(cudaGetLastError());
    return BPHOT_OK;
}

int bphot_pixwt_device(double xc, double yc, double r, const double* d_x, const double* d_y,
                       long long n, double* d_out, void* stream) {
    if (n < 0 || (n > 0 && (!d_x || !d_y || !d_out))) return fail(BPHOT_ERR_ARG, "bphot_pixwt: null pointer");
    if (n == 0) return BPHOT_OK;
    long long blocks = (n + 127) / 128;
    if (blocks > 148 * 16) blocks = 148 * 16;
    bphot::pixwt_kernel<<<(int)blocks, 128, 0, (cudaStream_t)stream>>>(xc, yc, r, d_x, d_y, n, d_out);
    g_launches++;
    CU_TRY(cudaGetLastError());
    return BPHOT_OK;
}

int bphot_psfmodel_device(const double* d_xx, const double* d_yy, long long n,
                          const double* d_gauss, const double* d_psf, int npsf, int table_coords,
                          double* d_out, void* stream) {
    if (n < 0 || npsf < 4 || (n > 0 && (!d_xx || !d_yy || !d_gauss || !d_psf || !d_out))) return fail(BPHOT_ERR_ARG, "bphot_psfmodel: null pointer");
    if (n == 0) return BPHOT_OK;
    long long blocks = (n + 127) / 128;
    if (blocks > 148 * 16) blocks = 148 * 16;
    bphot::psfmodel_kernel<<<(int)blocks, 128, 0, (cudaStream_t)stream>>>(d_xx, d_yy, n, d_gauss, d_psf, npsf, table_coords, d_out);
    g_launches++;
    CU_TRY(cudaGetLastError());
    return BPHOT_OK;
}

// ------------------------------------------------------------ host flavours
#define UP(buf, src, bytes)                                                     \
    do {                                                                        \
        CU_TRY(buf.alloc(bytes));                                               \
        CU_TRY(cudaMemcpyAsync(buf.p, src, bytes, cudaMemcpyHostToDevice, 0));  \
    } while (0)
#define DOWN(dst, buf, bytes) CU_TRY(cudaMemcpyAsync(dst, buf.p, bytes, cudaMemcpyDeviceToHost, 0))

int bphot_aper_host(const void* image, int image_dtype, int nrow, int ncol,
                    const double* xc, const double* yc, int nstars,
                    double phpadu, const double* apr, int naper, double zeropoint,
                    double skyrad_in, double skyrad_out, double badlo, double badhi,
                    int use_setsky, const double* setsky, int minsky, int exact,
                    double readnoise, int compat_py2,
                    double* mag, double* magerr, double* flux, double* fluxerr,
                    double* sky, double* skyerr, unsigned char* badflag,
                    int* nsky, int* mmm_iter, int device) {
    if (bphot_device_count() < 1) return fail(BPHOT_ERR_NODEVICE, "no CUDA device (libbphot has no CPU fallback)");
    if (!image || !apr || nstars < 0 || nrow <= 0 || ncol <= 0 || naper < 1 || naper > BPHOT_MAX_APER) return fail(BPHOT_ERR_ARG, "bphot_aper_host: bad argument");
    if (image_dtype != BPHOT_F32 && image_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_aper_host: image dtype %d", image_dtype);
    if (nstars == 0) return BPHOT_OK;
    CU_TRY(cudaSetDevice(device));
    const size_t nb = (size_t)nstars * sizeof(double), na = nb * naper;
    DevBuf dimg, dx, dy, dmag, dmagerr, dflux, dfluxerr, dsky, dskyerr, dbad, dnsky, dit;
    UP(dimg, image, (size_t)nrow * ncol * esize(image_dtype));
    UP(dx, xc, nb);
    UP(dy, yc, nb);
    CU_TRY(dmag.alloc(na)); CU_TRY(dmagerr.alloc(na)); CU_TRY(dflux.alloc(na)); CU_TRY(dfluxerr.alloc(na));
    CU_TRY(dsky.alloc(nb)); CU_TRY(dskyerr.alloc(nb)); CU_TRY(dbad.alloc((size_t)nstars * naper));
    CU_TRY(dnsky.alloc((size_t)nstars * 4)); CU_TRY(dit.alloc((size_t)nstars * 4));
    int rc = bphot_aper_device(dimg.p, image_dtype, nrow, ncol, ncol, dx.as<double>(), dy.as<double>(), nstars, phpadu, apr, naper,
                               zeropoint, skyrad_in, skyrad_out, badlo, badhi, use_setsky, setsky, minsky, exact, readnoise,
                               compat_py2, dmag.as<double>(), dmagerr.as<double>(), dflux.as<double>(), dfluxerr.as<double>(),
                               dsky.as<double>(), dskyerr.as<double>(), dbad.as<unsigned char>(), dnsky.as<int>(), dit.as<int>(), 0);
    if (rc) return rc;
    DOWN(mag, dmag, na); DOWN(magerr, dmagerr, na); DOWN(flux, dflux, na); DOWN(fluxerr, dfluxerr, na);
    DOWN(sky, dsky, nb); DOWN(skyerr, dskyerr, nb); DOWN(badflag, dbad, (size_t)nstars * naper);
    if (nsky) DOWN(nsky, dnsky, (size_t)nstars * 4);
    if (mmm_iter) DOWN(mmm_iter, dit, (size_t)nstars * 4);
    CU_TRY(cudaStreamSynchronize(0));
    return BPHOT_OK;
}

int bphot_pkfit_host(int variant, const void* image, int image_dtype,
                     const void* noise, int noise_dtype, const unsigned char* mask,
                     int nrow, int ncol,
                     const double* gauss, const double* psf, int npsf,
                     double ronois, double phpadu,
                     const double* scale, const double* x, const double* y, const double* sky,
                     double radius, int maxiter, int recenter, int nstars,
                     double* errmag, double* chi, double* sharp, int* niter,
                     double* scale_out, double* x_out, double* y_out, double* chi2,
                     int device) {
    if (bphot_device_count() < 1) return fail(BPHOT_ERR_NODEVICE, "no CUDA device (libbphot has no CPU fallback)");
    if (!image || !gauss || !psf || nstars < 0 || nrow <= 0 || ncol <= 0 || npsf < 4) return fail(BPHOT_ERR_ARG, "bphot_pkfit_host: bad argument");
    if (variant != BPHOT_PKFIT && (!noise || !mask)) return fail(BPHOT_ERR_ARG, "bphot_pkfit_host: noise and mask images are required for variant %d", variant);
    if (image_dtype != BPHOT_F32 && image_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_pkfit_host: image dtype %d", image_dtype);
    if (nstars == 0) return BPHOT_OK;
    CU_TRY(cudaSetDevice(device));
    const size_t nb = (size_t)nstars * sizeof(double), npx = (size_t)nrow * ncol;
    DevBuf dimg, dnz, dmk, dpsf, ds, dx, dy, dsk, o_err, o_chi, o_sh, o_it, o_sc, o_x, o_y, o_c2;
    UP(dimg, image, npx * esize(image_dtype));
    if (variant != BPHOT_PKFIT) {
        if (noise_dtype != BPHOT_F32 && noise_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_pkfit_host: noise dtype %d", noise_dtype);
        UP(dnz, noise, npx * esize(noise_dtype));
        UP(dmk, mask, npx);
    }
    UP(dpsf, psf, (size_t)npsf * npsf * 8);
    UP(ds, scale, nb); UP(dx, x, nb); UP(dy, y, nb); UP(dsk, sky, nb);
    CU_TRY(o_err.alloc(nb)); CU_TRY(o_chi.alloc(nb)); CU_TRY(o_sh.alloc(nb)); CU_TRY(o_it.alloc((size_t)nstars * 4));
    CU_TRY(o_sc.alloc(nb)); CU_TRY(o_x.alloc(nb)); CU_TRY(o_y.alloc(nb)); CU_TRY(o_c2.alloc(nb));
    int rc = bphot_pkfit_device(variant, dimg.p, image_dtype, dnz.p, noise_dtype, dmk.as<unsigned char>(), nrow, ncol, ncol, ncol, ncol,
                                gauss, dpsf.as<double>(), npsf, ronois, phpadu, ds.as<double>(), dx.as<double>(), dy.as<double>(),
                                dsk.as<double>(), radius, maxiter, recenter, nstars, o_err.as<double>(), o_chi.as<double>(),
                                o_sh.as<double>(), o_it.as<int>(), o_sc.as<double>(), o_x.as<double>(), o_y.as<double>(),
                                o_c2.as<double>(), 0);
    if (rc) return rc;
    DOWN(errmag, o_err, nb); DOWN(chi, o_chi, nb); DOWN(sharp, o_sh, nb); DOWN(niter, o_it, (size_t)nstars * 4);
    DOWN(scale_out, o_sc, nb);
    if (x_out) DOWN(x_out, o_x, nb);
    if (y_out) DOWN(y_out, o_y, nb);
    if (chi2) DOWN(chi2, o_c2, nb);
    CU_TRY(cudaStreamSynchronize(0));
    return BPHOT_OK;
}

}  // extern "C"
