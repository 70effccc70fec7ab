// bphot_capi.cu -- the C ABI of libbphot.so (include/bphot.h): argument checks, launch
// configuration, and the host-buffer convenience entry points.  Unity build: the kernels are
// included so that one nvcc invocation produces the library.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <vector>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include <cub/device/device_segmented_sort.cuh>
#include <thrust/iterator/transform_iterator.h>
#include <thrust/iterator/counting_iterator.h>

#include "../../include/bphot.h"
#include "aper_kernel.cu"
#include "pkfit_kernel.cu"
#include "find_kernel.cu"

// Scratch of bphot_aper_device_noalloc, the one entry point that allocates nothing: BPHOT_NOALLOC_SCRATCH_BYTES of device memory
// that every GPU loading this library holds until exit.  In host code this NAME is a host-side stand-in, not a device address:
// noalloc_scratch() below asks the runtime for the address, once per device.  All calls share it (overlapping calls race).
static __device__ __align__(256) unsigned char g_noalloc_scratch[BPHOT_NOALLOC_SCRATCH_BYTES];

namespace {

thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

// Optional per-kernel timing (bphot_profile / bphot_kernel_times): CUDA events recorded on the launch stream around
// each launch of the four hot kernels; the call that owns them synchronises the stream before returning.
enum { KT_APER_STAGE1 = 0, KT_APER_SUMS = 1, KT_APER_STAGE2 = 2, KT_PKFIT = 3, KT_APER_FINAL = 4, KT_N = 5 };
std::atomic<int> g_profile{0};
double g_kt_ms[KT_N] = {0, 0, 0, 0, 0};
long long g_kt_n[KT_N] = {0, 0, 0, 0, 0};
struct LaunchTimer {
    struct Rec { cudaEvent_t a, b; int cls; };
    Rec recs[64];
    int n = 0;
    cudaStream_t st;
    bool on;
    explicit LaunchTimer(cudaStream_t s) : st(s), on(g_profile.load() != 0) {}
    void begin(int cls) {
        if (!on) return;
        if (n == 64) flush();
        Rec& r = recs[n];
        r.cls = cls;
        cudaEventCreate(&r.a); cudaEventCreate(&r.b);
        cudaEventRecord(r.a, st);
    }
    void end() {
        if (!on) return;
        cudaEventRecord(recs[n].b, st);
        n++;
    }
    void flush() {
        if (!on || n == 0) return;
        cudaStreamSynchronize(st);
        for (int i = 0; i < n; i++) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, recs[i].a, recs[i].b) == cudaSuccess) { g_kt_ms[recs[i].cls] += ms; g_kt_n[recs[i].cls] += 1; }
            cudaEventDestroy(recs[i].a); cudaEventDestroy(recs[i].b);
        }
        n = 0;
    }
    ~LaunchTimer() { flush(); }
};

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

#ifndef BP_POOL_KEEP_GIB
#define BP_POOL_KEEP_GIB 8  // freed stream-ordered scratch the device pool keeps across synchronisations
#endif
#ifndef AP_BATCH
#define AP_BATCH 65536
#endif
#ifndef AP_MMM_INDEXED
#define AP_MMM_INDEXED 1  // 1 = mmm by one thread per star on the indexed sorted row; 0 = one warp per star on the plain row
#endif
#ifndef AP_MMM_OVERLAP
#define AP_MMM_OVERLAP 0  // 1 = the thread-per-star mmm of batch b runs on a side stream under stage 1 of batch b+1 (two scratch buffers); measured: no gain
#endif
#ifndef AP_SIDE_STREAM
#define AP_SIDE_STREAM 0  // measured: co-running aper_sums with stage 1/2 does not pay; kept as a build knob
#endif
#ifndef AP_SUMS_SIDE_CTAS
#define AP_SUMS_SIDE_CTAS 2  // CTAs per SM of the aperture-sum kernel while it shares the SMs with stage 1 / stage 2 (0 = no cap)
#endif
#ifndef AP_S2_CTAS
#define AP_S2_CTAS 0  // 0 = as many stage-2 CTAs per SM as fit
#endif
#ifndef AP_GROUP
#define AP_GROUP 4  // stage-1 launches per stage-2 (mmm) launch: the thread-per-star mmm wants many stars in flight
#endif
constexpr int APER_BATCH = AP_BATCH;  // stars per stage-1 launch (also the granularity of the banded-upload waits)

struct DevInfo {
    int sms;
    int smem_optin;
};
int dev_info(DevInfo* d) {
    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    {   // keep up to 8 GiB of freed stream-ordered scratch (sorted-annulus rows, find's h image) in the device's default
        // pool across synchronisations: with the default threshold of 0 every call re-maps ~0.7 GB from the driver (~15 ms)
        static std::atomic<unsigned long long> tuned{0};
        const unsigned long long bit = 1ull << (dev & 63);
        if (!(tuned.load() & bit)) {
            cudaMemPool_t pool;
            if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
                unsigned long long keep = (unsigned long long)BP_POOL_KEEP_GIB << 30;
                cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
            }
            cudaGetLastError();
            tuned.fetch_or(bit);
        }
    }
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

// Upper bound on the number of pixel centres strictly inside a disc of radius r, for ANY position of its centre: the
// centre's offset inside its pixel is cut into 16 x 16 cells and a lattice point counts for a cell if it is within r of
// the nearest point of the cell.  Rigorous (every centre lies in some cell) and within ~2 of the true maximum
// (82 for r = 5, true 81; the area bound pi (r + sqrt 2 / 2)^2 + 8 gave 110), which is what sizes the fit kernels'
// per-warp shared-memory lists.
int disc_lattice_bound(double r) {
    const int R = (int)ceil(r) + 1, NC = 16;
    const double h = 1.0 / NC, rr = r * r * (1.0 + 1e-12);
    int best = 0;
    for (int a = 0; a < NC; a++)
        for (int b = 0; b < NC; b++) {
            const double x0 = a * h, x1 = x0 + h, y0 = b * h, y1 = y0 + h;
            int c = 0;
            for (int i = -R; i <= R + 1; i++)
                for (int j = -R; j <= R + 1; j++) {
                    const double dx = i < x0 ? x0 - i : (i > x1 ? i - x1 : 0.0), dy = j < y0 ? y0 - j : (j > y1 ? j - y1 : 0.0);
                    if (dx * dx + dy * dy < rr) c++;
                }
            if (c > best) best = c;
        }
    return best;
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
// device-side copy out of page-locked host memory (bphot_fetch_host): 16-byte loads over PCIe, grid-stride
__global__ void fetch_host_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src, long long n16, unsigned char* dst8,
                                  const unsigned char* src8, long long tail0, long long nbytes) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += (long long)gridDim.x * blockDim.x) dst[i] = src[i];
    for (long long i = tail0 + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nbytes; i += (long long)gridDim.x * blockDim.x) dst8[i] = src8[i];
}
// largest y of the stars of each batch of the visiting order (one CTA per batch): which image rows the batch needs
__global__ void batch_maxrow_kernel(const double* y, const int* order, int n, int batch, float* out) {
    __shared__ float s[256];
    const int b = blockIdx.x;
    float m = -1e30f;
    for (int i = b * batch + threadIdx.x; i < n && i < (b + 1) * batch; i += blockDim.x) {
        const double v = y[order[i]];
        if (v == v) m = fmaxf(m, (float)v + 1.0f);  // +1: fp32 rounding of the row
    }
    s[threadIdx.x] = m;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) s[threadIdx.x] = fmaxf(s[threadIdx.x], s[threadIdx.x + o]);
        __syncthreads();
    }
    if (threadIdx.x == 0) out[b] = s[0];
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


// ---- find.py:106-183 on the host: everything that depends on fwhm alone
double np_sum_host(const double* a, int n) {  // numpy's pairwise add-reduce (blocks of 128, 8 accumulators)
    if (n < 8) {
        double r = 0.0;
        for (int i = 0; i < n; i++) r += a[i];
        return r;
    }
    if (n <= 128) {
        double r[8];
        int i;
        for (i = 0; i < 8; i++) r[i] = a[i];
        for (i = 8; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; j++) r[j] += a[i + j];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; i++) res += a[i];
        return res;
    }
    int n2 = n / 2;
    n2 -= n2 % 8;
    return np_sum_host(a, n2) + np_sum_host(a + n2, n - n2);
}

void find_constants(double fwhm, bphot::FindConsts* K) {
    using bphot::FIND_MAXBOX;
    memset(K, 0, sizeof(*K));
    double radius = 0.637 * fwhm;  // find.py:106-107
    if (radius < 2.001) radius = 2.001;
    const double radsq = radius * radius;
    int nhalf = (int)radius;
    if (nhalf > (FIND_MAXBOX - 1) / 2) nhalf = (FIND_MAXBOX - 1) / 2;
    const int nbox = 2 * nhalf + 1, nn = nbox * nbox;
    const double sigsq = (fwhm / 2.35482) * (fwhm / 2.35482);
    K->nhalf = nhalf; K->nbox = nbox; K->sigsq = sigsq;
    double g[FIND_MAXBOX * FIND_MAXBOX], t[FIND_MAXBOX * FIND_MAXBOX], wt[FIND_MAXBOX], v[FIND_MAXBOX];
    unsigned char mask[FIND_MAXBOX * FIND_MAXBOX];
    int pixels = 0;
    for (int r = 0; r < nbox; r++)
        for (int c = 0; c < nbox; c++) {
            const double d2 = (double)((r - nhalf) * (r - nhalf) + (c - nhalf) * (c - nhalf));  // find.py:123-126
            mask[r * nbox + c] = d2 <= radsq;
            pixels += mask[r * nbox + c];
            g[r * nbox + c] = exp(-0.5 * d2 / sigsq);
        }
    for (int i = 0; i < nbox; i++) wt[i] = (double)(nhalf - abs(i - nhalf) + 1);  // find.py:150-152
    for (int i = 0; i < nbox; i++) K->wt[i] = wt[i];
    for (int r = 0; r < nbox; r++) {  // sgx = sum(g*xwt, 1)
        for (int c = 0; c < nbox; c++) v[c] = g[r * nbox + c] * wt[c];
        K->sgx[r] = np_sum_host(v, nbox);
    }
    K->p = np_sum_host(wt, nbox);
    for (int c = 0; c < nbox; c++) {  // sgy = sum(g*ywt, 0): rows are added in order
        double s = g[c] * wt[0];
        for (int r = 1; r < nbox; r++) s += g[r * nbox + c] * wt[r];
        K->sgy[c] = s;
    }
#define BP_VSUM(expr) ([&]() { for (int i = 0; i < nbox; i++) v[i] = (expr); return np_sum_host(v, nbox); }())
    K->sumgx = BP_VSUM(wt[i] * K->sgy[i]);
    K->sumgy = BP_VSUM(wt[i] * K->sgx[i]);
    K->sumgsqy = BP_VSUM(wt[i] * K->sgy[i] * K->sgy[i]);
    K->sumgsqx = BP_VSUM(wt[i] * K->sgx[i] * K->sgx[i]);
    for (int i = 0; i < nbox; i++) {
        const double vec = (double)(nhalf - i);
        K->dgdx[i] = K->sgy[i] * vec;
        K->dgdy[i] = K->sgx[i] * vec;
    }
    K->sdgdxs = BP_VSUM(wt[i] * (K->dgdx[i] * K->dgdx[i]));
    K->sdgdx = BP_VSUM(wt[i] * K->dgdx[i]);
    K->sdgdys = BP_VSUM(wt[i] * (K->dgdy[i] * K->dgdy[i]));
    K->sdgdy = BP_VSUM(wt[i] * K->dgdy[i]);
    K->sgdgdx = BP_VSUM(wt[i] * K->sgy[i] * K->dgdx[i]);
    K->sgdgdy = BP_VSUM(wt[i] * K->sgx[i] * K->dgdy[i]);
    // convolution kernel, find.py:174-179
    for (int i = 0; i < nn; i++) K->c[i] = g[i] * mask[i];
    double sumc = np_sum_host(K->c, nn);
    for (int i = 0; i < nn; i++) t[i] = K->c[i] * K->c[i];
    const double sumcsq = np_sum_host(t, nn) - sumc * sumc / pixels;
    sumc = sumc / pixels;
    for (int i = 0; i < nn; i++)
        if (mask[i]) K->c[i] = (K->c[i] - sumc) / sumcsq;
    // 1-D kernel of the roundness statistic, find.py:180-183
    for (int i = 0; i < nbox; i++) v[i] = exp(-.5 * (double)((i - nhalf) * (i - nhalf)) / sigsq);
    double c1[FIND_MAXBOX];
    for (int i = 0; i < nbox; i++) c1[i] = v[i];
    const double sumc1 = np_sum_host(c1, nbox) / nbox;
    for (int i = 0; i < nbox; i++) v[i] = c1[i] * c1[i];
    const double sumc1sq = np_sum_host(v, nbox) - sumc1;
    for (int i = 0; i < nbox; i++) K->c1[i] = (c1[i] - sumc1) / sumc1sq;
#undef BP_VSUM
    for (int i = 0; i < nn; i++) K->ring[i] = mask[i];
    K->ring[nhalf * nbox + nhalf] = 0;  // find.py:199-200
    K->pixels = pixels - 1;
    {   // footprint offsets, nearest first: the local-maximum test usually fails on a close neighbour
        int m = 0;
        for (int d2 = 1; d2 <= 2 * nhalf * nhalf; d2++)
            for (int r = 0; r < nbox; r++)
                for (int c = 0; c < nbox; c++)
                    if (K->ring[r * nbox + c] && (r - nhalf) * (r - nhalf) + (c - nhalf) * (c - nhalf) == d2) {
                        K->rdy[m] = (signed char)(r - nhalf);
                        K->rdx[m] = (signed char)(c - nhalf);
                        m++;
                    }
    }
}

// one non-blocking side stream per device, created on first use and kept for the life of the process
int side_stream(cudaStream_t* out) {
    static std::atomic<cudaStream_t> streams[64];
    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    cudaStream_t s = streams[dev & 63].load();
    if (!s) {
        cudaStream_t created = nullptr;
        CU_TRY(cudaStreamCreateWithFlags(&created, cudaStreamNonBlocking));
        cudaStream_t expected = nullptr;
        if (streams[dev & 63].compare_exchange_strong(expected, created)) s = created;
        else { cudaStreamDestroy(created); s = expected; }
    }
    *out = s;
    return 0;
}

// the log10 table of bphot_math.cuh is filled once per device, on the caller's stream before the first mmm launch
int ensure_log10_table(cudaStream_t st) {
    static std::atomic<unsigned long long> done{0};
    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    const unsigned long long bit = 1ull << (dev & 63);
    if (done.load() & bit) return 0;
    bphot::log10_table_kernel<<<8, 256, 0, st>>>();
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaStreamSynchronize(st));  // other streams may launch mmm right after this call returns
    done.fetch_or(bit);
    g_launches += 1;
    return 0;
}

// device address of g_noalloc_scratch on the current device (cudaGetSymbolAddress once per device, that device current)
constexpr size_t NOALLOC_HEAD = 256;  // the first 256 bytes are the work counters, zeroed by every call
int noalloc_scratch(unsigned char** out) {
    static std::atomic<unsigned char*> ptrs[64];
    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(BPHOT_ERR_ARG, "bphot_aper_noalloc: device ordinal %d outside 0..63", dev);
    unsigned char* p = ptrs[dev].load();
    if (!p) {
        void* sym = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&sym, g_noalloc_scratch);
        if (e != cudaSuccess || !sym) return fail(BPHOT_ERR_CUDA, "bphot_aper_noalloc: cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
        p = reinterpret_cast<unsigned char*>(sym);
        ptrs[dev].store(p);
    }
    *out = p;
    return 0;
}

// ---- sky annuli above one CTA's shared memory (aper_gather_global_kernel): rows of unsorted keys in global scratch, one
// segmented device sort per batch of stars (CUB, one segment per star: [row start, row start + slots in use)), then the
// warp-per-star mmm / sigma-clipping kernel on the sorted rows, the aperture sums and the finalize kernel, all on `st`.
struct SegBegin {
    long long stride;
    __host__ __device__ int operator()(int i) const { return (int)(i * stride); }
};
struct SegEnd {
    const int* count;
    long long stride;
    __host__ __device__ int operator()(int i) const { return (int)(i * stride) + count[i]; }
};
constexpr long long BIG_BATCH_KEYS = 1ll << 27;  // keys per batch of stars: 0.5 GiB (fp32) per buffer, three buffers

template <typename T>
int aper_big_annulus_run(bphot::AperParams P, int nstars, int batch, T* buf_in, T* buf_out, void* tmp, size_t tmp_bytes,
                         int* seg_count, int* counters, size_t smem, cudaStream_t st) {
    using namespace bphot;
    int* next_counter = counters;
    for (int s0 = 0; s0 < nstars; s0 += batch) {
        int grid = 0;
        P.s0 = s0;
        P.scount = (nstars - s0) < batch ? (nstars - s0) : batch;
        P.row0 = 0;
        P.skeys = buf_in;
        int rc = persistent_grid(aper_gather_global_kernel<T>, AP_BIG_THREADS, smem, P.scount, &grid);
        if (rc) return rc;
        P.work_counter = next_counter++;
        aper_gather_global_kernel<T><<<grid, AP_BIG_THREADS, smem, st>>>(P, seg_count);
        auto begins = thrust::make_transform_iterator(thrust::counting_iterator<int>(0), SegBegin{P.sstride});
        auto ends = thrust::make_transform_iterator(thrust::counting_iterator<int>(0), SegEnd{seg_count, P.sstride});
        size_t need = 0;
        CU_TRY(cub::DeviceSegmentedSort::SortKeys(nullptr, need, buf_in, buf_out, (int)((long long)P.scount * P.sstride), P.scount, begins, ends, st));
        if (need > tmp_bytes) return fail(BPHOT_ERR_CUDA, "bphot_aper: segmented sort wants %zu B of temporary storage, %zu B reserved", need, tmp_bytes);
        CU_TRY(cub::DeviceSegmentedSort::SortKeys(tmp, need, buf_in, buf_out, (int)((long long)P.scount * P.sstride), P.scount, begins, ends, st));
        P.skeys = buf_out;
        rc = persistent_grid(aper_stage2_kernel<T>, AP_THREADS, 0, (P.scount + AP_WARPS - 1) / AP_WARPS, &grid);
        if (rc) return rc;
        P.work_counter = next_counter++;
        aper_stage2_kernel<T><<<grid, AP_THREADS, 0, st>>>(P);
        g_launches += 2;  // ours; the sort passes are CUB's
    }
    int grid = 0;
    P.s0 = 0; P.scount = nstars;
    int rc = persistent_grid(aper_sums_kernel<T>, AP_THREADS, 0, ((long long)nstars + AP_WARPS - 1) / AP_WARPS, &grid);
    if (rc) return rc;
    P.work_counter = next_counter++;
    aper_sums_kernel<T><<<grid, AP_THREADS, 0, st>>>(P);
    int gf = (nstars + 255) / 256;
    if (gf > 148 * 8) gf = 148 * 8;
    aper_finalize_kernel<<<gf, 256, 0, st>>>(P);
    g_launches += 2;
    CU_TRY(cudaGetLastError());
    return 0;
}

// P: everything but the scratch fields filled in by bphot_aper_device_banded; `pixels` = upper bound on annulus pixels per star
int aper_big_annulus(bphot::AperParams P, int image_dtype, long long pixels, int nstars, int smem_optin, cudaStream_t st) {
    using namespace bphot;
    if (P.span_rows > P.nrow) P.span_rows = P.nrow;  // a stamp never has more rows than the image
    if (pixels > (long long)P.nrow * P.ncol) pixels = (long long)P.nrow * P.ncol;
    const size_t smem = (size_t)P.span_rows * 20;
    P.sstride = (pixels + 15) / 16 * 16;
    if (smem + 2048 > (size_t)smem_optin || P.sstride > BIG_BATCH_KEYS)
        return fail(BPHOT_ERR_CAPACITY, "bphot_aper: sky annulus [%g,%g] on a %d x %d image: %lld pixels per star / %zu B of span tables "
                    "(limits %lld pixels, %d B)", P.skyrad_in, P.skyrad_out, P.nrow, P.ncol, P.sstride, smem, BIG_BATCH_KEYS, smem_optin - 2048);
    P.sort_cap = (int)P.sstride;
    P.idx_stride = 0; P.sidx = nullptr; P.order = nullptr;
    long long batch = BIG_BATCH_KEYS / P.sstride;
    if (batch > nstars) batch = nstars;
    if (batch > APER_BATCH) batch = APER_BATCH;
    const int nbatch = (int)((nstars + batch - 1) / batch);
    const size_t es = esize(image_dtype);
    const size_t row_bytes = ((size_t)batch * P.sstride * es + 255) & ~(size_t)255;
    size_t tmp_bytes = 0;
    {
        auto begins = thrust::make_transform_iterator(thrust::counting_iterator<int>(0), SegBegin{P.sstride});
        auto ends = thrust::make_transform_iterator(thrust::counting_iterator<int>(0), SegEnd{nullptr, P.sstride});
        if (image_dtype == BPHOT_F32)
            CU_TRY(cub::DeviceSegmentedSort::SortKeys(nullptr, tmp_bytes, (const float*)nullptr, (float*)nullptr, (int)(batch * P.sstride), (int)batch, begins, ends, st));
        else
            CU_TRY(cub::DeviceSegmentedSort::SortKeys(nullptr, tmp_bytes, (const double*)nullptr, (double*)nullptr, (int)(batch * P.sstride), (int)batch, begins, ends, st));
    }
    tmp_bytes = (tmp_bytes + 255) & ~(size_t)255;
    const size_t cnt_bytes = ((size_t)batch * sizeof(int) + 255) & ~(size_t)255;
    const size_t ctr_bytes = ((size_t)(2 * nbatch + 1) * sizeof(int) + 255) & ~(size_t)255;
    char* block = nullptr;
    CU_TRY(cudaMallocAsync((void**)&block, 2 * row_bytes + tmp_bytes + cnt_bytes + ctr_bytes, st));
    int* seg_count = reinterpret_cast<int*>(block + 2 * row_bytes + tmp_bytes);
    int* counters = reinterpret_cast<int*>(block + 2 * row_bytes + tmp_bytes + cnt_bytes);
    int rc = 0;
    if (cudaMemsetAsync(counters, 0, ctr_bytes, st) != cudaSuccess) rc = fail(BPHOT_ERR_CUDA, "bphot_aper: work counters: %s", cudaGetErrorString(cudaGetLastError()));
    if (!rc) rc = ensure_log10_table(st);
    if (!rc) {
        if (image_dtype == BPHOT_F32)
            rc = aper_big_annulus_run<float>(P, nstars, (int)batch, (float*)block, (float*)(block + row_bytes), block + 2 * row_bytes, tmp_bytes, seg_count, counters, smem, st);
        else
            rc = aper_big_annulus_run<double>(P, nstars, (int)batch, (double*)block, (double*)(block + row_bytes), block + 2 * row_bytes, tmp_bytes, seg_count, counters, smem, st);
    }
    cudaFreeAsync(block, st);
    return rc;
}

// bphot_mmm_device for vectors above one CTA's shared memory (mmm_prep_global_kernel / mmm_global_kernel)
template <typename T>
int mmm_big(const T* d_vec, const long long* d_offs, int nvec, double readnoise, int minsky, int mxiter, double* d_out3, int* d_iters,
            int sms, cudaStream_t st) {
    using namespace bphot;
    long long total = 0;
    CU_TRY(cudaMemcpyAsync(&total, d_offs + nvec, sizeof(long long), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    if (total < 0 || total > 0x7fffffffll) return fail(BPHOT_ERR_CAPACITY, "bphot_mmm: %lld elements in one call (limit 2^31 - 1 on the large-vector path)", total);
    size_t tmp_bytes = 0;
    CU_TRY(cub::DeviceSegmentedSort::SortKeys(nullptr, tmp_bytes, (const T*)nullptr, (T*)nullptr, (int)total, nvec, d_offs, d_offs + 1, st));
    tmp_bytes = (tmp_bytes + 255) & ~(size_t)255;
    const size_t col = ((size_t)total * sizeof(T) + 255) & ~(size_t)255;
    char* block = nullptr;
    CU_TRY(cudaMallocAsync((void**)&block, 2 * col + tmp_bytes + (size_t)nvec * sizeof(int), st));
    T* copy = reinterpret_cast<T*>(block), *sorted = reinterpret_cast<T*>(block + col);
    int* nvalid = reinterpret_cast<int*>(block + 2 * col + tmp_bytes);
    mmm_prep_global_kernel<T><<<nvec < sms * 8 ? nvec : sms * 8, 256, 0, st>>>(d_vec, d_offs, nvec, copy, nvalid);
    cudaError_t e = cub::DeviceSegmentedSort::SortKeys(block + 2 * col, tmp_bytes, (const T*)copy, sorted, (int)total, nvec, d_offs, d_offs + 1, st);
    if (e == cudaSuccess) {
        const int blocks = (nvec + AP_WARPS - 1) / AP_WARPS;
        mmm_global_kernel<T><<<blocks < sms * 8 ? blocks : sms * 8, AP_THREADS, 0, st>>>(sorted, d_offs, nvalid, nvec, readnoise, minsky, mxiter, d_out3, d_iters);
        e = cudaGetLastError();
    }
    cudaFreeAsync(block, st);
    g_launches += 2;
    if (e != cudaSuccess) return fail(BPHOT_ERR_CUDA, "bphot_mmm (large vectors): %s", cudaGetErrorString(e));
    return BPHOT_OK;
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
#define BP_STR2(x) #x
#define BP_STR(x) BP_STR2(x)
const char* bphot_build_flags(void) {
    return "AP_ASM_SELECT=" BP_STR(AP_ASM_SELECT) " AP_BATCH=" BP_STR(AP_BATCH) " AP_GROUP=" BP_STR(AP_GROUP) " AP_S1_CTAS=" BP_STR(AP_S1_CTAS) " AP_MMM_INDEXED=" BP_STR(AP_MMM_INDEXED) " AP_MMM_OVERLAP=" BP_STR(AP_MMM_OVERLAP)
           " AP_S2_MINB=" BP_STR(AP_S2_MINB) " AP_SUMS_CTAS=" BP_STR(AP_SUMS_CTAS) " AP_DYNAMIC=" BP_STR(AP_DYNAMIC)
           " PK_MINB_FORCED=" BP_STR(PK_MINB_FORCED) " PK_U=" BP_STR(PK_U) " PK_MINB_FIT=" BP_STR(PK_MINB_FIT) " PK_HOIST=" BP_STR(PK_HOIST) " PK_PSF_GLOBAL=" BP_STR(PK_PSF_GLOBAL) " AP_IDX_SKIP=" BP_STR(AP_IDX_SKIP) " AP_SIGNED_SHFL=" BP_STR(AP_SIGNED_SHFL) " AP_SMEM_UNIFORM=" BP_STR(AP_SMEM_UNIFORM) " PK_RECIP=" BP_STR(PK_RECIP) " PW_ONE_ATAN=" BP_STR(PW_ONE_ATAN) " PK_TSUM=" BP_STR(PK_TSUM) " PK_FMA_MODEL=" BP_STR(PK_FMA_MODEL) " PK_FMA_LOOP=" BP_STR(PK_FMA_LOOP) " PK_FAST_RCP=" BP_STR(PK_FAST_RCP);
}
const char* bphot_last_error(void) { return g_err; }
long long bphot_launch_count(void) { return g_launches.load(); }

int bphot_fetch_host(void* d_dst, const void* h_src, long long nbytes, void* stream) {
    if (nbytes < 0 || (nbytes > 0 && (!d_dst || !h_src))) return fail(BPHOT_ERR_ARG, "bphot_fetch_host: null pointer");
    if (nbytes == 0) return BPHOT_OK;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, h_src) != cudaSuccess || at.type != cudaMemoryTypeHost || !at.devicePointer) {
        cudaGetLastError();
        return fail(BPHOT_ERR_ARG, "bphot_fetch_host: source is not page-locked host memory");
    }
    const bool aligned = (((uintptr_t)d_dst | (uintptr_t)at.devicePointer) & 15) == 0;
    const long long n16 = aligned ? nbytes / 16 : 0;
    long long blocks = (nbytes / 16 + 255) / 256;
    if (blocks > 148 * 4) blocks = 148 * 4;
    if (blocks < 1) blocks = 1;
    fetch_host_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((uint4*)d_dst, (const uint4*)at.devicePointer, n16, (unsigned char*)d_dst,
                                                                    (const unsigned char*)at.devicePointer, n16 * 16, nbytes);
    g_launches += 1;
    CU_TRY(cudaGetLastError());
    return BPHOT_OK;
}

int bphot_pkfit_capacity(double radius) {
    if (!(radius > 0.) || !(radius < 64.)) return fail(BPHOT_ERR_ARG, "bphot_pkfit_capacity: radius %g outside (0, 64)", radius);
    return (disc_lattice_bound(radius) + 7) / 8 * 8;
}

int bphot_profile(int enable) {
    const int prev = g_profile.exchange(enable ? 1 : 0);
    if (enable) for (int k = 0; k < KT_N; k++) { g_kt_ms[k] = 0.; g_kt_n[k] = 0; }
    return prev;
}
int bphot_kernel_times(double* ms, long long* launches) {
    if (!ms || !launches) return fail(BPHOT_ERR_ARG, "bphot_kernel_times: null pointer");
    for (int k = 0; k < 4; k++) { ms[k] = g_kt_ms[k]; launches[k] = g_kt_n[k]; }  // the finalize kernel is not reported
    return BPHOT_OK;
}

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
                      double readnoise, int flags,
                      double* d_mag, double* d_magerr, double* d_flux, double* d_fluxerr,
                      double* d_sky, double* d_skyerr, unsigned char* d_badflag,
                      int* d_nsky, int* d_mmm_iter, void* stream) {
    return bphot_aper_device_banded(d_image, image_dtype, nrow, ncol, pitch, d_xc, d_yc, nstars, phpadu, apr, naper, zeropoint,
                                    skyrad_in, skyrad_out, badlo, badhi, use_setsky, setsky, minsky, exact, readnoise, flags,
                                    d_mag, d_magerr, d_flux, d_fluxerr, d_sky, d_skyerr, d_badflag, d_nsky, d_mmm_iter,
                                    nullptr, 0, 0, stream);
}

int bphot_aper_device_banded(const void* d_image, int image_dtype, int nrow, int ncol, long long pitch,
                             const double* d_xc, const double* d_yc, int nstars,
                             double phpadu, const double* apr, int naper, double zeropoint,
                             double skyrad_in, double skyrad_out, double badlo, double badhi,
                             int use_setsky, const double* setsky, int minsky, int exact,
                             double readnoise, int flags,
                             double* d_mag, double* d_magerr, double* d_flux, double* d_fluxerr,
                             double* d_sky, double* d_skyerr, unsigned char* d_badflag,
                             int* d_nsky, int* d_mmm_iter,
                             const void* const* band_events, int nbands, int band_rows, void* stream) {
    using namespace bphot;
    if (nbands < 0 || (nbands > 0 && (!band_events || band_rows <= 0))) return fail(BPHOT_ERR_ARG, "bphot_aper: bad band description");
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
    {   // (double)v < badlo  <=>  v < (smallest float >= badlo);  (double)v > badhi  <=>  v > (largest float <= badhi)
        float lo = (float)badlo, hi = (float)badhi;
        if ((double)lo < badlo) lo = nextafterf(lo, INFINITY);
        if ((double)hi > badhi) hi = nextafterf(hi, -INFINITY);
        P.badlo_t = lo; P.badhi_t = hi;
    }
    P.minsky = minsky; P.exact = exact ? 1 : 0; P.compat_py2 = (flags & BPHOT_APER_PY2) ? 1 : 0; P.mxiter = 50;
    P.sky_sigclip = (flags & BPHOT_APER_SIGMACLIP) ? 1 : 0;
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
    const size_t smem = key_bytes + (size_t)P.span_rows * 20;  // keys | span[4*rows] | rowoff[rows]
    DevInfo d;
    int rc = dev_info(&d);
    if (rc) return rc;
    if (smem + 16384 > (size_t)d.smem_optin) {  // the annulus does not fit a CTA: rows in global memory, segmented device sort
        if (nbands > 0) cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)band_events[nbands - 1], 0);  // needs the whole image
        return aper_big_annulus(P, image_dtype, bound, nstars, d.smem_optin, (cudaStream_t)stream);
    }
    // scratch rows of sorted annulus keys (stage 1 -> stage 2), one batch of stars at a time
    P.sstride = (bound + 15) / 16 * 16;  // rows are 64-byte aligned: the indexed mmm reads aligned groups of 16 keys
    const int batch = nstars < APER_BATCH ? nstars : APER_BATCH;
    cudaStream_t st = (cudaStream_t)stream;
    rc = ensure_log10_table(st);
    if (rc) return rc;
    void* scratch = nullptr;
    // thread-per-star mmm on an index of the sorted row (aper_kernel.cu sort_index): register-sort variants, mmm sky
    const bool indexed = AP_MMM_INDEXED && K > 0 && !use_setsky && !P.sky_sigclip;
    // scratch for `group` stage-1 launches: one mmm launch then works on up to 262 144 stars (5.7 -> 4.5 ms per 1 M stars: it
    // waits for memory and needs the warps); stage 1 keeps its 65 536-star launches
    const int group_max = (AP_MMM_OVERLAP ? 1 : AP_GROUP);
    const int group = (nstars + batch - 1) / batch < group_max ? (nstars + batch - 1) / batch : group_max;
    const long long gstars = (long long)batch * group < nstars ? (long long)batch * group : nstars;
    const size_t keys_bytes = ((size_t)gstars * P.sstride * es + 255) & ~(size_t)255;
    P.idx_stride = indexed ? (long long)sort_index_bytes(es) : 0;
    const int nbatch = (nstars + batch - 1) / batch;
    // AP_MMM_OVERLAP (off): the thread-per-star mmm waits for memory and leaves the issue slots idle, stage 1 is bound by issue
    // slots and leaves exactly 4 096 registers per SM free (8 CTAs x 96 threads x 80 registers), so mmm was run in 64-thread
    // blocks of <= 64 registers on a side stream, batch b under stage 1 of batch b+1, on a second scratch buffer.  Measured on
    // B200: aper 56.0 ms with, 55.7 ms without -- the side kernel's blocks do not get SM time while the persistent stage-1 grid
    // is resident -- so the knob stays off.  Per-kernel timing (bphot_profile) always keeps everything on one stream.
    const bool overlap = AP_MMM_OVERLAP && indexed && nbatch > 1 && g_profile.load() == 0;
    const int nbuf = overlap ? 2 : 1;
    const size_t buf_bytes = (keys_bytes + (size_t)gstars * (size_t)P.idx_stride + 255) & ~(size_t)255;
    if (!use_setsky) CU_TRY(cudaMallocAsync(&scratch, (size_t)nbuf * buf_bytes, st));
    P.skeys = scratch;
    P.sidx = indexed ? (void*)(reinterpret_cast<char*>(scratch) + keys_bytes) : nullptr;
    // one zeroed work counter per stage-1 / retry / stage-2 launch
    int* counters = nullptr;
    {   // a failure here must not leak the scratch block that is already allocated
        cudaError_t e = cudaMallocAsync((void**)&counters, (size_t)(3 * nbatch + 2) * sizeof(int), st);
        if (e == cudaSuccess) e = cudaMemsetAsync(counters, 0, (size_t)(3 * nbatch + 2) * sizeof(int), st);
        if (e != cudaSuccess) {
            if (counters) cudaFreeAsync(counters, st);
            if (scratch) cudaFreeAsync(scratch, st);
            return fail(BPHOT_ERR_CUDA, "bphot_aper: work counters: %s", cudaGetErrorString(e));
        }
    }
    int* next_counter = counters;
    StarOrder ord;
    rc = make_star_order(d_xc, d_yc, nstars, nrow, ncol, st, &ord);
    if (rc) { if (scratch) cudaFreeAsync(scratch, st); cudaFreeAsync(counters, st); return rc; }
    P.order = ord.order;
    // image still in flight (band_events): which band has to be resident before each batch of stars may start
    std::vector<int> batch_band;
    auto wait_band = [&](int band) { cudaStreamWaitEvent(st, (cudaEvent_t)band_events[band < nbands - 1 ? band : nbands - 1], 0); };
    if (nbands > 0) {
        batch_band.assign(nbatch, nbands - 1);
        if (ord.order) {
            float* d_max = nullptr;
            std::vector<float> h_max(nbatch, 0.f);
            cudaError_t e = cudaMallocAsync((void**)&d_max, nbatch * sizeof(float), st);
            if (e == cudaSuccess) {
                batch_maxrow_kernel<<<nbatch, 256, 0, st>>>(d_yc, ord.order, nstars, batch, d_max);
                e = cudaMemcpyAsync(h_max.data(), d_max, nbatch * sizeof(float), cudaMemcpyDeviceToHost, st);
                if (e == cudaSuccess) e = cudaStreamSynchronize(st);
                cudaFreeAsync(d_max, st);
                g_launches += 1;
            }
            if (e == cudaSuccess)
                for (int b = 0; b < nbatch; b++) {
                    const double last = (double)h_max[b] + skyrad_out + 2.0;  // lowest image row any stamp of the batch reads
                    const long long band = (long long)(last < 0 ? 0 : last) / band_rows;
                    batch_band[b] = (int)(band < nbands - 1 ? band : nbands - 1);
                }
        }
    }
    int grid1 = 0, grid2 = 0;
    // stage-1 CTA: 4 warps, or 3 when every annulus fits 96 threads x K keys (the 4th warp of the network would hold
    // only padding; without it 8 CTAs fit an SM instead of 6)
    // `bound` is an analytic upper limit (3 235 for skyrad=[40,50], true maximum 2 840): the narrow pass is chosen on
    // the expected count (annulus area + 3 %); a star that still exceeds 96 x K is marked and redone by a 128-thread pass
    const double expect = M_PI * (skyrad_out * skyrad_out - skyrad_in * skyrad_in) * 1.03;
    const int nt1 = (K > 0 && !use_setsky && (expect <= 96.0 * K || (flags & BPHOT_APER_NARROW_CTAS))) ? 96 : AP_THREADS;
    LaunchTimer lt(st);
#define BP_STAGE1(T, KK)                                                                          \
    do {                                                                                          \
        rc = persistent_grid(aper_stage1_kernel<T, KK>, nt1, smem, P.scount, &grid1);             \
        P.work_counter = next_counter++;                                                          \
        if (!rc) aper_stage1_kernel<T, KK><<<grid1, nt1, smem, st>>>(P);                          \
        if (!rc && nt1 != AP_THREADS) { /* stars whose annulus did not fit 96 x K keys */         \
            P.retry_wide = 1;                                                                     \
            P.work_counter = next_counter++;                                                      \
            rc = persistent_grid(aper_stage1_kernel<T, KK>, AP_THREADS, smem, P.scount, &grid1);  \
            if (!rc) aper_stage1_kernel<T, KK><<<grid1, AP_THREADS, smem, st>>>(P);               \
            P.retry_wide = 0;                                                                     \
            g_launches += 1;                                                                      \
        }                                                                                         \
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
        if (indexed) {                                                                                            \
            aper_mmm_kernel<T><<<(P.scount + AP_MMM_THREADS - 1) / AP_MMM_THREADS, AP_MMM_THREADS, 0, overlap ? s2 : st>>>(P); \
            break;                                                                                                \
        }                                                                                                         \
        rc = persistent_grid(aper_stage2_kernel<T>, AP_THREADS, 0, (P.scount + AP_WARPS - 1) / AP_WARPS, &grid2); \
        if (!rc && AP_S2_CTAS > 0 && grid2 > AP_S2_CTAS * d.sms) grid2 = AP_S2_CTAS * d.sms;                      \
        P.work_counter = next_counter++;                                                                          \
        if (!rc) aper_stage2_kernel<T><<<grid2, AP_THREADS, 0, st>>>(P);                                          \
    } while (0)
    // Aperture sums of every star (independent of the sky): one launch, warp per star.  The kernel is bound by the
    // fp64 pipe (atan in pixwt) while stage 1 is bound by integer/fp32 issue and shuffles, so it runs on a side stream,
    // concurrently with the stage-1 / stage-2 batches, on a few CTAs per SM; the finalize kernel joins the two.
    // While per-kernel timing is on (bphot_profile) everything stays on the caller's stream, one kernel at a time.
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    if (AP_SIDE_STREAM && !lt.on && nstars >= ORDER_MIN_STARS) {
        rc = side_stream(&side);
        if (rc) { if (scratch) cudaFreeAsync(scratch, st); cudaFreeAsync(counters, st); ord.release(st); return rc; }
        cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming);
        cudaEventRecord(ev_fork, st);
        cudaStreamWaitEvent(side, ev_fork, 0);
    }
    auto launch_sums = [&]() {
        int grida = 0;
        const long long nb = ((long long)nstars + AP_WARPS - 1) / AP_WARPS;
        cudaStream_t sa = side ? side : st;
        P.work_counter = next_counter++;
        lt.begin(KT_APER_SUMS);
        if (image_dtype == BPHOT_F32) {
            rc = persistent_grid(aper_sums_kernel<float>, AP_THREADS, 0, nb, &grida);
            if (!rc && side && AP_SUMS_SIDE_CTAS > 0 && grida > AP_SUMS_SIDE_CTAS * d.sms) grida = AP_SUMS_SIDE_CTAS * d.sms;
            if (!rc) aper_sums_kernel<float><<<grida, AP_THREADS, 0, sa>>>(P);
        } else {
            rc = persistent_grid(aper_sums_kernel<double>, AP_THREADS, 0, nb, &grida);
            if (!rc && side && AP_SUMS_SIDE_CTAS > 0 && grida > AP_SUMS_SIDE_CTAS * d.sms) grida = AP_SUMS_SIDE_CTAS * d.sms;
            if (!rc) aper_sums_kernel<double><<<grida, AP_THREADS, 0, sa>>>(P);
        }
        lt.end();
        g_launches += 1;
        if (side) cudaEventRecord(ev_join, side);
    };
    if (nbands == 0) launch_sums();  // otherwise after the batches: by then the whole image is resident
    cudaStream_t s2 = nullptr;
    std::vector<cudaEvent_t> ev_s1, ev_m;
    if (overlap) {
        rc = side_stream(&s2);
        if (rc) { if (scratch) cudaFreeAsync(scratch, st); cudaFreeAsync(counters, st); ord.release(st); return rc; }
        ev_s1.resize(nbatch); ev_m.resize(nbatch);
        for (int b = 0; b < nbatch; b++) {
            cudaEventCreateWithFlags(&ev_s1[b], cudaEventDisableTiming);
            cudaEventCreateWithFlags(&ev_m[b], cudaEventDisableTiming);
        }
    }
    for (int s0 = 0; s0 < nstars && !rc; s0 += batch) {
        const int b = s0 / batch, gb = b % group;
        P.s0 = s0;
        P.scount = (nstars - s0) < batch ? (nstars - s0) : batch;
        P.row0 = gb * batch;
        if (scratch) {
            char* buf = reinterpret_cast<char*>(scratch) + (size_t)((b / group) % nbuf) * buf_bytes;
            P.skeys = buf;
            P.sidx = indexed ? (void*)(buf + keys_bytes) : nullptr;
        }
        if (nbands > 0) wait_band(batch_band[b]);
        if (overlap && b >= nbuf) cudaStreamWaitEvent(st, ev_m[b - nbuf], 0);  // the buffer's previous reader is done
        lt.begin(KT_APER_STAGE1);
        if (image_dtype == BPHOT_F32) BP_STAGE1_K(float);
        else BP_STAGE1_K(double);
        lt.end();
        if (rc) break;
        if (overlap) {
            cudaEventRecord(ev_s1[b], st);
            cudaStreamWaitEvent(s2, ev_s1[b], 0);
        }
        if (gb != group - 1 && s0 + batch < nstars) { g_launches += 1; continue; }  // stage 2 after the group's last stage-1 launch
        P.s0 = (b - gb) * batch;
        P.scount = s0 + P.scount - P.s0;
        P.row0 = 0;
        lt.begin(KT_APER_STAGE2);
        if (image_dtype == BPHOT_F32) BP_STAGE2(float);
        else BP_STAGE2(double);
        lt.end();
        if (overlap) cudaEventRecord(ev_m[b], s2);
        g_launches += 2;
    }
    if (overlap) {
        for (int b = 0; b < nbatch; b++) cudaStreamWaitEvent(st, ev_m[b], 0);  // unrecorded events (error path) are no-ops
        for (int b = 0; b < nbatch; b++) { cudaEventDestroy(ev_s1[b]); cudaEventDestroy(ev_m[b]); }
    }
#undef BP_STAGE2
#undef BP_STAGE1_K
#undef BP_STAGE1
    if (nbands > 0 && !rc) {
        wait_band(nbands - 1);
        P.s0 = 0; P.scount = nstars;
        launch_sums();
    }
    if (side) {
        cudaStreamWaitEvent(st, ev_join, 0);
        cudaEventDestroy(ev_fork);
        cudaEventDestroy(ev_join);
    }
    if (!rc) {
        int gf = (nstars + 255) / 256;
        if (gf > d.sms * 8) gf = d.sms * 8;
        lt.begin(KT_APER_FINAL);
        aper_finalize_kernel<<<gf, 256, 0, st>>>(P);
        lt.end();
        g_launches += 1;
    }
    lt.flush();
    if (scratch) cudaFreeAsync(scratch, st);
    cudaFreeAsync(counters, st);
    ord.release(st);
    if (rc) return rc;
    CU_TRY(cudaGetLastError());
    return BPHOT_OK;
}

// ---- bphot_aper_device_noalloc: bphot_aper_device for callers that may not allocate (include/bphot.h).
// Its scratch (work counters | sorted annulus rows | row indexes) is g_noalloc_scratch, a file-scope device array that every GPU
// loading the library holds until exit.  One stage-1 launch (+ the wide retry), one stage-2 launch, the aperture sums and the
// finalize kernel, all on the caller's stream, stars in caller order: no visiting-order sort, no side stream, no events.
int bphot_aper_device_noalloc(const void* d_image, int image_dtype, int nrow, int ncol, long long pitch,
                              const double* d_xc, const double* d_yc, int nstars,
                              double phpadu, const double* apr, int naper, double zeropoint,
                              double skyrad_in, double skyrad_out, double badlo, double badhi,
                              int use_setsky, const double* setsky, int minsky, int exact,
                              double readnoise, int flags,
                              double* d_mag, double* d_magerr, double* d_flux, double* d_fluxerr,
                              double* d_sky, double* d_skyerr, unsigned char* d_badflag,
                              int* d_nsky, int* d_mmm_iter, void* stream) {
    using namespace bphot;
    if (!d_image || !apr || nstars < 0 || nrow <= 0 || ncol <= 0) return fail(BPHOT_ERR_ARG, "bphot_aper_noalloc: null image/apr or bad shape");
    if (naper < 1 || naper > BPHOT_MAX_APER) return fail(BPHOT_ERR_ARG, "bphot_aper_noalloc: naper=%d outside 1..%d", naper, BPHOT_MAX_APER);
    if (image_dtype != BPHOT_F32 && image_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_aper_noalloc: image dtype %d", image_dtype);
    if (use_setsky && !setsky) return fail(BPHOT_ERR_ARG, "bphot_aper_noalloc: use_setsky without setsky[3]");
    if (nstars == 0) return BPHOT_OK;
    if (!d_xc || !d_yc || !d_mag || !d_magerr || !d_flux || !d_fluxerr || !d_sky || !d_skyerr || !d_badflag || !d_nsky || !d_mmm_iter)
        return fail(BPHOT_ERR_ARG, "bphot_aper_noalloc: null star list or output pointer");
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
    if (!(skyrad_out >= 0.) || !(skyrad_in >= 0.) || !(skyrad_out < 1e4)) return fail(BPHOT_ERR_ARG, "bphot_aper_noalloc: skyrad out of range");
    P.skyrad_in = skyrad_in; P.skyrad_out = skyrad_out; P.badlo = badlo; P.badhi = badhi;
    {   // the fp32 form of the bad-pixel window, as in bphot_aper_device_banded
        float lo = (float)badlo, hi = (float)badhi;
        if ((double)lo < badlo) lo = nextafterf(lo, INFINITY);
        if ((double)hi > badhi) hi = nextafterf(hi, -INFINITY);
        P.badlo_t = lo; P.badhi_t = hi;
    }
    P.minsky = minsky; P.exact = exact ? 1 : 0; P.compat_py2 = (flags & BPHOT_APER_PY2) ? 1 : 0; P.mxiter = 50;
    P.sky_sigclip = (flags & BPHOT_APER_SIGMACLIP) ? 1 : 0;
    P.readnoise = readnoise;
    const long long bound = use_setsky ? 4 : annulus_bound(skyrad_in, skyrad_out);
    const int need = next_pow2(bound);
    const int K = need <= 1024 ? 8 : need <= 2048 ? 16 : need <= 4096 ? 32 : 0;
    P.sort_cap = K ? AP_THREADS * K : need;
    P.mag = d_mag; P.magerr = d_magerr; P.flux = d_flux; P.fluxerr = d_fluxerr;
    P.sky = d_sky; P.skyerr = d_skyerr; P.badflag = d_badflag; P.nsky = d_nsky; P.mmm_iter = d_mmm_iter;
    P.span_rows = (int)(2. * skyrad_out) + 4;
    const size_t es = esize(image_dtype);
    const size_t key_bytes = ((size_t)(K ? AP_THREADS * (K + 1) : need) * es + 15) & ~(size_t)15;
    const size_t smem = key_bytes + (size_t)P.span_rows * 20;
    int dev = 0, sms = 0, smem_optin = 0;
    CU_TRY(cudaGetDevice(&dev));
    CU_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CU_TRY(cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    if (smem + 16384 > (size_t)smem_optin)
        return fail(BPHOT_ERR_CAPACITY, "bphot_aper_noalloc: sky annulus [%g,%g] needs %zu B of shared memory per star (limit %d)",
                    skyrad_in, skyrad_out, smem, smem_optin);
    P.sstride = (bound + 15) / 16 * 16;
    const bool indexed = AP_MMM_INDEXED && K > 0 && !use_setsky && !P.sky_sigclip;
    P.idx_stride = indexed ? (long long)sort_index_bytes(es) : 0;
    // the limit: one stage-1 launch, and counters + rows + indexes inside the static array
    const size_t per_star = use_setsky ? 0 : (size_t)P.sstride * es + (size_t)P.idx_stride;
    const size_t keys_bytes = use_setsky ? 0 : (((size_t)nstars * P.sstride * es + 255) & ~(size_t)255);
    const size_t total = NOALLOC_HEAD + keys_bytes + (size_t)nstars * (size_t)P.idx_stride;
    if (nstars > APER_BATCH || total > (size_t)BPHOT_NOALLOC_SCRATCH_BYTES) {
        long long most = per_star ? (long long)(((size_t)BPHOT_NOALLOC_SCRATCH_BYTES - NOALLOC_HEAD - 256) / per_star) : APER_BATCH;
        if (most > APER_BATCH) most = APER_BATCH;
        return fail(BPHOT_ERR_CAPACITY, "bphot_aper_noalloc: %d stars need %zu B of scratch, the static array holds %llu B (%zu B per star "
                    "for sky annulus [%g,%g]: at most %lld stars per call); use bphot_aper_device for more",
                    nstars, total, (unsigned long long)BPHOT_NOALLOC_SCRATCH_BYTES, per_star, skyrad_in, skyrad_out, most);
    }
    cudaStream_t st = (cudaStream_t)stream;
    int rc = ensure_log10_table(st);
    if (rc) return rc;
    unsigned char* base = nullptr;
    rc = noalloc_scratch(&base);
    if (rc) return rc;
    // the array keeps the previous call's contents: the work counters must start at zero, rows and indexes are written before read
    CU_TRY(cudaMemsetAsync(base, 0, NOALLOC_HEAD, st));
    int* next_counter = reinterpret_cast<int*>(base);
    P.skeys = use_setsky ? nullptr : (void*)(base + NOALLOC_HEAD);
    P.sidx = indexed ? (void*)(base + NOALLOC_HEAD + keys_bytes) : nullptr;
    P.order = nullptr;
    P.s0 = 0; P.scount = nstars; P.row0 = 0;
    const double expect = M_PI * (skyrad_out * skyrad_out - skyrad_in * skyrad_in) * 1.03;
    const int nt1 = (K > 0 && !use_setsky && (expect <= 96.0 * K || (flags & BPHOT_APER_NARROW_CTAS))) ? 96 : AP_THREADS;
    int grid = 0;
#define BP_NA_STAGE1(T, KK)                                                                      \
    do {                                                                                         \
        rc = persistent_grid(aper_stage1_kernel<T, KK>, nt1, smem, P.scount, &grid);             \
        P.work_counter = next_counter++;                                                         \
        if (!rc) aper_stage1_kernel<T, KK><<<grid, nt1, smem, st>>>(P);                          \
        if (!rc && nt1 != AP_THREADS) {                                                          \
            P.retry_wide = 1;                                                                    \
            P.work_counter = next_counter++;                                                     \
            rc = persistent_grid(aper_stage1_kernel<T, KK>, AP_THREADS, smem, P.scount, &grid);  \
            if (!rc) aper_stage1_kernel<T, KK><<<grid, AP_THREADS, smem, st>>>(P);               \
            P.retry_wide = 0;                                                                    \
            g_launches += 1;                                                                     \
        }                                                                                        \
    } while (0)
#define BP_NA_STAGE1_K(T)                       \
    do {                                        \
        if (K == 8) BP_NA_STAGE1(T, 8);         \
        else if (K == 16) BP_NA_STAGE1(T, 16);  \
        else if (K == 32) BP_NA_STAGE1(T, 32);  \
        else BP_NA_STAGE1(T, 0);                \
    } while (0)
#define BP_NA_REST(T)                                                                                                \
    do {                                                                                                             \
        if (indexed) {                                                                                               \
            aper_mmm_kernel<T><<<(P.scount + AP_MMM_THREADS - 1) / AP_MMM_THREADS, AP_MMM_THREADS, 0, st>>>(P);      \
        } else {                                                                                                     \
            rc = persistent_grid(aper_stage2_kernel<T>, AP_THREADS, 0, (P.scount + AP_WARPS - 1) / AP_WARPS, &grid); \
            P.work_counter = next_counter++;                                                                         \
            if (!rc) aper_stage2_kernel<T><<<grid, AP_THREADS, 0, st>>>(P);                                          \
        }                                                                                                            \
        if (rc) break;                                                                                               \
        rc = persistent_grid(aper_sums_kernel<T>, AP_THREADS, 0, ((long long)nstars + AP_WARPS - 1) / AP_WARPS, &grid); \
        P.work_counter = next_counter++;                                                                             \
        if (!rc) aper_sums_kernel<T><<<grid, AP_THREADS, 0, st>>>(P);                                                \
    } while (0)
    if (image_dtype == BPHOT_F32) { BP_NA_STAGE1_K(float); if (!rc) BP_NA_REST(float); }
    else { BP_NA_STAGE1_K(double); if (!rc) BP_NA_REST(double); }
#undef BP_NA_REST
#undef BP_NA_STAGE1_K
#undef BP_NA_STAGE1
    if (rc) return rc;
    int gf = (nstars + 255) / 256;
    if (gf > sms * 8) gf = sms * 8;
    aper_finalize_kernel<<<gf, 256, 0, st>>>(P);
    g_launches += 4;
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
    if (variant < 0 || variant > 3) return fail(BPHOT_ERR_ARG, "bphot_pkfit: variant %d", variant);
    const bool analytic = (variant == BPHOT_PKFIT || variant == BPHOT_PKFIT_NORECENTER);
    if (!d_image || !gauss || !d_psf || npsf < 4 || nrow <= 0 || ncol <= 0 || nstars < 0) return fail(BPHOT_ERR_ARG, "bphot_pkfit: null image/gauss/psf or bad shape");
    if (!analytic && (!d_noise || !d_mask)) return fail(BPHOT_ERR_ARG, "bphot_pkfit: noise and mask images are required for variant %d", variant);
    if (image_dtype != BPHOT_F32 && image_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_pkfit: image dtype %d", image_dtype);
    if (!analytic && noise_dtype != BPHOT_F32 && noise_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_pkfit: noise dtype %d", noise_dtype);
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
    P.cap = (disc_lattice_bound(radius) + 7) / 8 * 8;  // lists are indexed q < ngood, nothing reads a whole 32-slot chunk
    P.errmag = d_errmag; P.chi = d_chi; P.sharp = d_sharp; P.niter = d_niter;
    P.scale_out = d_scale_out; P.x_out = d_x_out; P.y_out = d_y_out; P.chi2 = d_chi2;
    P.acap = 2 * (int)ceil(radius) + 4;  // int(x+r)+1 - int(x-r) + 1 <= 2r + 3
    const int narr = (variant == BPHOT_PKFIT_NORECENT) ? PK_NARR_FORCED : PK_NARR_FIT;
    const size_t tab = (PK_PSF_GLOBAL && variant == BPHOT_PKFIT_NORECENT) ? 0 : (size_t)npsf * npsf * 8;
    const size_t smem = tab + (size_t)PK_WARPS * ((size_t)P.cap * (narr * 8 + 4) + (size_t)P.acap * 32);
    DevInfo d;
    int rc = dev_info(&d);
    if (rc) return rc;
    if (smem + 1024 > (size_t)d.smem_optin)
        return fail(BPHOT_ERR_CAPACITY, "bphot_pkfit: radius %g / npsf %d need %zu B of shared memory (limit %d)", radius, npsf, smem, d.smem_optin);
    const long long nblocks = ((long long)nstars + PK_WARPS - 1) / PK_WARPS;
    int grid = 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int nd = analytic ? image_dtype : noise_dtype;
    StarOrder ord;
    rc = make_star_order(d_x, d_y, nstars, nrow, ncol, st, &ord);
    if (rc) return rc;
    P.order = ord.order;
    int* counter = nullptr;
    if (cudaMallocAsync((void**)&counter, sizeof(int), st) != cudaSuccess || cudaMemsetAsync(counter, 0, sizeof(int), st) != cudaSuccess) {
        ord.release(st);
        return fail(BPHOT_ERR_CUDA, "bphot_pkfit: work counter: %s", cudaGetErrorString(cudaGetLastError()));
    }
    P.work_counter = counter;
    LaunchTimer lt(st);
    lt.begin(KT_PKFIT);
#define BP_LAUNCH(V, TI, TN)                                                              \
    do {                                                                                  \
        rc = persistent_grid(pkfit_kernel<V, TI, TN>, PK_THREADS, smem, nblocks, &grid);  \
        if (rc) { ord.release(st); cudaFreeAsync(counter, st); return rc; }               \
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
    } else if (variant == BPHOT_PKFIT_NORECENTER) {
        if (image_dtype == BPHOT_F32) BP_LAUNCH(3, float, float);
        else BP_LAUNCH(3, double, double);
    } else if (variant == BPHOT_PKFIT_NOISE) BP_LAUNCH_T(1);
    else BP_LAUNCH_T(2);
#undef BP_LAUNCH_T
#undef BP_LAUNCH
    lt.end();
    lt.flush();
    ord.release(st);
    cudaFreeAsync(counter, st);
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
    int grid = 0;
    cudaStream_t st = (cudaStream_t)stream;
    rc = ensure_log10_table(st);
    if (rc) return rc;
    if (smem + 1024 > (size_t)d.smem_optin) {  // longer than one CTA can sort: copy, segmented device sort, warp per vector
        if (dtype == BPHOT_F32) return mmm_big<float>((const float*)d_vec, d_offs, nvec, readnoise, minsky, mxiter, d_out3, d_iters, d.sms, st);
        return mmm_big<double>((const double*)d_vec, d_offs, nvec, readnoise, minsky, mxiter, d_out3, d_iters, d.sms, st);
    }
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

int bphot_find_device(const void* d_image, int image_dtype, int nrow, int ncol, long long pitch,
                      double hmin, double fwhm, double round_lo, double round_hi, double sharp_lo, double sharp_hi,
                      int max_out, double* d_x, double* d_y, double* d_flux, double* d_sharp, double* d_round,
                      int* counts, void* stream) {
    using namespace bphot;
    if (!d_image || !counts || nrow <= 0 || ncol <= 0 || max_out < 0) return fail(BPHOT_ERR_ARG, "bphot_find: null image/counts or bad shape");
    if (image_dtype != BPHOT_F32 && image_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_find: image dtype %d", image_dtype);
    if (!(fwhm >= 0.5)) return fail(BPHOT_ERR_ARG, "bphot_find: FWHM must be at least 0.5 pixels (find.py:103)");
    if (max_out > 0 && (!d_x || !d_y || !d_flux || !d_sharp || !d_round)) return fail(BPHOT_ERR_ARG, "bphot_find: null output pointer");
    const long long npix = (long long)nrow * ncol;
    if (npix > 2147483647LL) return fail(BPHOT_ERR_ARG, "bphot_find: more than 2^31-1 pixels");
    FindConsts hK;
    find_constants(fwhm, &hK);
    if (nrow < 2 * hK.nbox || ncol < 2 * hK.nbox) return fail(BPHOT_ERR_ARG, "bphot_find: image smaller than two convolution boxes (%d px)", hK.nbox);
    for (int k = 0; k < 5; k++) counts[k] = 0;
    cudaStream_t st = (cudaStream_t)stream;
    FindParams P;
    memset(&P, 0, sizeof(P));
    P.img = d_image; P.nrow = nrow; P.ncol = ncol; P.pitch = pitch > 0 ? pitch : ncol;
    P.hmin = hmin; P.round_lo = round_lo; P.round_hi = round_hi; P.sharp_lo = sharp_lo; P.sharp_hi = sharp_hi;
    const int tiles = ((ncol + FIND_TX - 1) / FIND_TX) * ((nrow + FIND_TY - 1) / FIND_TY);
    P.nblocks = tiles;
    // one stream-ordered block for the fixed-size scratch: h | blockmin | minh | consts | counts | flags
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t o_h = 0, o_bm = o_h + up((size_t)npix * 8), o_mh = o_bm + up((size_t)tiles * 8), o_K = o_mh + 256,
                 o_cnt = o_K + up(sizeof(FindConsts)), o_fl = o_cnt + 256, total = o_fl + up((size_t)npix);
    char* base = nullptr;
    CU_TRY(cudaMallocAsync((void**)&base, total, st));
    void* extra = nullptr;  // candidate-sized scratch, allocated once the count is known
    auto cleanup = [&]() { if (extra) cudaFreeAsync(extra, st); cudaFreeAsync(base, st); };
#define CU_TRY_F(expr)                                                                                   \
    do {                                                                                                 \
        cudaError_t e_ = (expr);                                                                         \
        if (e_ != cudaSuccess) { cleanup(); return fail(BPHOT_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); } \
    } while (0)
    P.h = reinterpret_cast<double*>(base + o_h); P.blockmin = reinterpret_cast<double*>(base + o_bm);
    P.minh = reinterpret_cast<double*>(base + o_mh); P.K = reinterpret_cast<const FindConsts*>(base + o_K);
    P.counts = reinterpret_cast<int*>(base + o_cnt); P.flags = reinterpret_cast<unsigned char*>(base + o_fl);
    CU_TRY_F(cudaMemcpyAsync(base + o_K, &hK, sizeof(FindConsts), cudaMemcpyHostToDevice, st));
    const size_t tile_bytes = (size_t)(FIND_TX + 2 * hK.nhalf) * (FIND_TY + 2 * hK.nhalf) * 8;
    DevInfo di;
    if (dev_info(&di)) { cleanup(); return BPHOT_ERR_CUDA; }
    int cgrid = tiles < di.sms * 8 ? tiles : di.sms * 8;  // persistent: 8 CTAs of 256 threads per SM
    P.ngrid = cgrid;
    if (image_dtype == BPHOT_F32) find_convolve_kernel<float><<<cgrid, FIND_THREADS, tile_bytes, st>>>(P);
    else find_convolve_kernel<double><<<cgrid, FIND_THREADS, tile_bytes, st>>>(P);
    find_min_kernel<<<1, FIND_THREADS, 0, st>>>(P);
    long long cb = (npix + FIND_THREADS - 1) / FIND_THREADS;
    if (cb > di.sms * 8) cb = di.sms * 8;
    find_candidate_kernel<<<(int)cb, FIND_THREADS, 0, st>>>(P);
    g_launches += 3;
    int hc[6] = {0, 0, 0, 0, 0, 0};
    CU_TRY_F(cudaMemcpyAsync(hc, P.counts, sizeof(hc), cudaMemcpyDeviceToHost, st));
    CU_TRY_F(cudaStreamSynchronize(st));
    const int ncand = hc[0];
    counts[1] = ncand;
    if (ncand == 0) { cleanup(); CU_TRY(cudaGetLastError()); return BPHOT_OK; }
    // candidate-sized scratch: cand | sel | nsel | cx cy cflux csharp cround | cok | cub temp
    size_t tmp1 = 0, tmp2 = 0;
    thrust::counting_iterator<int> iota(0);
    CU_TRY_F(cub::DeviceSelect::Flagged(nullptr, tmp1, iota, (const unsigned char*)nullptr, (int*)nullptr, (int*)nullptr, (int)npix, st));
    CU_TRY_F(cub::DeviceSelect::Flagged(nullptr, tmp2, iota, (const unsigned char*)nullptr, (int*)nullptr, (int*)nullptr, ncand, st));
    const size_t tmpb = up(tmp1 > tmp2 ? tmp1 : tmp2), col = up((size_t)ncand * 8);
    const size_t e_cand = 0, e_sel = e_cand + up((size_t)ncand * 4), e_n = e_sel + up((size_t)ncand * 4), e_col = e_n + 256,
                 e_ok = e_col + 5 * col, e_tmp = e_ok + up((size_t)ncand), etotal = e_tmp + tmpb;
    CU_TRY_F(cudaMallocAsync(&extra, etotal, st));
    char* eb = reinterpret_cast<char*>(extra);
    int* d_cand = reinterpret_cast<int*>(eb + e_cand);
    int* d_sel = reinterpret_cast<int*>(eb + e_sel);
    int* d_n = reinterpret_cast<int*>(eb + e_n);
    size_t tb = tmpb;
    CU_TRY_F(cub::DeviceSelect::Flagged(eb + e_tmp, tb, iota, P.flags, d_cand, d_n, (int)npix, st));
    P.cand = d_cand; P.ncand = ncand;
    P.cx = reinterpret_cast<double*>(eb + e_col); P.cy = reinterpret_cast<double*>(eb + e_col + col);
    P.cflux = reinterpret_cast<double*>(eb + e_col + 2 * col); P.csharp = reinterpret_cast<double*>(eb + e_col + 3 * col);
    P.cround = reinterpret_cast<double*>(eb + e_col + 4 * col); P.cok = reinterpret_cast<unsigned char*>(eb + e_ok);
    int sb = (ncand + 127) / 128;
    if (sb > 148 * 16) sb = 148 * 16;
    if (image_dtype == BPHOT_F32) find_stats_kernel<float><<<sb, 128, 0, st>>>(P);
    else find_stats_kernel<double><<<sb, 128, 0, st>>>(P);
    g_launches += 1;
    tb = tmpb;
    CU_TRY_F(cub::DeviceSelect::Flagged(eb + e_tmp, tb, iota, P.cok, d_sel, d_n, ncand, st));
    CU_TRY_F(cudaMemcpyAsync(hc, P.counts, sizeof(hc), cudaMemcpyDeviceToHost, st));
    CU_TRY_F(cudaStreamSynchronize(st));
    counts[0] = hc[4]; counts[2] = hc[1]; counts[3] = hc[2]; counts[4] = hc[3];
    const int nstars = hc[4];
    if (nstars > max_out) {
        cleanup();
        return fail(BPHOT_ERR_CAPACITY, "bphot_find: %d stars found, output arrays hold %d (counts[0] has the size needed)", nstars, max_out);
    }
    if (nstars > 0) {
        P.sel = d_sel; P.nsel = nstars;
        P.ox = d_x; P.oy = d_y; P.oflux = d_flux; P.osharp = d_sharp; P.oround = d_round;
        int gb = (nstars + 255) / 256;
        if (gb > 148 * 8) gb = 148 * 8;
        find_gather_kernel<<<gb, 256, 0, st>>>(P);
        g_launches += 1;
        CU_TRY_F(cudaStreamSynchronize(st));
    }
    cleanup();
#undef CU_TRY_F
    CU_TRY(cudaGetLastError());
    return BPHOT_OK;
}

int bphot_cntrd_device(const void* d_image, int image_dtype, int nrow, int ncol, long long pitch,
                       const double* d_x, const double* d_y, int npts, double fwhm,
                       int extend_x, int extend_y, int keepcenter,
                       double* d_xcen, double* d_ycen, int* d_status, void* stream) {
    using namespace bphot;
    if (!d_image || nrow <= 0 || ncol <= 0 || npts < 0 || extend_x < 0 || extend_y < 0) return fail(BPHOT_ERR_ARG, "bphot_cntrd: null image or bad shape");
    if (image_dtype != BPHOT_F32 && image_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_cntrd: image dtype %d", image_dtype);
    if (npts == 0) return BPHOT_OK;
    if (!d_x || !d_y || !d_xcen || !d_ycen || !d_status) return fail(BPHOT_ERR_ARG, "bphot_cntrd: null star list or output pointer");
    CntrdConsts hK;
    memset(&hK, 0, sizeof(hK));
    int nhalf = (int)(0.637 * fwhm);  // cntrd.py:91-93
    if (nhalf < 2) nhalf = 2;
    const int nbox = 2 * nhalf + 1;
    if (nbox > CNTRD_MAXBOX) return fail(BPHOT_ERR_CAPACITY, "bphot_cntrd: fwhm %g needs a %d-pixel box (limit %d)", fwhm, nbox, CNTRD_MAXBOX);
    hK.nhalf = nhalf; hK.nbox = nbox; hK.ir = nhalf - 1 < 1 ? 1 : nhalf - 1;
    hK.ex = extend_x; hK.ey = extend_y; hK.keepcenter = keepcenter ? 1 : 0;
    double t[CNTRD_MAXBOX];
    for (int k = 0; k < nbox - 1; k++) {  // cntrd.py:165-170
        hK.dd[k] = (double)k + 0.5 - nhalf;
        hK.w[k] = 1. - 0.5 * (fabs(hK.dd[k]) - 0.5) / (nhalf - 0.5);
        hK.wdd[k] = hK.w[k] * hK.dd[k];
        t[k] = hK.w[k] * (hK.dd[k] * hK.dd[k]);
    }
    hK.sumc = np_sum_host(hK.w, nbox - 1);
    hK.sumxsq = np_sum_host(t, nbox - 1);
    cudaStream_t st = (cudaStream_t)stream;
    CntrdConsts* dK = nullptr;
    CU_TRY(cudaMallocAsync((void**)&dK, sizeof(CntrdConsts), st));
    cudaError_t e = cudaMemcpyAsync(dK, &hK, sizeof(CntrdConsts), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);  // hK lives on this stack frame
    if (e != cudaSuccess) { cudaFreeAsync(dK, st); return fail(BPHOT_ERR_CUDA, "bphot_cntrd: %s", cudaGetErrorString(e)); }
    CntrdParams P;
    P.img = d_image; P.nrow = nrow; P.ncol = ncol; P.pitch = pitch > 0 ? pitch : ncol;
    P.x = d_x; P.y = d_y; P.npts = npts; P.K = dK; P.xcen = d_xcen; P.ycen = d_ycen; P.status = d_status;
    int blocks = (npts + 127) / 128;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (image_dtype == BPHOT_F32) cntrd_kernel<float><<<blocks, 128, 0, st>>>(P);
    else cntrd_kernel<double><<<blocks, 128, 0, st>>>(P);
    g_launches += 1;
    cudaFreeAsync(dK, st);
    CU_TRY(cudaGetLastError());
    return BPHOT_OK;
}

int bphot_find_host(const void* image, int image_dtype, int nrow, int ncol,
                    double hmin, double fwhm, double round_lo, double round_hi, double sharp_lo, double sharp_hi,
                    int max_out, double* x, double* y, double* flux, double* sharp, double* roundness,
                    int* counts, int device) {
    if (bphot_device_count() < 1) return fail(BPHOT_ERR_NODEVICE, "no CUDA device (libbphot has no CPU fallback)");
    if (!image || nrow <= 0 || ncol <= 0 || max_out < 0 || !counts) return fail(BPHOT_ERR_ARG, "bphot_find_host: bad argument");
    if (image_dtype != BPHOT_F32 && image_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_find_host: image dtype %d", image_dtype);
    CU_TRY(cudaSetDevice(device));
    DevBuf dimg, dout;
    CU_TRY(dimg.alloc((size_t)nrow * ncol * esize(image_dtype)));
    CU_TRY(cudaMemcpyAsync(dimg.p, image, (size_t)nrow * ncol * esize(image_dtype), cudaMemcpyHostToDevice, 0));
    const size_t col = (size_t)max_out * 8;
    CU_TRY(dout.alloc(5 * col));
    double* o = dout.as<double>();
    int rc = bphot_find_device(dimg.p, image_dtype, nrow, ncol, ncol, hmin, fwhm, round_lo, round_hi, sharp_lo, sharp_hi, max_out,
                               o, o + max_out, o + 2 * (size_t)max_out, o + 3 * (size_t)max_out, o + 4 * (size_t)max_out, counts, 0);
    if (rc) return rc;
    const size_t nb = (size_t)counts[0] * 8;
    if (nb) {
        CU_TRY(cudaMemcpy(x, o, nb, cudaMemcpyDeviceToHost));
        CU_TRY(cudaMemcpy(y, o + max_out, nb, cudaMemcpyDeviceToHost));
        CU_TRY(cudaMemcpy(flux, o + 2 * (size_t)max_out, nb, cudaMemcpyDeviceToHost));
        CU_TRY(cudaMemcpy(sharp, o + 3 * (size_t)max_out, nb, cudaMemcpyDeviceToHost));
        CU_TRY(cudaMemcpy(roundness, o + 4 * (size_t)max_out, nb, cudaMemcpyDeviceToHost));
    }
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
                    double readnoise, int flags,
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
                               flags, dmag.as<double>(), dmagerr.as<double>(), dflux.as<double>(), dfluxerr.as<double>(),
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
    const bool analytic = (variant == BPHOT_PKFIT || variant == BPHOT_PKFIT_NORECENTER);
    if (!analytic && (!noise || !mask)) return fail(BPHOT_ERR_ARG, "bphot_pkfit_host: noise and mask images are required for variant %d", variant);
    if (image_dtype != BPHOT_F32 && image_dtype != BPHOT_F64) return fail(BPHOT_ERR_ARG, "bphot_pkfit_host: image dtype %d", image_dtype);
    if (nstars == 0) return BPHOT_OK;
    CU_TRY(cudaSetDevice(device));
    const size_t nb = (size_t)nstars * sizeof(double), npx = (size_t)nrow * ncol;
    DevBuf dimg, dnz, dmk, dpsf, ds, dx, dy, dsk, o_err, o_chi, o_sh, o_it, o_sc, o_x, o_y, o_c2;
    UP(dimg, image, npx * esize(image_dtype));
    if (!analytic) {
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
