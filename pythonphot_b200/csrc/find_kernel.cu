// find_kernel.cu -- DAOFIND star detection (sm_100a): PythonPhot/find.py:89-394.
//
// The reference convolves the frame with a lowered, truncated Gaussian (scipy.ndimage.convolve,
// find.py:188), keeps pixels above `hmin` that are >= every neighbour inside the kernel footprint
// (a Python loop over footprint offsets, find.py:220-233), then runs a per-candidate Python loop for
// the sharpness / roundness statistics and the marginal-Gaussian centroid (find.py:259-346).  Here:
//   K1 find_convolve_kernel   32x16-pixel tiles staged with their halo in shared memory (coalesced
//                             row loads, reflected borders like scipy's default mode), fp64
//                             accumulation in scipy's footprint order, + a NaN-propagating block minimum
//   K2 find_min_kernel        min over blocks = np.min(h) (find.py:190), NaN if any h is NaN
//   K3 find_candidate_kernel  threshold + local-maximum test per pixel against the bordered h
//   --  cub::DeviceSelect     row-major ordered compaction (the order np.where gives, find.py:211)
//   K4 find_stats_kernel      one thread per candidate: box sums, statistics, centroid, verdict
//   K5 find_gather_kernel     accepted candidates -> output columns, order kept
#include "bphot_device.cuh"

namespace bphot {

constexpr int FIND_MAXBOX = 13;  // find.py:93
constexpr int FIND_TX = 64, FIND_TY = 32, FIND_THREADS = 256;  // output tile of the convolution

struct FindConsts {  // everything find.py:106-183 derives from fwhm alone; filled on the host (bphot_capi.cu)
    int nhalf, nbox, pixels;  // pixels: footprint size without the centre (find.py:200)
    double sigsq, p;
    double c[FIND_MAXBOX * FIND_MAXBOX];            // convolution kernel, 0 outside the footprint
    unsigned char ring[FIND_MAXBOX * FIND_MAXBOX];  // footprint minus centre
    signed char rdy[FIND_MAXBOX * FIND_MAXBOX], rdx[FIND_MAXBOX * FIND_MAXBOX];  // its offsets, nearest first (`pixels` of them)
    double c1[FIND_MAXBOX], wt[FIND_MAXBOX], sgx[FIND_MAXBOX], sgy[FIND_MAXBOX], dgdx[FIND_MAXBOX], dgdy[FIND_MAXBOX];
    double sumgx, sumgy, sumgsqx, sumgsqy, sdgdxs, sdgdx, sdgdys, sdgdy, sgdgdx, sgdgdy;
};

struct FindParams {
    const void* img;
    int nrow, ncol;
    long long pitch;
    const FindConsts* K;  // device copy
    double* h;            // [nrow, ncol] convolved image
    double* blockmin;     // [grid of K1]
    int nblocks;          // tiles of K1
    int ngrid;            // CTAs of K1 (= entries of blockmin)
    double* minh;         // [1]
    double hmin, round_lo, round_hi, sharp_lo, sharp_hi;
    unsigned char* flags; // [nrow*ncol] candidate flags
    int* counts;          // [6]: ncand, badsharp, badround, badcntrd, nstars, spare
    const int* cand;      // [ncand] flat pixel indices, row-major ascending
    int ncand;
    double* cx; double* cy; double* cflux; double* csharp; double* cround;  // per candidate
    unsigned char* cok;   // per candidate: 1 = meets every criterion
    const int* sel;       // [nstars] indices into the candidate arrays
    int nsel;
    double* ox; double* oy; double* oflux; double* osharp; double* oround;  // outputs
};

// np.min semantics: NaN wins
BP_D double nanmin(double a, double b) { return (a != a) ? a : ((b != b) ? b : fmin(a, b)); }

// scipy.ndimage 'reflect' extension (d c b a | a b c d | d c b a)
BP_D int reflect_index(int i, int n) {
    if ((unsigned)i < (unsigned)n) return i;
    if (n == 1) return 0;
    const int period = 2 * n;
    i %= period;
    if (i < 0) i += period;
    return i < n ? i : period - 1 - i;
}

template <typename T>
__global__ void __launch_bounds__(FIND_THREADS) find_convolve_kernel(const FindParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ FindConsts sK;
    __shared__ double s_min[FIND_THREADS / 32];
    double* tile = reinterpret_cast<double*>(smem_raw);
    const int tid = threadIdx.x;
    {   // constants -> shared memory (word copies)
        const int* src = reinterpret_cast<const int*>(P.K);
        int* dst = reinterpret_cast<int*>(&sK);
        for (int i = tid; i < (int)(sizeof(FindConsts) / 4); i += FIND_THREADS) dst[i] = src[i];
    }
    __syncthreads();
    const int nhalf = sK.nhalf, nbox = sK.nbox;
    const int tw = FIND_TX + 2 * nhalf, th = FIND_TY + 2 * nhalf;
    const int tiles_x = (P.ncol + FIND_TX - 1) / FIND_TX;
    const T* __restrict__ img = reinterpret_cast<const T*>(P.img);
    double mn = __longlong_as_double(0x7ff0000000000000LL);  // +inf
    const int lx = tid & 31;
    // persistent CTAs: the constants are staged once, then the CTA walks tiles (neighbouring CTAs take neighbouring
    // tiles of a tile row, so their halos overlap in L2)
    for (int tl = blockIdx.x; tl < P.nblocks; tl += gridDim.x) {
    const int x0 = (tl % tiles_x) * FIND_TX, y0 = (tl / tiles_x) * FIND_TY;
    __syncthreads();  // the previous tile's readers are done
    for (int ty = tid >> 5; ty < th; ty += FIND_THREADS / 32) {
        const T* __restrict__ grow = img + (size_t)reflect_index(y0 + ty - nhalf, P.nrow) * P.pitch;
        for (int tx = lx; tx < tw; tx += 32) tile[ty * tw + tx] = (double)grow[reflect_index(x0 + tx - nhalf, P.ncol)];
    }
    __syncthreads();
    for (int l = tid; l < FIND_TX * FIND_TY; l += FIND_THREADS) {
        const int ly = l / FIND_TX, lxx = l % FIND_TX;  // powers of two
        const int gx = x0 + lxx, gy = y0 + ly;
        if (gx < P.ncol && gy < P.nrow) {
            // scipy accumulates over the footprint (non-zero weights) in row-major order, find.py:188
            double acc = 0.0;
            for (int ky = 0; ky < nbox; ky++) {
                const double* trow = tile + (ly + ky) * tw + lxx;
                const double* crow = sK.c + ky * nbox;
                for (int kx = 0; kx < nbox; kx++) {
                    const double w = crow[kx];
                    if (w != 0.0) acc = acc + trow[kx] * w;
                }
            }
            P.h[(size_t)gy * P.ncol + gx] = acc;
            mn = nanmin(mn, acc);
        }
    }
    }  // tiles
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mn = nanmin(mn, __shfl_xor_sync(FULL, mn, o));
    if ((tid & 31) == 0) s_min[tid >> 5] = mn;
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < FIND_THREADS / 32; w++) mn = nanmin(mn, s_min[w]);
        P.blockmin[blockIdx.x] = mn;
    }
}

__global__ void __launch_bounds__(FIND_THREADS) find_min_kernel(const FindParams P) {
    __shared__ double s_min[FIND_THREADS / 32];
    double mn = __longlong_as_double(0x7ff0000000000000LL);
    for (int i = threadIdx.x; i < P.ngrid; i += FIND_THREADS) mn = nanmin(mn, P.blockmin[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mn = nanmin(mn, __shfl_xor_sync(FULL, mn, o));
    if ((threadIdx.x & 31) == 0) s_min[threadIdx.x >> 5] = mn;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < FIND_THREADS / 32; w++) mn = nanmin(mn, s_min[w]);
        P.minh[0] = mn;
        for (int k = 0; k < 6; k++) P.counts[k] = 0;
    }
}

// h after the border assignment of find.py:191-192.  (sic) rows n_y-nhalf .. n_y-2 are overwritten, the last
// row keeps its convolved values between the side borders.
BP_D double h_bordered(const FindParams& P, int nhalf, double minh, int y, int x) {
    const bool border = x < nhalf || x >= P.ncol - nhalf || y < nhalf || (y >= P.nrow - nhalf && y < P.nrow - 1);
    return border ? minh : P.h[(size_t)y * P.ncol + x];
}

// find.py:211-233: h >= hmin, every footprint neighbour inside the frame, and h >= each of them.
__global__ void __launch_bounds__(FIND_THREADS) find_candidate_kernel(const FindParams P) {
    __shared__ FindConsts sK;
    {
        const int* src = reinterpret_cast<const int*>(P.K);
        int* dst = reinterpret_cast<int*>(&sK);
        for (int i = threadIdx.x; i < (int)(sizeof(FindConsts) / 4); i += FIND_THREADS) dst[i] = src[i];
    }
    __syncthreads();
    const int nhalf = sK.nhalf;
    const double minh = P.minh[0];
    // one warp per 32-pixel row segment: coalesced reads of h, no 64-bit division per pixel
    const int segs_x = (P.ncol + 31) / 32;
    const long long nseg = (long long)segs_x * P.nrow;
    const int lane = threadIdx.x & 31;
    int found = 0;
    for (long long sg = blockIdx.x * (long long)(FIND_THREADS / 32) + (threadIdx.x >> 5); sg < nseg; sg += (long long)gridDim.x * (FIND_THREADS / 32)) {
        const int y = (int)(sg / segs_x), x = (int)(sg - (long long)y * segs_x) * 32 + lane;
        if (x >= P.ncol) continue;
        const double v = h_bordered(P, nhalf, minh, y, x);
        bool cand = v >= P.hmin;
        if (cand) {
            for (int k = 0; k < sK.pixels && cand; k++) {
                const int ny = y + sK.rdy[k], nx = x + sK.rdx[k];
                if (ny < 0 || ny >= P.nrow || nx < 0 || nx >= P.ncol) cand = false;
                else cand = v >= h_bordered(P, nhalf, minh, ny, nx);
            }
        }
        P.flags[(size_t)y * P.ncol + x] = cand ? 1 : 0;
        found += cand;
    }
    if (found) atomicAdd(&P.counts[0], found);
}

// numpy's pairwise add-reduce of n <= 13*13 contiguous float64 values (loops_utils.h.src: blocks of 128,
// 8 accumulators), so that np.sum() over a box or a marginal rounds the way the reference's does.
__device__ __noinline__ double np_sum_block(const double* a, int n) {
    if (n < 8) {
        double r = 0.0;
        for (int i = 0; i < n; i++) r = r + a[i];
        return r;
    }
    double r0 = a[0], r1 = a[1], r2 = a[2], r3 = a[3], r4 = a[4], r5 = a[5], r6 = a[6], r7 = a[7];
    int i = 8;
    for (; i < n - (n % 8); i += 8) {
        r0 = r0 + a[i]; r1 = r1 + a[i + 1]; r2 = r2 + a[i + 2]; r3 = r3 + a[i + 3];
        r4 = r4 + a[i + 4]; r5 = r5 + a[i + 5]; r6 = r6 + a[i + 6]; r7 = r7 + a[i + 7];
    }
    double res = ((r0 + r1) + (r2 + r3)) + ((r4 + r5) + (r6 + r7));
    for (; i < n; i++) res = res + a[i];
    return res;
}
BP_D double np_sum(const double* a, int n) {
    if (n <= 128) return np_sum_block(a, n);
    int n2 = n / 2;
    n2 -= n2 % 8;
    return np_sum_block(a, n2) + np_sum_block(a + n2, n - n2);  // n <= 169: both halves <= 128
}

// One thread per candidate: find.py:259-346.  The box (<= 13x13 fp64) and the marginals live in local memory.
template <typename T>
__global__ void __launch_bounds__(128) find_stats_kernel(const FindParams P) {
    __shared__ FindConsts sK;
    {
        const int* src = reinterpret_cast<const int*>(P.K);
        int* dst = reinterpret_cast<int*>(&sK);
        for (int i = threadIdx.x; i < (int)(sizeof(FindConsts) / 4); i += 128) dst[i] = src[i];
    }
    __syncthreads();
    const int nhalf = sK.nhalf, nbox = sK.nbox, nn = nbox * nbox;
    const double minh = P.minh[0];
    const T* __restrict__ img = reinterpret_cast<const T*>(P.img);
    double temp[FIND_MAXBOX * FIND_MAXBOX], work[FIND_MAXBOX * FIND_MAXBOX], sd[FIND_MAXBOX], v[FIND_MAXBOX];
    for (int i = blockIdx.x * 128 + threadIdx.x; i < P.ncand; i += gridDim.x * 128) {
        const int flat = P.cand[i];
        const int iy = flat / P.ncol, ix = flat - iy * P.ncol;
        // every footprint neighbour is in the frame (K3), so the nbox x nbox box is too
        const T* __restrict__ box = img + (size_t)(iy - nhalf) * P.pitch + (ix - nhalf);
        for (int r = 0; r < nbox; r++)
            for (int c = 0; c < nbox; c++) temp[r * nbox + c] = (double)box[(size_t)r * P.pitch + c];
        const double d = h_bordered(P, nhalf, minh, iy, ix);
        int verdict = 0;  // 1 sharpness, 2 roundness, 3 centroid
        double around = 0., ddx = 0., ddy = 0.;
        // sharpness, find.py:265 (mask*temp is summed over the whole box: a NaN under a zero mask entry still poisons it)
        for (int k = 0; k < nn; k++) work[k] = (double)sK.ring[k] * temp[k];
        const double sharp1 = (temp[nhalf * nbox + nhalf] - np_sum(work, nn) / sK.pixels) / d;
        if (sharp1 < P.sharp_lo || sharp1 > P.sharp_hi) verdict = 1;
        if (!verdict) {  // roundness, find.py:272-281; np.sum(temp, axis=0) adds the rows in order, axis=1 is pairwise per row
            for (int c = 0; c < nbox; c++) {
                double s = temp[c];
                for (int r = 1; r < nbox; r++) s = s + temp[r * nbox + c];
                v[c] = s * sK.c1[c];
            }
            const double dx = np_sum(v, nbox);
            for (int r = 0; r < nbox; r++) v[r] = np_sum(temp + r * nbox, nbox) * sK.c1[r];
            const double dy = np_sum(v, nbox);
            if (dx <= 0. || dy <= 0.) verdict = 2;
            else {
                around = 2 * (dx - dy) / (dx + dy);
                if (around < P.round_lo || around > P.round_hi) verdict = 2;
            }
        }
        if (!verdict) {  // x centroid from the marginal along y, find.py:293-313
            for (int c = 0; c < nbox; c++) {
                double s = temp[c] * sK.wt[0];
                for (int r = 1; r < nbox; r++) s = s + temp[r * nbox + c] * sK.wt[r];
                sd[c] = s;
            }
            for (int c = 0; c < nbox; c++) v[c] = sK.wt[c] * sK.sgy[c] * sd[c];
            const double sumgd = np_sum(v, nbox);
            for (int c = 0; c < nbox; c++) v[c] = sK.wt[c] * sd[c];
            const double sumd = np_sum(v, nbox);
            for (int c = 0; c < nbox; c++) v[c] = sK.wt[c] * sd[c] * sK.dgdx[c];
            const double sddgdx = np_sum(v, nbox);
            const double hx = (sumgd - sK.sumgx * sumd / sK.p) / (sK.sumgsqy - sK.sumgx * sK.sumgx / sK.p);
            if (hx <= 0.) verdict = 3;
            else {
                const double skylvl = (sumd - hx * sK.sumgx) / sK.p;
                ddx = (sK.sgdgdx - (sddgdx - sK.sdgdx * (hx * sK.sumgx + skylvl * sK.p))) / (hx * sK.sdgdxs / sK.sigsq);
                if (fabs(ddx) >= nhalf) verdict = 3;
            }
        }
        if (!verdict) {  // y centroid, find.py:317-338
            for (int r = 0; r < nbox; r++) {
                for (int c = 0; c < nbox; c++) work[c] = temp[r * nbox + c] * sK.wt[c];
                sd[r] = np_sum(work, nbox);
            }
            for (int r = 0; r < nbox; r++) v[r] = sK.wt[r] * sK.sgx[r] * sd[r];
            const double sumgd = np_sum(v, nbox);
            for (int r = 0; r < nbox; r++) v[r] = sK.wt[r] * sd[r];
            const double sumd = np_sum(v, nbox);
            for (int r = 0; r < nbox; r++) v[r] = sK.wt[r] * sd[r] * sK.dgdy[r];
            const double sddgdy = np_sum(v, nbox);
            const double hy = (sumgd - sK.sumgy * sumd / sK.p) / (sK.sumgsqx - sK.sumgy * sK.sumgy / sK.p);
            if (hy <= 0.) verdict = 3;
            else {
                const double skylvl = (sumd - hy * sK.sumgy) / sK.p;
                ddy = (sK.sgdgdy - (sddgdy - sK.sdgdy * (hy * sK.sumgy + skylvl * sK.p))) / (hy * sK.sdgdys / sK.sigsq);
                if (fabs(ddy) >= nhalf) verdict = 3;
            }
        }
        P.cok[i] = verdict == 0;
        if (verdict) atomicAdd(&P.counts[verdict], 1);
        else {
            atomicAdd(&P.counts[4], 1);
            P.cx[i] = ix + ddx; P.cy[i] = iy + ddy; P.cflux[i] = d; P.csharp[i] = sharp1; P.cround[i] = around;
        }
    }
}

__global__ void find_gather_kernel(const FindParams P) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P.nsel; i += gridDim.x * blockDim.x) {
        const int j = P.sel[i];
        P.ox[i] = P.cx[j]; P.oy[i] = P.cy[j]; P.oflux[i] = P.cflux[j]; P.osharp[i] = P.csharp[j]; P.oround[i] = P.cround[j];
    }
}

// ------------------------------------------------------------------ cntrd
// cntrd.cntrd (PythonPhot/cntrd.py:84-231): derivative-search centroid around approximate positions, one thread
// per star (the boxes are a few dozen pixels).  Status codes are BPHOT_CNTRD_* of include/bphot.h.
constexpr int CNTRD_MAXBOX = 63;  // nbox = 2*max(int(0.637 fwhm), 2) + 1
struct CntrdConsts {  // cntrd.py:91-100, 165-171: everything that depends on fwhm alone (host-computed)
    int nhalf, nbox, ir, ex, ey, keepcenter;
    double w[CNTRD_MAXBOX], dd[CNTRD_MAXBOX], wdd[CNTRD_MAXBOX];  // weights, offsets, w*dd
    double sumc, sumxsq;
};
struct CntrdParams {
    const void* img;
    int nrow, ncol;
    long long pitch;
    const double* x;
    const double* y;
    int npts;
    const CntrdConsts* K;
    double* xcen;
    double* ycen;
    int* status;
};

template <typename T>
__global__ void __launch_bounds__(128) cntrd_kernel(const CntrdParams P) {
    __shared__ CntrdConsts sK;
    {
        const int* src = reinterpret_cast<const int*>(P.K);
        int* dst = reinterpret_cast<int*>(&sK);
        for (int i = threadIdx.x; i < (int)(sizeof(CntrdConsts) / 4); i += 128) dst[i] = src[i];
    }
    __syncthreads();
    const int nhalf = sK.nhalf, nbox = sK.nbox, ir = sK.ir, nd = nbox - 1;
    const T* __restrict__ img = reinterpret_cast<const T*>(P.img);
    double d[CNTRD_MAXBOX], v[CNTRD_MAXBOX];
    for (int i = blockIdx.x * 128 + threadIdx.x; i < P.npts; i += gridDim.x * 128) {
        const double x = P.x[i], y = P.y[i];
        const double ix = rint(x), iy = rint(y);  // np.round: half to even (cntrd.py:107-108)
        auto fail = [&](int code) { P.xcen[i] = -1.0; P.ycen[i] = -1.0; P.status[i] = code; };
        double xmax = ix, ymax = iy;
        if (!sK.keepcenter) {  // cntrd.py:112-141: brightest pixel of the (extended) search box
            const int hx = nhalf + sK.ex, hy = nhalf + sK.ey;
            if (!(ix >= hx) || !(ix + hx <= P.ncol - 1) || !(iy >= hy) || !(iy + hy <= P.nrow - 1)) { fail(1); continue; }
            const T* __restrict__ big = img + (size_t)((long long)iy - hy) * P.pitch + ((long long)ix - hx);
            double mx = 0.0;
            bool any = false;
            for (int r = 0; r <= 2 * hy; r++)
                for (int c = 0; c <= 2 * hx; c++) {
                    const double val = (double)big[(size_t)r * P.pitch + c];
                    if (val == val && (!any || val > mx)) { mx = val; any = true; }
                }
            if (!any) { fail(7); continue; }  // the reference raises on an all-NaN box
            long long sx = 0, sy = 0;
            int nmax = 0;
            for (int r = 0; r <= 2 * hy; r++)
                for (int c = 0; c <= 2 * hx; c++)
                    if ((double)big[(size_t)r * P.pitch + c] == mx) { sx += c; sy += r; nmax++; }
            const double jx = nmax > 1 ? rint((double)sx / nmax) : (double)sx;  // cntrd.py:133-138
            const double jy = nmax > 1 ? rint((double)sy / nmax) : (double)sy;
            xmax = ix - hx + jx;
            ymax = iy - hy + jy;
        }
        if (!(xmax >= nhalf) || !(xmax + nhalf <= P.ncol - 1) || !(ymax >= nhalf) || !(ymax + nhalf <= P.nrow - 1)) { fail(2); continue; }
        const T* __restrict__ box = img + (size_t)((long long)ymax - nhalf) * P.pitch + ((long long)xmax - nhalf);
        auto at = [&](int r, int c) -> double { return (double)box[(size_t)r * P.pitch + c]; };
        // X: first differences along x, summed over the 2*ir+1 central rows in row order (cntrd.py:175-177)
        for (int c = 0; c < nd; c++) {
            double s = at(nhalf - ir, c + 1) - at(nhalf - ir, c);
            for (int r = nhalf - ir + 1; r <= nhalf + ir; r++) s = s + (at(r, c + 1) - at(r, c));
            d[c] = s;
        }
        for (int c = 0; c < nd; c++) v[c] = sK.w[c] * d[c];
        double sumd = np_sum(v, nd);
        for (int c = 0; c < nd; c++) v[c] = sK.wdd[c] * d[c];  // (w*dd)*deriv, cntrd.py:179
        double sumxd = np_sum(v, nd);
        if (sumxd >= 0) { fail(3); continue; }
        const double dx = sK.sumxsq * sumd / (sK.sumc * sumxd);
        if (fabs(dx) > nhalf) { fail(4); continue; }
        // Y (cntrd.py:196-211)
        for (int r = 0; r < nd; r++) {  // np.sum(deriv, 1): pairwise along each row of 2*ir+1 differences
            for (int c = 0; c <= 2 * ir; c++) v[c] = at(r + 1, nhalf - ir + c) - at(r, nhalf - ir + c);
            d[r] = np_sum(v, 2 * ir + 1);
        }
        for (int r = 0; r < nd; r++) v[r] = sK.w[r] * d[r];
        sumd = np_sum(v, nd);
        for (int r = 0; r < nd; r++) v[r] = sK.w[r] * d[r] * sK.dd[r];  // (w*deriv)*dd, cntrd.py:200
        sumxd = np_sum(v, nd);
        if (sumxd >= 0) { fail(5); continue; }
        const double dy = sK.sumxsq * sumd / (sK.sumc * sumxd);
        if (fabs(dy) > nhalf) { fail(6); continue; }
        P.xcen[i] = xmax - dx;
        P.ycen[i] = ymax - dy;
        P.status[i] = 0;
    }
}

}  // namespace bphot
