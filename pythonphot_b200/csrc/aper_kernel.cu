// aper_kernel.cu -- concentric-aperture photometry + mmm sky, one CTA per star (sm_100a).
//
// Replaces the per-star Python loop of PythonPhot/aper.py:217-417 and its callees
// mmm.mmm (mmm.py:7-314) and pixwt.Pixwt (pixwt.py:8-235).  Per star, one 256-thread CTA
//   P1  gathers the sky-annulus pixels of the (2*skyrad_out+1)^2 stamp into shared memory
//       (row-wise, coalesced, candidate column span per row; aper.py:240-267),
//   P2  sorts them in shared memory (bitonic network; np.sort at mmm.py:119),
//   P3  warp 0 runs the mmm clipping loop on the sorted vector (binary-searched brackets in
//       place of the index walk, fp64 range sums by warp shuffle) WHILE warps 1..7 accumulate
//       the aperture sums (exact circle/pixel overlap weights or the polygon approximation),
//   P4  combines both into flux / error / magnitude (aper.py:383-403).
// A persistent grid strides over the star list.
#include "bphot_device.cuh"
#include "bphot_params.h"

namespace bphot {

constexpr int APER_THREADS = 256;
constexpr int APER_WARPS = APER_THREADS / 32;
constexpr int APER_WORKERS = APER_WARPS - 1;  // warps that do aperture sums during mmm

// ----------------------------------------------------------------- sort
// Bitonic sorting network over m = 2^k keys in shared memory (ascending).
template <typename T>
BP_D void cta_bitonic_sort(T* keys, int m, int tid) {
    for (int k = 2; k <= m; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < (m >> 1); t += APER_THREADS) {
                const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // lower index of the pair
                const int p = i | j;
                const bool up = (i & k) == 0;
                const T a = keys[i], b = keys[p];
                if ((a > b) == up) { keys[i] = b; keys[p] = a; }
            }
            __syncthreads();
        }
    }
}

// ------------------------------------------------------------------ mmm
// mmm.py:106-314 on an already sorted, NaN-free vector, executed by ONE warp: the control code is
// mmm_sorted() of bphot_math.cuh (all lanes carry identical scalars); the range sums are
// lane-strided fp64 partials combined by a butterfly, so every lane sees the same bits.
template <typename T>
BP_D MmmResult mmm_warp(const T* sky, int nsky, double readnoise, int minsky, int mxiter, int lane) {
    auto at = [sky](int i) -> double { return (double)sky[i]; };
    auto delta_sums = [&](int lo, int hi, double skymid, double* s1, double* s2) {
        double a = 0.0, q = 0.0;
        for (int i = lo + lane; i < hi; i += 32) {
            const double d = at(i) - skymid;
            a = a + d;
            q = q + d * d;
        }
        *s1 = warp_sum(a);
        *s2 = warp_sum(q);
    };
    auto plain_sum = [&](int lo, int hi) -> double {
        double a = 0.0;
        for (int i = lo + lane; i < hi; i += 32) a = a + at(i);
        return warp_sum(a);
    };
    return mmm_sorted(at, delta_sums, plain_sum, nsky, readnoise, minsky, mxiter);
}

// ----------------------------------------------------------------- aper
struct ApPartial {  // one worker warp's share of one aperture
    double sfull;   // sum of pixel values with weight exactly 1 (approx mode) / sum v*w (exact mode)
    double spart;   // approx mode: sum v*f over partial pixels
    double wpart;   // approx mode: sum f over partial pixels
    int nfull, npart, flags;  // flags: bit0 non-finite pixel seen, bit1 bad-range pixel seen
};

template <typename T>
__global__ void __launch_bounds__(APER_THREADS) aper_kernel(const AperParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* keys = reinterpret_cast<T*>(smem_raw);
    __shared__ int s_cnt[2];
    __shared__ ApPartial s_part[BPHOT_MAX_APER][APER_WORKERS];
    __shared__ MmmResult s_mmm;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const T* __restrict__ img = reinterpret_cast<const T*>(P.img);
    const int naper = P.naper;
    const double rinsq = P.skyrad_in * P.skyrad_in, routsq = P.skyrad_out * P.skyrad_out;

    for (int s = blockIdx.x; s < P.nstars; s += gridDim.x) {
        const double x = P.xc[s], y = P.yc[s];
        // aper.py:209-210 (a NaN position crashes the reference; it is flagged bad here)
        const bool badstar = !(x >= 0.5) || !(x <= P.ncol - 1.5) || !(y >= 0.5) || !(y <= P.nrow - 1.5);
        auto write_bad = [&](double skyv, double skyerrv, int nskyv, int iters) {
            for (int k = tid; k < naper; k += APER_THREADS) {
                const size_t o = (size_t)s * naper + k;
                P.mag[o] = nan64(); P.magerr[o] = nan64(); P.flux[o] = nan64(); P.fluxerr[o] = nan64();
                P.badflag[o] = 1;
            }
            if (tid == 0) { P.sky[s] = skyv; P.skyerr[s] = skyerrv; P.nsky[s] = nskyv; P.mmm_iter[s] = iters; }
        };
        if (badstar) { write_bad(0.0, 0.0, 0, 0); continue; }

        // aper.py:188-208: stamp limits (int truncation, clamp to the frame), centre in stamp coords
        long long lx = (long long)(x - P.skyrad_out), ly = (long long)(y - P.skyrad_out);
        long long ux = (long long)(x + P.skyrad_out), uy = (long long)(y + P.skyrad_out);
        if (lx < 0) lx = 0;
        if (ux > P.ncol - 1) ux = P.ncol - 1;
        if (ly < 0) ly = 0;
        if (uy > P.nrow - 1) uy = P.nrow - 1;
        const int nx = (int)(ux - lx + 1), ny = (int)(uy - ly + 1);
        const double dx = x - (double)lx, dy = y - (double)ly;
        double edge = dx - 0.5;
        if (nx + 0.5 - dx < edge) edge = nx + 0.5 - dx;
        if (dy - 0.5 < edge) edge = dy - 0.5;
        if (ny + 0.5 - dy < edge) edge = ny + 0.5 - dy;
        const T* __restrict__ stamp = img + (size_t)ly * P.pitch + lx;

        int nsky_all = 0, nvalid = 0;
        if (!P.use_setsky) {
            // ---- P1: gather the annulus (aper.py:240-267)
            if (tid == 0) { s_cnt[0] = 0; s_cnt[1] = 0; }
            __syncthreads();
            for (int r = warp; r < ny; r += APER_WARPS) {
                const double dyv = (double)r - dy, dysq = dyv * dyv;
                const double rem = routsq - dysq;
                if (rem < 0.0) continue;  // rsq >= dysq > routsq for the whole row
                const double hw = sqrt(rem) + 1.0;
                int c0 = (int)floor(dx - hw), c1 = (int)ceil(dx + hw);
                if (c0 < 0) c0 = 0;
                if (c1 > nx - 1) c1 = nx - 1;
                const T* __restrict__ row = stamp + (size_t)r * P.pitch;
                for (int cb = c0; cb <= c1; cb += 32) {
                    const int c = cb + lane;
                    bool sel = false, valid = false;
                    T val = T(0);
                    if (c <= c1) {
                        const double dxv = (double)c - dx;
                        const double rsq = dxv * dxv + dysq;
                        if (rsq >= rinsq && rsq <= routsq) {
                            val = row[c];
                            const double dv = (double)val;
                            sel = !(P.badlo != P.badhi && (dv < P.badlo || dv > P.badhi));
                            valid = sel && (val == val);  // mmm.py:106 drops NaNs; aper's nsky keeps them
                        }
                    }
                    const unsigned m_all = __ballot_sync(FULL, sel), m_val = __ballot_sync(FULL, valid);
                    if (m_all == 0u) continue;
                    int base = 0;
                    if (lane == 0) {
                        atomicAdd(&s_cnt[0], __popc(m_all));
                        base = atomicAdd(&s_cnt[1], __popc(m_val));
                    }
                    base = __shfl_sync(FULL, base, 0);
                    if (valid) {
                        const int idx = base + __popc(m_val & ((1u << lane) - 1u));
                        if (idx < P.sort_cap) keys[idx] = val;
                    }
                }
            }
            __syncthreads();
            nsky_all = s_cnt[0];
            nvalid = s_cnt[1] < P.sort_cap ? s_cnt[1] : P.sort_cap;
            if (nsky_all < P.minsky) {  // aper.py:269-272
                write_bad(0.0, 0.0, nsky_all, 0);
                __syncthreads();  // s_cnt is reset by the next star
                continue;
            }

            // ---- P2: sort (mmm.py:119); pad to a power of two with +inf
            int m = 2;
            while (m < nvalid) m <<= 1;
            for (int i = nvalid + tid; i < m; i += APER_THREADS) keys[i] = KeyTraits<T>::inf();
            __syncthreads();
            cta_bitonic_sort(keys, m, tid);
        }

        // ---- P3: warp 0 -> mmm ; warps 1.. -> aperture sums
        if (warp == 0) {
            if (!P.use_setsky) {
                MmmResult R = mmm_warp(keys, nvalid, P.readnoise, P.minsky, P.mxiter, lane);
                if (lane == 0) s_mmm = R;
            }
        } else {
            const int w = warp - 1;
            for (int k = 0; k < naper; k++) {
                const double ap = P.apr[k];
                ApPartial A;
                A.sfull = 0.0; A.spart = 0.0; A.wpart = 0.0; A.nfull = 0; A.npart = 0; A.flags = 0;
                if (edge >= ap) {
                    const double bigrad = ap + 0.5, smallrad = ap / sqrt(2.) - 0.5;  // aper.py:179-181
                    int c0 = (int)floor(dx - bigrad) - 1, c1 = (int)ceil(dx + bigrad) + 1;
                    int r0 = (int)floor(dy - bigrad) - 2, r1 = (int)ceil(dy + bigrad) + 1;
                    if (c0 < 0) c0 = 0;
                    if (c1 > nx - 1) c1 = nx - 1;
                    if (r0 < 0) r0 = 0;
                    if (r1 > ny - 1) r1 = ny - 1;
                    for (int r = r0 + w; r <= r1; r += APER_WORKERS) {
                        const T* __restrict__ row = stamp + (size_t)r * P.pitch;
                        for (int c = c0 + lane; c <= c1; c += 32) {
                            double wgt;
                            bool full = false;
                            if (P.exact) {  // aper.py:313-333
                                const double xxp = (double)c;
                                const double yyp = P.compat_py2 ? (double)r
                                                                : (double)((long long)r * nx + c) / (double)nx;  // aper.py:246-248
                                const double x1 = fabs(xxp - dx), y1 = fabs(yyp - dy);
                                if (x1 > bigrad || y1 > bigrad) continue;
                                if (x1 < smallrad && y1 < smallrad) wgt = 1.0;
                                else {
                                    wgt = pixwt(dx, dy, ap, xxp, yyp);
                                    if (wgt < 0.0) wgt = 0.0;
                                }
                                if (!(wgt > 0.0)) continue;
                            } else {  // aper.py:336-349
                                const double dxv = (double)c - dx, dyv = (double)r - dy;
                                const double rr = sqrt(dxv * dxv + dyv * dyv) - 0.5;
                                if (!(rr < ap)) continue;
                                wgt = ap - rr;
                                if (wgt > 1.) wgt = 1.;
                                if (wgt < 0.) wgt = 0.;
                                full = (wgt == 1.0);
                            }
                            double v = (double)row[c];
                            if (!isfinite(v)) { A.flags |= 1; v = 0.0; }  // aper.py:359-366
                            if (P.badlo < P.badhi && (v <= P.badlo || v >= P.badhi)) { A.flags |= 2; v = 0.0; }  // aper.py:367-378
                            if (P.exact || full) {
                                A.sfull = A.sfull + v * wgt;
                                A.nfull += 1;
                            } else {
                                A.spart = A.spart + v * wgt;
                                A.wpart = A.wpart + wgt;
                                A.npart += 1;
                            }
                        }
                    }
                    A.sfull = warp_sum(A.sfull);
                    A.spart = warp_sum(A.spart);
                    A.wpart = warp_sum(A.wpart);
                    A.nfull = warp_sum_int(A.nfull);
                    A.npart = warp_sum_int(A.npart);
                    A.flags = __reduce_or_sync(FULL, (unsigned)A.flags);
                }
                if (lane == 0) s_part[k][w] = A;
            }
        }
        __syncthreads();

        // ---- P4: combine (aper.py:290-300, 383-403)
        double skymod, skysig, skyvar, sigsq;
        int iters = 0, nsky_out;
        if (!P.use_setsky) {
            skymod = s_mmm.skymod; skysig = s_mmm.sigma; iters = s_mmm.iters;
            skyvar = skysig * skysig;
            sigsq = skyvar / nsky_all;
            nsky_out = nsky_all;
            if (skysig < 0.0) {  // aper.py:293-296, 411-412
                write_bad(skymod, skysig, nsky_all, iters);
                __syncthreads();
                continue;
            }
            if (skysig > 999.99) skysig = 999;  // aper.py:298
        } else {  // aper.py:302-308
            skymod = P.setsky[0]; skysig = P.setsky[1];
            skyvar = skysig * skysig;
            sigsq = skyvar / P.setsky[2];
            nsky_out = (int)P.setsky[2];
        }
        if (tid < naper) {
            const int k = tid;
            const size_t o = (size_t)s * naper + k;
            const double ap = P.apr[k];
            if (!(edge >= ap)) {  // aper.py:311, 350-353
                P.mag[o] = nan64(); P.magerr[o] = nan64(); P.flux[o] = nan64(); P.fluxerr[o] = nan64();
                P.badflag[o] = 1;
            } else {
                const double area = M_PI * ap * ap;
                double sfull = 0.0, spart = 0.0, wpart = 0.0;
                int nfull = 0, npart = 0, flags = 0;
                for (int w = 0; w < APER_WORKERS; w++) {
                    const ApPartial& A = s_part[k][w];
                    sfull = sfull + A.sfull; spart = spart + A.spart; wpart = wpart + A.wpart;
                    nfull += A.nfull; npart += A.npart; flags |= A.flags;
                }
                double total = sfull;
                if (!P.exact && npart > 0) total = sfull + ((area - nfull) / wpart) * spart;  // aper.py:343-349
                if (flags & 1) total = 0.0;  // fractn = 0 for the whole aperture, aper.py:366
                const double apflux = total - skymod * area;  // aper.py:391
                const double err = sqrt(area * skyvar + fabs(apflux) / P.phpadu + sigsq * (area * area));  // aper.py:394-397
                P.flux[o] = apflux;
                P.fluxerr[o] = err;
                if (apflux > 0.0) {  // aper.py:399-403
                    P.magerr[o] = 1.0857 * err / apflux;
                    P.mag[o] = P.zeropoint - 2.5 * log10(apflux);
                } else {
                    P.magerr[o] = nan64();
                    P.mag[o] = nan64();
                }
                P.badflag[o] = flags ? 1 : 0;
            }
        }
        if (tid == 0) { P.sky[s] = skymod; P.skyerr[s] = skysig; P.nsky[s] = nsky_out; P.mmm_iter[s] = iters; }
        __syncthreads();  // keys / s_part / s_mmm are reused by the next star
    }
}

template __global__ void aper_kernel<float>(const AperParams);
template __global__ void aper_kernel<double>(const AperParams);

// ---------------------------------------------------- unit-test kernels
// Batched mmm over CSR-packed vectors (row a4 of the scope table): one CTA per vector.
template <typename T>
__global__ void __launch_bounds__(APER_THREADS) mmm_kernel(const T* vec, const long long* offs, int nvec, double readnoise,
                                                           int minsky, int mxiter, int sort_cap, double* out3, int* iters) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* keys = reinterpret_cast<T*>(smem_raw);
    __shared__ int s_n;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int v = blockIdx.x; v < nvec; v += gridDim.x) {
        const long long b = offs[v], e = offs[v + 1];
        if (tid == 0) s_n = 0;
        __syncthreads();
        for (long long ib = b + (long long)warp * 32; ib < e; ib += APER_THREADS) {
            const long long i = ib + lane;
            T val = T(0);
            bool ok = false;
            if (i < e) { val = vec[i]; ok = (val == val); }
            const unsigned mk = __ballot_sync(FULL, ok);
            int base = 0;
            if (lane == 0 && mk) base = atomicAdd(&s_n, __popc(mk));
            base = __shfl_sync(FULL, base, 0);
            if (ok) {
                const int idx = base + __popc(mk & ((1u << lane) - 1u));
                if (idx < sort_cap) keys[idx] = val;
            }
        }
        __syncthreads();
        const int n = s_n < sort_cap ? s_n : sort_cap;
        int m = 2;
        while (m < n) m <<= 1;
        for (int i = n + tid; i < m; i += APER_THREADS) keys[i] = KeyTraits<T>::inf();
        __syncthreads();
        cta_bitonic_sort(keys, m, tid);
        if (warp == 0) {
            MmmResult R = mmm_warp(keys, n, readnoise, minsky, mxiter, lane);
            if (lane == 0) {
                out3[3 * (size_t)v + 0] = R.skymod; out3[3 * (size_t)v + 1] = R.sigma; out3[3 * (size_t)v + 2] = R.skew;
                iters[v] = R.iters;
            }
        }
        __syncthreads();
    }
}
template __global__ void mmm_kernel<float>(const float*, const long long*, int, double, int, int, int, double*, int*);
template __global__ void mmm_kernel<double>(const double*, const long long*, int, double, int, int, int, double*, int*);

// Pixwt(xc, yc, r, x[], y[]) -- pixwt.py:8 (row a5).
__global__ void pixwt_kernel(double xc, double yc, double r, const double* x, const double* y, long long n, double* out) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = pixwt(xc, yc, r, x[i], y[i]);
}

}  // namespace bphot
