// aper_kernel.cu -- concentric-aperture photometry + mmm sky (sm_100a).
//
// Replaces the per-star Python loop of PythonPhot/aper.py:217-417 and its callees
// mmm.mmm (mmm.py:7-314) and pixwt.Pixwt (pixwt.py:8-235), in three kernels.
// Stage 1, one 128-thread CTA per star:
//   P1  the four warps gather the sky-annulus pixels of the (2*skyrad_out+1)^2 stamp into shared
//       memory, row by row, visiting exactly the two column spans of each row that lie in the
//       annulus (coalesced, every pixel has its own key slot; aper.py:240-267);
//   P2  the keys are sorted (np.sort at mmm.py:119) by a bitonic network that lives in REGISTERS:
//       each thread holds K consecutive keys, compare-exchanges between a thread's own keys are
//       FMNMX pairs, exchanges inside a warp are shuffles, and only the 3 (K=32) stages that cross
//       warps go through shared memory; the sorted vector goes to a global scratch row.
// Aperture sums, one WARP per star: pixels are classified geometrically (inside / outside / crossed
// by the circle) and only crossed pixels go through pixwt, four lanes per pixel (one per side).
// Stage 2, one WARP per star: the mmm clipping loop on the sorted row (32-way ballot searches in
// place of the reference's index walk, fp64 range sums by warp shuffle), then flux / error /
// magnitude (aper.py:383-403).
// K = 0 selects the fallback for annuli of more than 4096 pixels: shared-memory bitonic sort.
#include "bphot_device.cuh"
#include "bphot_params.h"

namespace bphot {

constexpr int AP_THREADS = 128;
constexpr int AP_WARPS = AP_THREADS / 32;
#ifndef AP_S1_CTAS
#define AP_S1_CTAS 6  // resident stage-1 CTAs per SM the register budget is compiled for (K = 32, fp32 keys)
#endif
#ifndef AP_S1_CTAS_F64
#define AP_S1_CTAS_F64 1  // fp64 keys: measured 220 ms per 1 M stars at 255 registers (1), 250 at 3 CTAs, 260 at 4 -- not occupancy-bound
#endif
#ifndef AP_PREFETCH
#define AP_PREFETCH 0  // aperture sums: prefetch the aperture box at the start of the star; measured 10.3 -> 11.1 ms, off
#endif
#ifndef AP_RECIP
#define AP_RECIP 1  // aperture sums: per-pixel divisions by per-star constants as multiplications (values move by an ulp; fluxes are held to 1e-9)
#endif
#ifndef AP_GATHER_U
#define AP_GATHER_U 4
#endif
#ifndef AP_DYNAMIC
#define AP_DYNAMIC 1  // 1 = stage-1 CTAs / stage-2 warps draw stars from an atomic counter; 0 = static grid stride
#endif
#ifndef AP_S2_MINB
#define AP_S2_MINB 9  // minimum resident stage-2 CTAs per SM the register budget is compiled for
#endif
#ifndef AP_SUMS_CTAS
#define AP_SUMS_CTAS 8  // 6 -> 8 (64 registers): aper_sums 10.86 -> 10.32 ms per 1 M stars
#endif

BP_D float kmin(float a, float b) { return fminf(a, b); }
BP_D float kmax(float a, float b) { return fmaxf(a, b); }
BP_D double kmin(double a, double b) { return fmin(a, b); }
BP_D double kmax(double a, double b) { return fmax(a, b); }

// Compare-exchange against a value that came from another thread: the lower index of the pair keeps the minimum.
// ptxas has no single-instruction form for "min or max by predicate": `p ? fminf : fmaxf` becomes FMNMX into a fresh
// register, a predicated FMNMX over it and an IMAD.MOV back (ncu, v8: 3 939 of stage 1's 28 800 warp-instructions per
// star are such moves).  AP_ASM_SELECT=1 (default) writes the two predicated min/max in place (no move): stage 1
// 37.45 -> 33.59 ms per 1 M stars, parity suite green (profiles/r02_asmselect_ab.txt).
#ifndef AP_ASM_SELECT
#define AP_ASM_SELECT 1
#endif
#if AP_ASM_SELECT
BP_D float ksel(float a, float b, bool keepmin) {
    asm("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %2, 0;\n\t@p min.f32 %0, %0, %1;\n\t@!p max.f32 %0, %0, %1;\n\t}"
        : "+f"(a) : "f"(b), "r"((int)keepmin));
    return a;
}
BP_D double ksel(double a, double b, bool keepmin) { return keepmin ? kmin(a, b) : kmax(a, b); }
#define AP_KSEL(dst, other, keepmin) dst = ksel(dst, other, keepmin)
#else
#define AP_KSEL(dst, other, keepmin) dst = keepmin ? kmin(dst, other) : kmax(dst, other)
#endif

// ----------------------------------------------------------------- sort
// Element e of the 128*K-key vector lives in register e % K of thread e / K.  The network is the
// all-ascending form of bitonic sort: phase k first pairs e with its mirror e ^ (k-1) inside each
// k-block, then half-cleans with partners e ^ j, j = k/4 .. 1; the lower index keeps the minimum.
template <typename T>
BP_D void ce(T& a, T& b) {
    const T lo = kmin(a, b), hi = kmax(a, b);
    a = lo;
    b = hi;
}
template <typename T, int K, int J>
BP_D void half_clean(T (&r)[K]) {
    if constexpr (J >= 1) {
#pragma unroll
        for (int q = 0; q < K; q++)
            if ((q & J) == 0) ce(r[q], r[q ^ J]);
        half_clean<T, K, J / 2>(r);
    }
}
template <typename T, int K, int KK>
BP_D void local_phases(T (&r)[K]) {  // phases k = KK .. K, entirely inside one thread
    if constexpr (KK <= K) {
#pragma unroll
        for (int q = 0; q < K; q++)
            if (q < (q ^ (KK - 1))) ce(r[q], r[q ^ (KK - 1)]);
        half_clean<T, K, KK / 4>(r);
        local_phases<T, K, KK * 2>(r);
    }
}
// partner thread tid ^ m inside the warp; mirror stages pair register q with the partner's K-1-q
template <typename T, int K>
BP_D void shfl_stage(T (&r)[K], int m, bool keepmin) {
#pragma unroll
    for (int q = 0; q < K; q++) {
        const T o = __shfl_xor_sync(FULL, r[q], m);
        AP_KSEL(r[q], o, keepmin);
    }
}
template <typename T, int K>
BP_D void shfl_mirror(T (&r)[K], int m, bool keepmin) {
#pragma unroll
    for (int q = 0; q < K / 2; q++) {
        const T oa = __shfl_xor_sync(FULL, r[K - 1 - q], m);
        const T ob = __shfl_xor_sync(FULL, r[q], m);
        AP_KSEL(r[q], oa, keepmin);
        AP_KSEL(r[K - 1 - q], ob, keepmin);
    }
}
// AP_SIGNED_SHFL (fp32 keys): the min/max pipe is what bounds the sort (FMNMX issues every other cycle per scheduler; 78
// stages x 32 keys), and a cross-lane compare-exchange written as two predicated FMNMX costs two of them.  Instead a thread
// whose side of the exchange keeps the MAXIMUM holds its keys NEGATED for that stage: both sides then execute the same
// r = min(r, -partner) -- one FMNMX, the negation is an operand modifier -- because min(-a, -b) = -max(a, b).  Between stages
// every register is multiplied by +-1 (exact; FMUL runs on the other pipe) to move to the next stage's sign pattern, and once
// more before the in-thread stages, which need plain keys.  `s` = +1 / -1: this lane's sign change on entry.
// Measured on B200 (1 M stars): stage 1 36.96 ms with, 36.61 ms without -- the FMNMX count drops (static 994 -> 862) but the
// instruction count does not (FMUL + SHFL + FMNMX per key instead of SHFL + 2 FMNMX, plus one scaling pass per phase): the sort
// is bound by issue slots, not by the min/max pipe.  Parity suite green with it; the knob stays off.
#ifndef AP_SIGNED_SHFL
#define AP_SIGNED_SHFL 0
#endif
#ifndef AP_SMEM_UNIFORM
#define AP_SMEM_UNIFORM 1  // shared-memory stages: one FMNMX per key behind a warp-uniform branch instead of two predicated ones
#endif
template <int K>
BP_D void shfl_stage_signed(float (&r)[K], int m, float s) {
#pragma unroll
    for (int q = 0; q < K; q++) {
        const float v = r[q] * s;
        const float o = __shfl_xor_sync(FULL, v, m);
        r[q] = fminf(v, -o);
    }
}
template <int K>
BP_D void shfl_mirror_signed(float (&r)[K], int m, float s) {
#pragma unroll
    for (int q = 0; q < K / 2; q++) {
        const float va = r[q] * s, vb = r[K - 1 - q] * s;
        const float oa = __shfl_xor_sync(FULL, vb, m);
        const float ob = __shfl_xor_sync(FULL, va, m);
        r[q] = fminf(va, -oa);
        r[K - 1 - q] = fminf(vb, -ob);
    }
}
template <int K>
BP_D void scale_keys(float (&r)[K], float s) {
#pragma unroll
    for (int q = 0; q < K; q++) r[q] = r[q] * s;
}

// partner thread in another warp: through shared memory, row stride K+1 keeps both the writes and
// the partner reads conflict-free.  (Measured dead end, round 2: letting only the two partner warps meet at a named barrier
// `bar.sync id, 64` instead of the CTA barrier made stage 1 slower, 36.5 -> 46.8 ms per 1 M stars, and hung the unit kernel.)  `doit` is warp-uniform: false for warps that hold only padding, or whose
// partner warp does (then the exchange cannot change anything); every thread still reaches the barriers.
template <typename T, int K>
BP_D void smem_stage(T (&r)[K], int m, bool mirror, bool keepmin, T* xbuf, int tid, bool doit) {
    T* mine = xbuf + tid * (K + 1);
    if (doit) {
#pragma unroll
        for (int q = 0; q < K; q++) mine[q] = r[q];
    }
    __syncthreads();
    if (doit) {
        const T* theirs = xbuf + (tid ^ m) * (K + 1);
        // m >= 32: keepmin is the same for the whole warp, so one FMNMX per key behind a uniform branch
        if (!AP_SMEM_UNIFORM) {
#pragma unroll
            for (int q = 0; q < K; q++) {
                const T o = theirs[mirror ? K - 1 - q : q];
                AP_KSEL(r[q], o, keepmin);
            }
        } else if (keepmin) {
#pragma unroll
            for (int q = 0; q < K; q++) r[q] = kmin(r[q], theirs[mirror ? K - 1 - q : q]);
        } else {
#pragma unroll
            for (int q = 0; q < K; q++) r[q] = kmax(r[q], theirs[mirror ? K - 1 - q : q]);
        }
    }
    __syncthreads();
}
// `nwarps` = warps that hold real keys.  The network is standard (every comparator leaves the minimum at the lower
// index), so threads above them, which start with +inf in every register, can never receive anything else: their
// warps skip the whole network and their partners skip the exchanges with them.  With skyrad=[40,50] the annulus
// has <= 2840 pixels = 89 threads x 32 keys: three warps sort, the fourth only keeps the barriers.
template <typename T, int K>
BP_D void cta_sort_regs(T (&r)[K], T* xbuf, int tid, int nwarps) {
    const int warp = tid >> 5;
    const bool wact = warp < nwarps;
    if (wact) local_phases<T, K, 2>(r);
#pragma unroll 1
    for (int k = 2 * K; k <= AP_THREADS * K; k <<= 1) {
        const int m = k / K - 1;
        const bool keepmin = (tid & (k / (2 * K))) == 0;
        if constexpr (AP_SIGNED_SHFL && sizeof(T) == 4) {
            float sg = 1.f;  // sign this thread's registers carry (shuffle stages only; shared-memory stages come first, on plain keys)
            if (m < 32) {
                sg = keepmin ? 1.f : -1.f;
                if (wact) shfl_mirror_signed<K>(r, m, sg);
            } else smem_stage<T, K>(r, m, true, keepmin, xbuf, tid, wact && (warp ^ (m >> 5)) < nwarps);
#pragma unroll 1
            for (int j = k >> 2; j >= K; j >>= 1) {
                const int mm = j / K;
                const bool km = (tid & mm) == 0;
                if (mm < 32) {
                    const float want = km ? 1.f : -1.f;
                    if (wact) shfl_stage_signed<K>(r, mm, sg * want);
                    sg = want;
                } else smem_stage<T, K>(r, mm, false, km, xbuf, tid, wact && (warp ^ (mm >> 5)) < nwarps);
            }
            if (wact) { scale_keys<K>(r, sg); half_clean<T, K, K / 2>(r); }
            continue;
        }
        if (m < 32) { if (wact) shfl_mirror<T, K>(r, m, keepmin); }
        else smem_stage<T, K>(r, m, true, keepmin, xbuf, tid, wact && (warp ^ (m >> 5)) < nwarps);
#pragma unroll 1
        for (int j = k >> 2; j >= K; j >>= 1) {
            const int mm = j / K;
            const bool km = (tid & mm) == 0;
            if (mm < 32) { if (wact) shfl_stage<T, K>(r, mm, km); }
            else smem_stage<T, K>(r, mm, false, km, xbuf, tid, wact && (warp ^ (mm >> 5)) < nwarps);
        }
        if (wact) half_clean<T, K, K / 2>(r);
    }
}

// Fallback: bitonic sorting network over m = 2^k keys in shared memory (ascending).
template <typename T>
BP_D void cta_bitonic_sort(T* keys, int m, int tid) {
    for (int k = 2; k <= m; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < (m >> 1); t += (int)blockDim.x) {
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
// Sorted keys in shared memory; element i sits at i + (i >> shift) (row padding of the register
// sort's write-back; shift = 31 means unpadded).
template <typename T>
struct SortedView {
    const T* p;
    int shift;
    BP_D double operator()(int i) const { return (double)p[i + (i >> shift)]; }
};
// Sorted keys in global memory (written by an earlier launch, read-only here).
template <typename T>
struct GlobalView {
    const T* __restrict__ p;
    BP_D double operator()(int i) const { return (double)__ldg(p + i); }
};

// First index in [0, n) whose element satisfies the monotone predicate (n if none): 32 probes
// per round instead of one, so 4096 keys take 3 dependent shared-memory reads, not 12.
template <typename Pred>
BP_D int warp_first_true(const Pred& pred, int n, int lane) {
    int lo = 0, hi = n;
    while (hi - lo > 32) {
        const int step = (hi - lo + 31) >> 5;
        int idx = lo + (lane + 1) * step - 1;
        if (idx > hi - 1) idx = hi - 1;
        const unsigned b = __ballot_sync(FULL, pred(idx));
        if (b == 0u) return hi;
        const int L = __ffs(b) - 1;
        int nhi = lo + (L + 1) * step - 1;
        if (nhi > hi - 1) nhi = hi - 1;
        lo = lo + L * step;
        hi = nhi;  // pred(hi) is known true
    }
    const int idx = lo + lane;
    const unsigned b = __ballot_sync(FULL, idx < hi && pred(idx));
    return b ? lo + __ffs(b) - 1 : hi;
}

// mmm.py:106-314 on an already sorted, NaN-free vector, executed by ONE warp: the control code is
// mmm_sorted() of bphot_math.cuh (all lanes carry identical scalars); the range sums are
// lane-strided fp64 partials combined by a butterfly, so every lane sees the same bits.
template <typename View>
BP_D MmmResult mmm_warp(const View at, int nsky, double readnoise, int minsky, int mxiter, int lane) {
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
    auto first = [&](double v, bool strict) -> int {
        return warp_first_true([&](int i) { const double a = at(i); return strict ? a > v : a >= v; }, nsky, lane);
    };
    auto first2 = [&](double v1, bool st1, bool need1, double v2, bool st2, bool need2, int* r1, int* r2) {
        if (need1) *r1 = first(v1, st1);  // need* are warp-uniform
        if (need2) *r2 = first(v2, st2);
    };
    auto sums3 = [&](int lo0, int hi0, bool need0, int lo1, int hi1, bool need1, int lo2, int hi2, double mid, double* o5) {
        if (need0) delta_sums(lo0, hi0, mid, &o5[0], &o5[1]);  // need* are warp-uniform
        if (need1) delta_sums(lo1, hi1, mid, &o5[2], &o5[3]);
        o5[4] = plain_sum(lo2, hi2);
    };
    return mmm_sorted(at, delta_sums, plain_sum, first2, sums3, nsky, readnoise, minsky, mxiter);
}

// scipy.stats.sigmaclip(skybuf, low, high) followed by np.mean / np.std of the survivors (aper.py:275-282), by one
// warp on the SORTED annulus: the survivors of every pass are a contiguous index range, found with the same 32-way
// searches as mmm's brackets; mean and the two-pass variance are lane-strided fp64 sums + butterfly.
template <typename View>
BP_D MmmResult sigclip_warp(const View at, int n, double low, double high, int lane) {
    MmmResult R;
    R.skymod = nan64(); R.sigma = nan64(); R.skew = 0.0; R.iters = 0;
    int lo = 0, hi = n;  // survivors: [lo, hi)
    while (hi > lo) {
        R.iters++;
        const int cnt = hi - lo;
        double a = 0.0;
        for (int i = lo + lane; i < hi; i += 32) a = a + at(i);
        const double mean = warp_sum(a) / cnt;
        double q = 0.0;
        for (int i = lo + lane; i < hi; i += 32) { const double d = at(i) - mean; q = q + d * d; }
        const double sd = sqrt(warp_sum(q) / cnt);
        R.skymod = mean; R.sigma = sd;
        const double critlo = mean - sd * low, crithi = mean + sd * high;
        // first index with x >= critlo, first index with x > crithi, both inside [lo, hi)
        const int nlo = lo + warp_first_true([&](int i) { return at(lo + i) >= critlo; }, cnt, lane);
        const int nhi = lo + warp_first_true([&](int i) { return at(lo + i) > crithi; }, cnt, lane);
        const int delta = cnt - (nhi - nlo);
        lo = nlo; hi = nhi;
        if (delta == 0) break;
        if (hi <= lo) { R.skymod = nan64(); R.sigma = nan64(); }  // nothing survived: mean of an empty array
    }
    return R;
}

// ----------------------------------------------------------------- aper
struct ApPartial {  // one worker warp's share of one aperture
    double sfull;   // sum of pixel values with weight exactly 1 (approx mode) / sum v*w (exact mode)
    double spart;   // approx mode: sum v*f over partial pixels
    double wpart;   // approx mode: sum f over partial pixels
    int nfull, npart, flags;  // flags: bit0 non-finite pixel seen, bit1 bad-range pixel seen
};

struct StarGeom {  // aper.py:188-208
    int nx, ny;
    double dx, dy, edge;
    size_t origin;  // element offset of the stamp's first pixel
};

BP_D StarGeom star_geom(const AperParams& P, double x, double y) {
    long long lx = (long long)(x - P.skyrad_out), ly = (long long)(y - P.skyrad_out);  // int truncation, then clamp
    long long ux = (long long)(x + P.skyrad_out), uy = (long long)(y + P.skyrad_out);
    if (lx < 0) lx = 0;
    if (ux > P.ncol - 1) ux = P.ncol - 1;
    if (ly < 0) ly = 0;
    if (uy > P.nrow - 1) uy = P.nrow - 1;
    StarGeom g;
    g.nx = (int)(ux - lx + 1);
    g.ny = (int)(uy - ly + 1);
    g.dx = x - (double)lx;
    g.dy = y - (double)ly;
    double edge = g.dx - 0.5;
    if (g.nx + 0.5 - g.dx < edge) edge = g.nx + 0.5 - g.dx;
    if (g.dy - 0.5 < edge) edge = g.dy - 0.5;
    if (g.ny + 0.5 - g.dy < edge) edge = g.ny + 0.5 - g.dy;
    g.edge = edge;
    g.origin = (size_t)ly * P.pitch + lx;
    return g;
}

// P1a: EXACT column spans of every stamp row, one row per thread (aper.py:240-267).  In floating point
// rsq(c) = (c-dx)^2 + dysq is monotone in |c-dx| (every operation rounds monotonically), so the columns of row r
// with rsq <= rout^2 form one interval [a0, a1] and those with rsq < rin^2 one interval [b0, b1] inside it; both
// are found from a sqrt estimate and then corrected with the reference's own fp64 test, at most a few steps.
// The annulus pixels of the row are [a0, b0) and (b1, a1]; span[4r..4r+3] = a0, b0, b1, a1 and rowoff[r] = number of
// annulus pixels in the rows above (filled by row_offsets), so pixel t of row r owns key slot rowoff[r] + t.
BP_D void row_spans(const StarGeom& g, double rinsq, double routsq, int* span, int* rowoff, int tid) {
    for (int r = tid; r < g.ny; r += (int)blockDim.x) {
        const double dyv = (double)r - g.dy, dysq = dyv * dyv;
        auto rsq_at = [&](int c) { const double dxv = (double)c - g.dx; return dxv * dxv + dysq; };
        int a0 = 0, a1 = -1, b0 = 0, b1 = -1;
        const double rem = routsq - dysq;
        if (rem >= 0.0) {  // otherwise rsq >= dysq > routsq for the whole row
            const double hw = sqrt(rem);
            int lo = (int)floor(g.dx - hw) - 1, hi = (int)ceil(g.dx + hw) + 1;  // both surely outside
            while (lo <= hi && !(rsq_at(lo) <= routsq)) lo++;
            while (hi >= lo && !(rsq_at(hi) <= routsq)) hi--;
            int e0 = 1, e1 = 0;  // columns strictly inside the inner circle
            const double remi = rinsq - dysq;
            if (remi > 0.0) {
                const double hwi = sqrt(remi);
                e0 = (int)floor(g.dx - hwi) - 1;
                e1 = (int)ceil(g.dx + hwi) + 1;
                while (e0 <= e1 && !(rsq_at(e0) < rinsq)) e0++;
                while (e1 >= e0 && !(rsq_at(e1) < rinsq)) e1--;
            }
            if (lo < 0) lo = 0;
            if (hi > g.nx - 1) hi = g.nx - 1;
            if (lo <= hi) {
                a0 = lo; a1 = hi;
                if (e0 < lo) e0 = lo;
                if (e1 > hi) e1 = hi;
                if (e0 <= e1) { b0 = e0; b1 = e1; }
                else { b0 = hi + 1; b1 = hi; }
            }
        }
        span[4 * r + 0] = a0; span[4 * r + 1] = b0; span[4 * r + 2] = b1; span[4 * r + 3] = a1;
        rowoff[r] = (b0 - a0) + (a1 - b1);
    }
}

// Exclusive prefix sum of the per-row pixel counts, by one warp; returns the total (all lanes).
// Four consecutive rows per lane, so a stamp of up to 128 rows (skyrad[1] <= 62) is ONE warp scan: the other warps of the CTA
// wait at a barrier while this runs, and four dependent scans of 32 rows were 5 % of stage 1's stall samples.
BP_D int row_offsets(int* rowoff, int ny, int lane) {
    int carry = 0;
    for (int base = 0; base < ny; base += 128) {
        const int r = base + 4 * lane;
        const int v0 = r < ny ? rowoff[r] : 0, v1 = r + 1 < ny ? rowoff[r + 1] : 0;
        const int v2 = r + 2 < ny ? rowoff[r + 2] : 0, v3 = r + 3 < ny ? rowoff[r + 3] : 0;
        const int s = (v0 + v1) + (v2 + v3);
        int inc = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int up = __shfl_up_sync(FULL, inc, o);
            if (lane >= o) inc += up;
        }
        const int ex = carry + inc - s;
        if (r < ny) rowoff[r] = ex;
        if (r + 1 < ny) rowoff[r + 1] = ex + v0;
        if (r + 2 < ny) rowoff[r + 2] = ex + v0 + v1;
        if (r + 3 < ny) rowoff[r + 3] = ex + v0 + v1 + v2;
        carry += __shfl_sync(FULL, inc, 31);
    }
    return carry;
}

// P1b: one warp's rows of the annulus gather: coalesced loads along the row, every pixel goes to its own slot, so
// there is no ballot, no atomic and no dependence between loads.  Pixels the reference drops (outside the bad-pixel
// window, aper.py:260-266; NaN, mmm.py:106) leave +inf in their slot -- the sort does not care where the padding
// sits.  Returns, per lane, partial counts: x = pixels rejected by the window, y = selected pixels that are NaN.
template <typename T>
BP_D int2 gather_rows(const AperParams& P, const T* __restrict__ stamp, const StarGeom& g, const int* span, const int* rowoff,
                      T* __restrict__ keys, int warp, int lane) {
    int nrej = 0, nnan = 0;
    const bool window = P.badlo != P.badhi;
    const int nw = (int)(blockDim.x >> 5);  // 4 warps, or 3 when the annulus fits 96 threads x K keys
    constexpr int U = AP_GATHER_U;  // rows in flight per warp: U independent loads per lane before the first use
    for (int r0 = warp; r0 < g.ny; r0 += U * nw) {
        int a0[U], nl[U], skip[U], tot[U], off[U];
        const T* __restrict__ row[U];
        int maxtot = 0;
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int r = r0 + u * nw;
            tot[u] = 0;
            if (r < g.ny) {
                const int4 sp = *reinterpret_cast<const int4*>(span + 4 * r);
                a0[u] = sp.x;
                nl[u] = sp.y - sp.x;
                skip[u] = sp.z + 1 - sp.y;  // width of the excluded inner span
                tot[u] = nl[u] + (sp.w - sp.z);
                off[u] = rowoff[r];
                row[u] = stamp + (size_t)r * P.pitch;
            }
            maxtot = tot[u] > maxtot ? tot[u] : maxtot;
        }
        for (int t = lane; t < maxtot; t += 32) {
            T val[U];
#pragma unroll
            for (int u = 0; u < U; u++)
                if (t < tot[u]) val[u] = row[u][a0[u] + t + (t < nl[u] ? 0 : skip[u])];
#pragma unroll
            for (int u = 0; u < U; u++) {
                if (t < tot[u]) {
                    // aper.py:260-266 in the key type: for fp32 pixels badlo_t / badhi_t are the window rounded INWARD to
                    // fp32 on the host, so val < badlo_t <=> (double)val < badlo exactly, without a conversion per pixel
                    const bool rej = window && (val[u] < (T)(sizeof(T) == 4 ? P.badlo_t : P.badlo) || val[u] > (T)(sizeof(T) == 4 ? P.badhi_t : P.badhi));
                    const bool isnan_ = !rej && !(val[u] == val[u]);
                    nrej += rej;
                    nnan += isnan_;
                    const int idx = off[u] + t;
                    if (idx < P.sort_cap) keys[idx] = (rej || isnan_) ? KeyTraits<T>::inf() : val[u];
                }
            }
        }
    }
    return make_int2(nrej, nnan);
}

// Per-warp list of pixels the aperture circle crosses (they need pixwt): centre, stamp position, weight.
constexpr int AP_LIST = 64;
struct CrossList {
    double x[AP_LIST], y[AP_LIST], w[AP_LIST];
    int rc[AP_LIST];
};

// One pixel value into the running sums, with the reference's bad-pixel bookkeeping (aper.py:359-378).
template <typename T>
BP_D void add_pixel(const AperParams& P, const T* __restrict__ stamp, int r, int c, double wgt, bool full, ApPartial& A) {
    double v = (double)stamp[(size_t)r * P.pitch + c];
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

// Exact overlap weights of the listed pixels: four lanes per pixel, one per Oneside term, recombined in
// Intarea's order of addition (pixwt.py:234-235); then every lane adds its pixels to the sums.
template <typename T>
BP_D void flush_cross_list(const AperParams& P, const T* __restrict__ stamp, const StarGeom& g, double ap,
                           CrossList* L, int cnt, int lane, ApPartial& A) {
    __syncwarp();
    for (int g0 = 0; g0 < cnt; g0 += 8) {
        const int j = g0 + (lane >> 2), side = lane & 3;
        double sd = 0.0;
        if (j < cnt) sd = pixwt_side(g.dx, g.dy, ap, L->x[j], L->y[j], side);
        const double s1 = __shfl_down_sync(FULL, sd, 1), s2 = __shfl_down_sync(FULL, sd, 2), s3 = __shfl_down_sync(FULL, sd, 3);
        if (side == 0 && j < cnt) L->w[j] = ((sd + s1) + s2) + s3;
    }
    __syncwarp();
    for (int i = lane; i < cnt; i += 32) {
        const double wgt = L->w[i];  // negative weights count as 0 (aper.py:329); only positive ones are used (:330)
        if (wgt > 0.0) add_pixel<T>(P, stamp, L->rc[i] >> 16, L->rc[i] & 65535, wgt, false, A);
    }
    __syncwarp();
}

// This warp's share (chunks w, w+nw, ...) of one aperture's pixel sums (aper.py:310-380).
template <typename T>
BP_D ApPartial aperture_warp(const AperParams& P, const T* __restrict__ stamp, const StarGeom& g, double ap,
                             int w, int nw, int lane, CrossList* L) {
    ApPartial A;
    A.sfull = 0.0; A.spart = 0.0; A.wpart = 0.0; A.nfull = 0; A.npart = 0; A.flags = 0;
    const double bigrad = ap + 0.5, smallrad = ap / sqrt(2.) - 0.5;  // aper.py:179-181
    // candidate box: every pixel the reference can give a positive weight lies inside it in either compat mode
    // (py3: yy = row + col/nx shifts a pixel's y by < 1 row, aper.py:246-248)
    int c0 = (int)floor(g.dx - bigrad), c1 = (int)ceil(g.dx + bigrad);
    int r0 = (int)floor(g.dy - bigrad) - (P.exact && !P.compat_py2 ? 1 : 0), r1 = (int)ceil(g.dy + bigrad);
    if (c0 < 0) c0 = 0;
    if (c1 > g.nx - 1) c1 = g.nx - 1;
    if (r0 < 0) r0 = 0;
    if (r1 > g.ny - 1) r1 = g.ny - 1;
    const int bw = c1 - c0 + 1, bh = r1 - r0 + 1;
    const int npx = (bw > 0 && bh > 0) ? bw * bh : 0;
    const unsigned lt = (1u << lane) - 1u;
    int cnt = 0;
    // box position (row rr, column cc) of this lane's pixel, advanced by the loop stride without a division per pixel
    const int step = 32 * nw, bwd = bw > 0 ? bw : 1, krow = step / bwd, kcol = step - krow * bwd;
    int rr = (w * 32 + lane) / bwd, cc = (w * 32 + lane) - rr * bwd;
    const double inv_nx = 1.0 / (double)g.nx;  // AP_RECIP: the Python-3 pixel y-centre (row + col / nx) by one reciprocal per star
    for (int base = w * 32; base < npx; base += step) {
        const int p = base + lane;
        int r = 0, c = 0, cls = 0;
        double xxp = 0., yyp = 0., wgt = 0.;
        bool full = false;
        if (p < npx) {
            r = r0 + rr;
            c = c0 + cc;
            if (P.exact) {  // aper.py:313-333
                xxp = (double)c;
                yyp = P.compat_py2 ? (double)r : AP_RECIP ? (double)(r * g.nx + c) * inv_nx : (double)(r * g.nx + c) / (double)g.nx;  // r * nx + c < 2^31 (stamp <= 20 000 px wide)
                const double x1 = fabs(xxp - g.dx), y1 = fabs(yyp - g.dy);
                if (!(x1 > bigrad || y1 > bigrad)) cls = (x1 < smallrad && y1 < smallrad) ? 1 : pixel_class(x1, y1, ap);
                if (cls == 1) wgt = 1.0;
            } else {  // aper.py:336-349
                const double dxv = (double)c - g.dx, dyv = (double)r - g.dy;
                const double rad = sqrt(dxv * dxv + dyv * dyv) - 0.5;
                if (rad < ap) {
                    wgt = ap - rad;
                    if (wgt > 1.) wgt = 1.;
                    if (wgt < 0.) wgt = 0.;
                    full = (wgt == 1.0);
                    cls = 3;
                }
            }
        }
        const unsigned bmask = __ballot_sync(FULL, cls == 2);
        if (cls == 2) {
            const int q = cnt + __popc(bmask & lt);
            L->x[q] = xxp; L->y[q] = yyp; L->rc[q] = r * 65536 + c;
        }
        cnt += __popc(bmask);
        if (cls == 1 || cls == 3) add_pixel<T>(P, stamp, r, c, wgt, full, A);
        if (cnt > AP_LIST - 32) {
            flush_cross_list<T>(P, stamp, g, ap, L, cnt, lane, A);
            cnt = 0;
        }
        rr += krow; cc += kcol;
        if (cc >= bwd) { cc -= bwd; rr++; }
    }
    if (cnt > 0) flush_cross_list<T>(P, stamp, g, ap, L, cnt, lane, A);
    A.sfull = warp_sum(A.sfull);
    A.spart = warp_sum(A.spart);
    A.wpart = warp_sum(A.wpart);
    A.nfull = warp_sum_int(A.nfull);
    A.npart = warp_sum_int(A.npart);
    A.flags = __reduce_or_sync(FULL, (unsigned)A.flags);
    return A;
}

// ------------------------------------------------- index of the sorted row
// What mmm reads from the sorted annulus is (a) single keys by rank, (b) rank queries "first key > / >= v" and
// (c) sums of (key - skymid) and (key - skymid)^2 over rank ranges (mmm.py:140-145, 217-220, 244-247) plus the
// plain sum of the central 20 % (mmm.py:288).  Stage 1 holds the sorted keys in registers, K consecutive ones per
// thread, so it leaves a small index next to the sorted row (bphot_math.cuh IndexedView): a 16-ary search tree
// (L1[i] = key[256 i], L2[j] = key[16 j]) and, for every group of 16 keys, the EXCLUSIVE prefix sums
// PT[g] = sum over keys [0, 16 g) of (d, d^2), d = key - skymid, skymid = mmm.py:121's median.  With it a range sum is
// two table reads plus two 16-key groups and a rank query three 16-key groups, all independent loads, so mmm becomes
// scalar work that waits for memory a few times per clipping pass: aper_mmm_kernel runs it with one THREAD per
// star (32 stars per warp) instead of one warp per star.
// Index layout per star (idx_stride bytes, 64-byte aligned): T L1[16] | T L2[256] | double2 PT[257].
constexpr size_t sort_index_bytes(size_t key_size) { return ((size_t)(16 + 256) * key_size + (size_t)257 * 16 + 63) & ~(size_t)63; }

#ifndef AP_MED_SMEM
#define AP_MED_SMEM 0  // 1 = the two median keys reach sort_index through shared memory instead of the scratch row; measured: stage 1 36.4 -> 38.0 ms (the 2 x K predicated selects and the divergent branch cost more than the L2 round trip the other CTAs hide)
#endif
#ifndef AP_IDX_SKIP
#define AP_IDX_SKIP 0  // timing experiments only (wrong results): 1 = no per-key sums, 2 = no scan, 4 = no writes, 8 = no median read
#endif
template <typename T, int K>
BP_D void sort_index(const AperParams& P, const T (&r)[K], const T* __restrict__ dst, int nvalid, size_t row, int tid, double (*s_wsum)[2], T* s_med) {
    static_assert(K == 8 || K == 16 || K == 32, "index groups are 16 keys");
    const int lane = tid & 31, warp = tid >> 5, nthr = (int)blockDim.x;
    unsigned char* base = reinterpret_cast<unsigned char*>(P.sidx) + row * (size_t)P.idx_stride;
    T* __restrict__ L1 = reinterpret_cast<T*>(base);
    T* __restrict__ L2 = L1 + 16;
    double2* __restrict__ PT = reinterpret_cast<double2*>(L2 + 256);
    // the two median keys (ranks (n-1)/2 and n/2) sit in registers of one or two threads of this CTA: they go through shared
    // memory (AP_MED_SMEM; reading them back from the scratch row just written cost an L2 round trip with every thread waiting)
    double skymid = 0.0;
#if AP_MED_SMEM
    (void)dst;
    if (nvalid > 0) {
        const int m1 = (nvalid - 1) / 2, m2 = nvalid / 2;
        if (m1 / K == tid || m2 / K == tid) {  // at most two threads, usually one
            const int q1 = m1 / K == tid ? m1 % K : -1, q2 = m2 / K == tid ? m2 % K : -1;
#pragma unroll
            for (int q = 0; q < K; q++) {
                if (q == q1) s_med[0] = r[q];
                if (q == q2) s_med[1] = r[q];
            }
        }
    }
    __syncthreads();
    if (!(AP_IDX_SKIP & 8) && nvalid > 0) skymid = 0.5 * (double)s_med[0] + 0.5 * (double)s_med[1];  // mmm.py:121
#else
    __syncthreads();
    if (!(AP_IDX_SKIP & 8) && nvalid > 0) skymid = 0.5 * (double)__ldcg(dst + (nvalid - 1) / 2) + 0.5 * (double)__ldcg(dst + nvalid / 2);  // mmm.py:121
#endif
    const int e0 = tid * K;
    constexpr int H = K > 16 ? 16 : K;  // keys of this thread in its first index group
    // four independent chains of K/4 keys each (the adds of one chain wait for each other), folded pairwise
    constexpr int C = K / 4;
    double ca[4] = {0.0, 0.0, 0.0, 0.0}, cq[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int q = 0; q < C; q++) {
#pragma unroll
        for (int c = 0; c < 4; c++) {
            if (!(AP_IDX_SKIP & 1) && e0 + c * C + q < nvalid) {
                const double d = bp_widen(r[c * C + q]) - skymid;
                ca[c] = ca[c] + d;
                cq[c] = __fma_rn(d, d, cq[c]);
            }
        }
    }
    const double ha = ca[0] + ca[1], hq = cq[0] + cq[1];  // K = 32: sums of this thread's first group of 16
    const double a = ha + (ca[2] + ca[3]), qq = hq + (cq[2] + cq[3]);
    (void)H;
    // inclusive scan over the threads of the warp, then over the warps
    double ia_ = a, iq_ = qq;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        if (AP_IDX_SKIP & 2) break;
        const double ua = __shfl_up_sync(FULL, ia_, o), uq = __shfl_up_sync(FULL, iq_, o);
        if (lane >= o) { ia_ = ua + ia_; iq_ = uq + iq_; }
    }
    if (lane == 31) { s_wsum[warp][0] = ia_; s_wsum[warp][1] = iq_; }
    double xa = __shfl_up_sync(FULL, ia_, 1), xq = __shfl_up_sync(FULL, iq_, 1);  // exclusive within the warp
    if (lane == 0) { xa = 0.0; xq = 0.0; }
    __syncthreads();
    double ba = 0.0, bq = 0.0;
    for (int w = 0; w < warp; w++) { ba = ba + s_wsum[w][0]; bq = bq + s_wsum[w][1]; }
    const double pa = ba + xa, pq = bq + xq;  // sums over keys [0, e0)
    if (AP_IDX_SKIP & 4) { if (pa == 1.2345) PT[0] = make_double2(pa + ha, pq + hq); return; }
    if constexpr (K == 32) {
        PT[2 * tid] = make_double2(pa, pq);
        PT[2 * tid + 1] = make_double2(pa + ha, pq + hq);
        L2[2 * tid] = r[0];
        L2[2 * tid + 1] = r[16];
        if ((tid & 7) == 0) L1[tid >> 3] = r[0];
    } else if constexpr (K == 16) {
        PT[tid] = make_double2(pa, pq);
        L2[tid] = r[0];
        if ((tid & 15) == 0) L1[tid >> 4] = r[0];
    } else {
        if ((tid & 1) == 0) { PT[tid >> 1] = make_double2(pa, pq); L2[tid >> 1] = r[0]; }
        if ((tid & 31) == 0) L1[tid >> 5] = r[0];
    }
    if (tid == nthr - 1) PT[nthr * K / 16] = make_double2(ba + ia_, bq + iq_);
}

// ------------------------------------------------------------ the two stages
// The per-star work splits into a wide part (gather, sort, aperture sums: thousands of independent
// pixels) and a narrow one (mmm: ~20 dependent clipping iterations on the sorted vector, enough work for
// one warp only).  Run in one CTA, three of four warps idle behind the mmm warp for most of the star's
// lifetime, and registers/shared memory of the sort cap the SM at 5 stars in flight.  So the stages are
// separate launches over batches of stars: stage 1 (CTA per star) leaves each star's sorted annulus in a
// global scratch row (11 KB, mostly L2-resident) and its raw aperture sums in the output arrays; stage 2
// (WARP per star, ~40 stars in flight per SM) runs mmm on that row and finishes flux/error/magnitude.
//
// Stage-1 -> stage-2 hand-over, in place in the outputs: flux[s,k] = sum of weight*pixel (aper.py:380),
// badflag[s,k] = raw flag bits, nsky[s] = annulus pixel count, mmm_iter[s] = number of sorted keys, or
// -1 (star off the frame, aper.py:209-233) / -2 (nsky < minsky, aper.py:269-272).

// K > 0: register sort of 128*K keys (P.sort_cap == 128*K).  K == 0: shared-memory sort of P.sort_cap keys.
template <typename T, int K>
__global__ void __launch_bounds__(AP_THREADS, (K == 32 && sizeof(T) == 4) ? AP_S1_CTAS : (K == 32 ? AP_S1_CTAS_F64 : 1)) aper_stage1_kernel(const AperParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* keys = reinterpret_cast<T*>(smem_raw);
    const size_t key_bytes = ((size_t)(K > 0 ? AP_THREADS * (K + 1) : P.sort_cap) * sizeof(T) + 15) & ~(size_t)15;
    int* span = reinterpret_cast<int*>(smem_raw + key_bytes);
    int* rowoff = span + 4 * P.span_rows;
    __shared__ int s_ntot;
    __shared__ int s_nrej[AP_WARPS], s_nnan[AP_WARPS];
    __shared__ double s_wsum[AP_WARPS][2];
    __shared__ T s_med[2];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = (int)(blockDim.x >> 5);
    const T* __restrict__ img = reinterpret_cast<const T*>(P.img);
    T* __restrict__ skeys = reinterpret_cast<T*>(P.skeys);
    const double rinsq = P.skyrad_in * P.skyrad_in, routsq = P.skyrad_out * P.skyrad_out;

#if AP_DYNAMIC
    // Stars are drawn from a counter instead of a fixed stride: a launch ends when the last star does, not when the
    // unluckiest CTA has worked off its share (static stride left ~19 % of the warp slots idle in the tail of a launch).
    __shared__ int s_next[2];
    for (int it = 0;; it++) {
        if (tid == 0) s_next[it & 1] = atomicAdd(P.work_counter, 1);
        __syncthreads();  // also orders this star's use of keys / tables after the previous star's
        const int i = P.s0 + s_next[it & 1];
        if (i >= P.s0 + P.scount) break;
#else
    for (int i = P.s0 + blockIdx.x; i < P.s0 + P.scount; i += gridDim.x) {
#endif
        const int s = P.order ? P.order[i] : i;  // stars are visited in image-tile order, results land at s
        if (P.retry_wide && P.mmm_iter[s] != -3) continue;  // CTA-uniform
        const double x = P.xc[s], y = P.yc[s];
        // aper.py:209-210 (a NaN position crashes the reference; it is flagged bad here)
        const bool badstar = !(x >= 0.5) || !(x <= P.ncol - 1.5) || !(y >= 0.5) || !(y <= P.nrow - 1.5);
        if (badstar) {
            if (tid == 0) { P.nsky[s] = 0; P.mmm_iter[s] = -1; }
            continue;
        }
        const StarGeom g = star_geom(P, x, y);
        const T* __restrict__ stamp = img + g.origin;

        int nsky_all = 0, nvalid = 0;
        if (!P.use_setsky) {
            // ---- P1: gather the annulus
            row_spans(g, rinsq, routsq, span, rowoff, tid);
            __syncthreads();
            if (warp == 0) {
                const int tot = row_offsets(rowoff, g.ny, lane);
                if (lane == 0) s_ntot = tot;
            }
            __syncthreads();
            if (K > 0 && (int)blockDim.x * K < s_ntot && s_ntot <= P.sort_cap) {
                // the annulus needs the fourth warp of the network: left to the 128-thread pass that follows
                if (tid == 0) P.mmm_iter[s] = -3;
                __syncthreads();
                continue;
            }
            const int2 cnt = gather_rows<T>(P, stamp, g, span, rowoff, keys, warp, lane);
            const int wrej = warp_sum_int(cnt.x), wnan = warp_sum_int(cnt.y);
            if (lane == 0) { s_nrej[warp] = wrej; s_nnan[warp] = wnan; }
            __syncthreads();
            int nrej = 0, nnan = 0;
            for (int w = 0; w < nw; w++) { nrej += s_nrej[w]; nnan += s_nnan[w]; }
            int nslots = s_ntot < P.sort_cap ? s_ntot : P.sort_cap;  // key slots in use (some hold +inf)
            if (nslots > (int)P.sstride) nslots = (int)P.sstride;
            nsky_all = s_ntot - nrej;       // aper.py:267 counts NaNs
            nvalid = nsky_all - nnan;       // mmm.py:106 drops them
            if (nvalid > nslots) nvalid = nslots;
            if (nsky_all < P.minsky) {  // aper.py:269-272
                if (tid == 0) { P.nsky[s] = nsky_all; P.mmm_iter[s] = -2; }
                __syncthreads();  // the tables are rewritten by the next star
                continue;
            }
            // ---- P2: sort (mmm.py:119); the tail is padded with +inf; sorted keys go to the scratch row
            T* __restrict__ dst = skeys + (size_t)(i - P.s0 + P.row0) * P.sstride;
            if constexpr (K > 0) {
                T r[K];
                const int nwarps = ((nslots + K - 1) / K + 31) >> 5, nthr = nwarps * 32;  // warps / threads that get real keys
#pragma unroll
                for (int q = 0; q < K; q++) {
                    const int idx = q * nthr + tid;  // any bijection onto the active threads will do before sorting
                    r[q] = (tid < nthr && idx < nslots) ? keys[idx] : KeyTraits<T>::inf();
                }
                __syncthreads();
                cta_sort_regs<T, K>(r, keys, tid, nwarps);
                constexpr int VEC = 16 / sizeof(T);  // sstride is a multiple of VEC, rows are 16-byte aligned
#pragma unroll
                for (int q = 0; q < K; q += VEC) {
                    const int e = tid * K + q;
                    if (e < nvalid) {
                        if constexpr (sizeof(T) == 4) *reinterpret_cast<float4*>(dst + e) = make_float4(r[q], r[q + 1], r[q + 2], r[q + 3]);
                        else *reinterpret_cast<double2*>(dst + e) = make_double2(r[q], r[q + 1]);
                    }
                }
                if (P.sidx) sort_index<T, K>(P, r, dst, nvalid, (size_t)(i - P.s0 + P.row0), tid, s_wsum, s_med);
            } else {
                int m = 2;
                while (m < nslots) m <<= 1;
                for (int i = nslots + tid; i < m; i += (int)blockDim.x) keys[i] = KeyTraits<T>::inf();
                __syncthreads();
                cta_bitonic_sort(keys, m, tid);
                for (int i = tid; i < nvalid; i += (int)blockDim.x) dst[i] = keys[i];
            }
        }

        if (tid == 0) { P.nsky[s] = nsky_all; P.mmm_iter[s] = nvalid; }
        __syncthreads();  // keys / tables / counters are reused by the next star
    }
}

// Stage 1 for annuli that do not fit one CTA's shared memory (more than 32 768 fp32 / 16 384 fp64 pixels, e.g.
// skyrad=[100,400]; aper.py:257-288 has no size limit): the same exact spans and slot-addressed gather, but the keys go
// UNSORTED into the star's row of the global scratch (P.skeys, P.sstride = P.sort_cap slots) and seg_count[i - s0] gets the
// number of slots in use; a segmented device sort (bphot_capi.cu) orders every row, aper_stage2_kernel runs mmm on it.
// Only the span tables are in shared memory (20 bytes per stamp row).
constexpr int AP_BIG_THREADS = 256;
template <typename T>
__global__ void __launch_bounds__(AP_BIG_THREADS) aper_gather_global_kernel(const AperParams P, int* __restrict__ seg_count) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int* span = reinterpret_cast<int*>(smem_raw);
    int* rowoff = span + 4 * P.span_rows;
    __shared__ int s_ntot, s_star;
    __shared__ int s_nrej[AP_BIG_THREADS / 32], s_nnan[AP_BIG_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = (int)(blockDim.x >> 5);
    const T* __restrict__ img = reinterpret_cast<const T*>(P.img);
    T* __restrict__ skeys = reinterpret_cast<T*>(P.skeys);
    const double rinsq = P.skyrad_in * P.skyrad_in, routsq = P.skyrad_out * P.skyrad_out;
    for (;;) {
        __syncthreads();  // the previous star's tables and counters are no longer read
        if (tid == 0) s_star = atomicAdd(P.work_counter, 1);
        __syncthreads();
        const int i = P.s0 + s_star;
        if (i >= P.s0 + P.scount) break;
        const int s = P.order ? P.order[i] : i;
        const double x = P.xc[s], y = P.yc[s];
        if (tid == 0) seg_count[i - P.s0] = 0;
        if (!(x >= 0.5) || !(x <= P.ncol - 1.5) || !(y >= 0.5) || !(y <= P.nrow - 1.5)) {  // aper.py:209-210
            if (tid == 0) { P.nsky[s] = 0; P.mmm_iter[s] = -1; }
            continue;
        }
        const StarGeom g = star_geom(P, x, y);
        row_spans(g, rinsq, routsq, span, rowoff, tid);
        __syncthreads();
        if (warp == 0) {
            const int tot = row_offsets(rowoff, g.ny, lane);
            if (lane == 0) s_ntot = tot;
        }
        __syncthreads();
        T* __restrict__ dst = skeys + (size_t)(i - P.s0) * P.sstride;
        const int2 cnt = gather_rows<T>(P, img + g.origin, g, span, rowoff, dst, warp, lane);
        const int wrej = warp_sum_int(cnt.x), wnan = warp_sum_int(cnt.y);
        if (lane == 0) { s_nrej[warp] = wrej; s_nnan[warp] = wnan; }
        __syncthreads();
        if (tid == 0) {
            int nrej = 0, nnan = 0;
            for (int w = 0; w < nw; w++) { nrej += s_nrej[w]; nnan += s_nnan[w]; }
            const int nslots = s_ntot < P.sort_cap ? s_ntot : P.sort_cap;  // the bound is rigorous: always s_ntot
            const int nsky_all = s_ntot - nrej;                              // aper.py:267 counts NaNs
            int nvalid = nsky_all - nnan;                                    // mmm.py:106 drops them (they hold +inf here)
            if (nvalid > nslots) nvalid = nslots;
            if (nsky_all < P.minsky) { P.nsky[s] = nsky_all; P.mmm_iter[s] = -2; }  // aper.py:269-272
            else { P.nsky[s] = nsky_all; P.mmm_iter[s] = nvalid; seg_count[i - P.s0] = nslots; }
        }
    }
}

// Aperture sums, one WARP per star (aper.py:310-380): no CTA-wide barrier, a small code footprint of its own, and
// ~40 stars in flight per SM.  Leaves flux[s,k] = sum of weight*pixel and badflag[s,k] = raw flag bits for stage 2.
template <typename T>
__global__ void __launch_bounds__(AP_THREADS, AP_SUMS_CTAS) aper_sums_kernel(const AperParams P) {
    __shared__ CrossList s_list[AP_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    (void)warp;
    const T* __restrict__ img = reinterpret_cast<const T*>(P.img);
    const int naper = P.naper;
#if AP_DYNAMIC
    for (;;) {
        int i = 0;
        if (lane == 0) i = atomicAdd(P.work_counter, 1);
        i = __shfl_sync(FULL, i, 0);
        if (i >= P.nstars) break;
#else
    for (int i = blockIdx.x * AP_WARPS + warp; i < P.nstars; i += gridDim.x * AP_WARPS) {
#endif
        const int s = P.order ? P.order[i] : i;
        const double x = P.xc[s], y = P.yc[s];
        if (!(x >= 0.5) || !(x <= P.ncol - 1.5) || !(y >= 0.5) || !(y <= P.nrow - 1.5)) continue;  // aper.py:209-210, flagged by stage 1
        const StarGeom g = star_geom(P, x, y);
        const T* __restrict__ stamp = img + g.origin;
#if AP_PREFETCH
        {   // the rows of the largest aperture's box, requested before the first aperture is walked
            double amax = P.apr[0];
            for (int k = 1; k < naper; k++) amax = fmax(amax, P.apr[k]);
            int pr0 = (int)floor(g.dy - amax - 0.5) - 1, pr1 = (int)ceil(g.dy + amax + 0.5);
            int pc0 = (int)floor(g.dx - amax - 0.5), pc1 = (int)ceil(g.dx + amax + 0.5);
            if (pr0 < 0) pr0 = 0;
            if (pr1 > g.ny - 1) pr1 = g.ny - 1;
            if (pc0 < 0) pc0 = 0;
            if (pc1 > g.nx - 1) pc1 = g.nx - 1;
            if (pr1 >= pr0 && pc1 >= pc0)
                prefetch_rows(stamp, (size_t)P.pitch * sizeof(T), pr0, pr1 - pr0 + 1, (size_t)pc0 * sizeof(T), (size_t)(pc1 + 1) * sizeof(T), lane);
        }
#endif
        for (int k = 0; k < naper; k++) {
            const double ap = P.apr[k];
            double total = 0.0;
            int flags = 0;
            if (g.edge >= ap) {  // aper.py:311
                const ApPartial A = aperture_warp<T>(P, stamp, g, ap, 0, 1, lane, &s_list[warp]);
                const double area = M_PI * ap * ap;
                total = A.sfull;
                if (!P.exact && A.npart > 0) total = A.sfull + ((area - A.nfull) / A.wpart) * A.spart;  // aper.py:343-349
                flags = A.flags;
                if (flags & 1) total = 0.0;  // fractn = 0 for the whole aperture, aper.py:366
            }
            if (lane == 0) {
                const size_t o = (size_t)s * naper + k;
                P.flux[o] = total;
                P.badflag[o] = (uint8_t)flags;
            }
        }
    }
}

// Stage 2: one warp per star -- mmm on the sorted scratch row (aper.py:288).  Writes the per-star sky columns only:
// sky = skymod, skyerr = mmm's sigma UNCLAMPED (aper_finalize_kernel needs it for skyvar and clamps it afterwards),
// nsky, mmm_iter.  The per-aperture columns belong to aper_sums_kernel (raw sums, possibly running concurrently on
// another stream) and aper_finalize_kernel.
template <typename T>
__global__ void __launch_bounds__(AP_THREADS, AP_S2_MINB) aper_stage2_kernel(const AperParams P) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    (void)warp;
    const T* __restrict__ skeys = reinterpret_cast<const T*>(P.skeys);
#if AP_DYNAMIC
    for (;;) {
        int i = 0;
        if (lane == 0) i = atomicAdd(P.work_counter, 1);
        i = P.s0 + __shfl_sync(FULL, i, 0);
        if (i >= P.s0 + P.scount) break;
#else
    for (int i = P.s0 + blockIdx.x * AP_WARPS + warp; i < P.s0 + P.scount; i += gridDim.x * AP_WARPS) {
#endif
        const int s = P.order ? P.order[i] : i;
        const int status = P.mmm_iter[s], nsky_all = P.nsky[s];
        double skymod = 0.0, skysig = 0.0;
        int iters = 0, nsky_out = nsky_all;
        if (status == -1) nsky_out = 0;   // star off the frame, aper.py:221-233: sky = skyerr = 0
        else if (status == -2) {}         // nsky < minsky, aper.py:269-272: sky = skyerr = 0
        else if (!P.use_setsky) {
            GlobalView<T> view;
            view.p = skeys + (size_t)(i - P.s0) * P.sstride;
            MmmResult R;
            if (!P.sky_sigclip) R = mmm_warp(view, status, P.readnoise, P.minsky, P.mxiter, lane);
            else if (status != nsky_all) { R.skymod = nan64(); R.sigma = nan64(); R.iters = 0; }  // a NaN in the annulus: scipy propagates it
            else R = sigclip_warp(view, status, 4.0, 4.0, lane);  // aper.py:275-282
            skymod = R.skymod; skysig = R.sigma; iters = R.iters;
        } else {  // aper.py:302-308
            skymod = P.setsky[0]; skysig = P.setsky[1];
            nsky_out = (int)P.setsky[2];
        }
        if (lane == 0) { P.sky[s] = skymod; P.skyerr[s] = skysig; P.nsky[s] = nsky_out; P.mmm_iter[s] = iters; }
    }
}

// Stage 2, indexed flavour: one THREAD per star runs mmm (mmm.py:121-314) on the sorted row through its index
// (IndexedView).  32 stars per warp: ~1.5 k warp-instructions per star instead of the warp-per-star kernel's 14 k.
// Range sums come from prefix differences, so their last bits differ from a direct sum (1e-13 relative); counts and
// iteration numbers are decided by the same comparisons on the same keys.
#ifndef AP_MMM_OVERLAP
#define AP_MMM_OVERLAP 0
#endif
// with AP_MMM_OVERLAP: 64 threads x <= 64 registers = the 4 096 registers per SM that stage 1's 8 CTAs leave free
#ifndef AP_MMM_MINB
#define AP_MMM_MINB 1  // resident CTAs per SM the thread-per-star mmm is compiled for
#endif
constexpr int AP_MMM_THREADS = AP_MMM_OVERLAP ? 64 : AP_THREADS;
template <typename T>
__global__ void __launch_bounds__(AP_MMM_THREADS, AP_MMM_OVERLAP ? 16 : AP_MMM_MINB) aper_mmm_kernel(const AperParams P) {
    const T* __restrict__ skeys = reinterpret_cast<const T*>(P.skeys);
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < P.scount; j += gridDim.x * blockDim.x) {
        const int i = P.s0 + j;
        const int s = P.order ? P.order[i] : i;
        const int status = P.mmm_iter[s], nsky_all = P.nsky[s];
        double skymod = 0.0, skysig = 0.0;
        int iters = 0, nsky_out = nsky_all;
        if (status == -1) nsky_out = 0;   // star off the frame, aper.py:221-233: sky = skyerr = 0
        else if (status == -2) {}         // nsky < minsky, aper.py:269-272: sky = skyerr = 0
        else {
            IndexedView<T> V;
            const unsigned char* base = reinterpret_cast<const unsigned char*>(P.sidx) + (size_t)j * (size_t)P.idx_stride;
            V.k = skeys + (size_t)j * P.sstride;
            V.L1 = reinterpret_cast<const T*>(base);
            V.L2 = V.L1 + 16;
            V.PT = reinterpret_cast<const double*>(V.L2 + 256);
            V.n = status;
            V.skymid = status > 0 ? 0.5 * V((status - 1) / 2) + 0.5 * V(status / 2) : 0.0;
            const MmmResult R = mmm_sorted(
                V, [&](int lo, int hi, double, double* s1, double* s2) { V.delta_sums(lo, hi, s1, s2); },
                [&](int lo, int hi) { return V.plain_sum(lo, hi); },
                [&](double v1, bool st1, bool need1, double v2, bool st2, bool need2, int* r1, int* r2) {
                    V.first2(v1, st1, need1, v2, st2, need2, r1, r2);
                },
                [&](int lo0, int hi0, bool need0, int lo1, int hi1, bool need1, int lo2, int hi2, double, double* o5) {
                    V.sums3(lo0, hi0, need0, lo1, hi1, need1, lo2, hi2, o5);
                },
                status, P.readnoise, P.minsky, P.mxiter);
            skymod = R.skymod; skysig = R.sigma; iters = R.iters;
        }
        P.sky[s] = skymod; P.skyerr[s] = skysig; P.nsky[s] = nsky_out; P.mmm_iter[s] = iters;
    }
}

// Last step, one thread per star: sky-subtracted flux, error, magnitude and flags from the raw aperture sums and the
// sky columns (aper.py:290-300, 383-403, 411-417).
__global__ void __launch_bounds__(256) aper_finalize_kernel(const AperParams P) {
    const int naper = P.naper;
    for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < P.nstars; s += gridDim.x * blockDim.x) {
        const double x = P.xc[s], y = P.yc[s];
        const bool badstar = !(x >= 0.5) || !(x <= P.ncol - 1.5) || !(y >= 0.5) || !(y <= P.nrow - 1.5);  // aper.py:209-210
        const double skymod = P.sky[s];
        double skysig = P.skyerr[s];
        bool bad = badstar;
        double skyvar = 0.0, sigsq = 0.0;
        if (!bad) {
            if (!P.use_setsky) {
                const int nsky_all = P.nsky[s];
                if (nsky_all < P.minsky) bad = true;  // aper.py:269-272
                else {
                    skyvar = skysig * skysig;  // aper.py:290-291
                    sigsq = skyvar / nsky_all;
                    if (skysig < 0.0) bad = true;  // aper.py:293-296: sky and skyerr = -1 stay as mmm left them
                    else if (skysig > 999.99) P.skyerr[s] = 999;  // aper.py:298
                }
            } else {
                skyvar = skysig * skysig;
                sigsq = skyvar / P.setsky[2];
            }
        }
        StarGeom g;
        if (!bad) g = star_geom(P, x, y);
        for (int k = 0; k < naper; k++) {
            const size_t o = (size_t)s * naper + k;
            const double ap = P.apr[k];
            if (bad || !(g.edge >= ap)) {  // aper.py:221-233, 311, 350-353
                P.mag[o] = nan64(); P.magerr[o] = nan64(); P.flux[o] = nan64(); P.fluxerr[o] = nan64();
                P.badflag[o] = 1;
                continue;
            }
            const double area = M_PI * ap * ap;
            const double apflux = P.flux[o] - skymod * area;  // aper.py:391
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
            P.badflag[o] = P.badflag[o] ? 1 : 0;
        }
    }
}

// ---------------------------------------------------- unit-test kernels
// Batched mmm over CSR-packed vectors (row a4 of the scope table): one CTA per vector, same sort
// and clipping code as aper_kernel (K as there; sort_cap = 128*K when K > 0).
template <typename T, int K>
__global__ void __launch_bounds__(AP_THREADS) mmm_kernel(const T* vec, const long long* offs, int nvec, double readnoise,
                                                         int minsky, int mxiter, int sort_cap, double* out3, int* iters) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* keys = reinterpret_cast<T*>(smem_raw);
    __shared__ int s_n;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int v = blockIdx.x; v < nvec; v += gridDim.x) {
        const long long b = offs[v], e = offs[v + 1];
        if (tid == 0) s_n = 0;
        __syncthreads();
        for (long long ib = b + (long long)warp * 32; ib < e; ib += AP_THREADS) {
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
        SortedView<T> view;
        view.p = keys;
        view.shift = 31;
        if constexpr (K > 0) {
            T r[K];
#pragma unroll
            for (int q = 0; q < K; q++) {
                const int idx = q * AP_THREADS + tid;
                r[q] = idx < n ? keys[idx] : KeyTraits<T>::inf();
            }
            __syncthreads();
            cta_sort_regs<T, K>(r, keys, tid, AP_WARPS);
#pragma unroll
            for (int q = 0; q < K; q++) keys[tid * (K + 1) + q] = r[q];
            view.shift = __ffs(K) - 1;
        } else {
            int m = 2;
            while (m < n) m <<= 1;
            for (int i = n + tid; i < m; i += AP_THREADS) keys[i] = KeyTraits<T>::inf();
            __syncthreads();
            cta_bitonic_sort(keys, m, tid);
        }
        __syncthreads();
        if (warp == 0) {
            MmmResult R = mmm_warp(view, n, readnoise, minsky, mxiter, lane);
            if (lane == 0) {
                out3[3 * (size_t)v + 0] = R.skymod; out3[3 * (size_t)v + 1] = R.sigma; out3[3 * (size_t)v + 2] = R.skew;
                iters[v] = R.iters;
            }
        }
        __syncthreads();
    }
}

// Pixwt(xc, yc, r, x[], y[]) -- pixwt.py:8 (row a5).
// mmm over vectors that do not fit one CTA's shared memory (mmm.py has no size limit): mmm_prep_global_kernel copies every
// vector into scratch with NaNs replaced by +inf (mmm.py:106 drops them) and counts the rest, a segmented device sort orders
// the copies (bphot_capi.cu), mmm_global_kernel runs the clipping loop with one warp per vector on the sorted copy.
template <typename T>
__global__ void __launch_bounds__(256) mmm_prep_global_kernel(const T* __restrict__ vec, const long long* __restrict__ offs, int nvec,
                                                              T* __restrict__ out, int* __restrict__ nvalid) {
    __shared__ int s_cnt[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int v = blockIdx.x; v < nvec; v += gridDim.x) {
        const long long b = offs[v], e = offs[v + 1];
        int cnt = 0;
        for (long long i = b + threadIdx.x; i < e; i += blockDim.x) {
            const T val = vec[i];
            const bool ok = (val == val);
            cnt += ok;
            out[i] = ok ? val : KeyTraits<T>::inf();
        }
        cnt = warp_sum_int(cnt);
        __syncthreads();
        if (lane == 0) s_cnt[warp] = cnt;
        __syncthreads();
        if (threadIdx.x == 0) {
            int t = 0;
            for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += s_cnt[w];
            nvalid[v] = t;
        }
    }
}
template <typename T>
__global__ void __launch_bounds__(AP_THREADS) mmm_global_kernel(const T* __restrict__ sorted, const long long* __restrict__ offs,
                                                                const int* __restrict__ nvalid, int nvec, double readnoise, int minsky,
                                                                int mxiter, double* out3, int* iters) {
    const int lane = threadIdx.x & 31;
    for (int v = blockIdx.x * AP_WARPS + (threadIdx.x >> 5); v < nvec; v += gridDim.x * AP_WARPS) {
        GlobalView<T> view;
        view.p = sorted + offs[v];
        const MmmResult R = mmm_warp(view, nvalid[v], readnoise, minsky, mxiter, lane);
        if (lane == 0) {
            out3[3 * (size_t)v + 0] = R.skymod; out3[3 * (size_t)v + 1] = R.sigma; out3[3 * (size_t)v + 2] = R.skew;
            iters[v] = R.iters;
        }
    }
}

__global__ void pixwt_kernel(double xc, double yc, double r, const double* x, const double* y, long long n, double* out) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = pixwt(xc, yc, r, x[i], y[i]);
}

}  // namespace bphot
