// bphot_device.cuh -- warp/CTA helpers shared by the kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "bphot_math.cuh"

namespace bphot {

constexpr unsigned FULL = 0xffffffffu;

// Butterfly sum: every lane ends with the same bits (a+b == b+a), so control flow that
// depends on the result stays warp-uniform.
BP_D double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = v + __shfl_xor_sync(FULL, v, o);
    return v;
}
// Sixteen butterfly sums at once.  Level by level (lane distance 16, 8, 4, 2) every lane passes on the half of its values
// its partner keeps and adds what it receives to the half it keeps, so a level moves 8, 4, 2, 1 values instead of 16; the
// last level (distance 1) is a plain butterfly and value k ends up complete in lanes 2k and 2k+1, from where it is broadcast.
// Every value goes through the same additions, in the same order, as in warp_sum(): the totals have the same bits, and
// every lane gets the same ones.  60 double shuffles + 16 additions instead of 80 + 80.
BP_D void warp_sum16(double (&v)[16], int lane) {
#pragma unroll
    for (int h = 8; h >= 1; h >>= 1) {
        const int bit = 2 * h;  // lane distance of this level: 16, 8, 4, 2
        const bool up = (lane & bit) != 0;
#pragma unroll
        for (int j = 0; j < h; j++) {
            const double send = up ? v[j] : v[j + h];
            const double keep = up ? v[j + h] : v[j];
            v[j] = keep + __shfl_xor_sync(FULL, send, bit);
        }
    }
    const double tot = v[0] + __shfl_xor_sync(FULL, v[0], 1);
#pragma unroll
    for (int k = 0; k < 16; k++) v[k] = __shfl_sync(FULL, tot, 2 * k);
}
BP_D int warp_sum_int(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = v + __shfl_xor_sync(FULL, v, o);
    return v;
}
BP_D double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(FULL, v, o));
    return v;
}
BP_D double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(FULL, v, o));
    return v;
}

// Asks L1/L2 for the 128-byte lines of rows [row0, row0 + nrows) x bytes [b0, b1) of a pitched array, one row per lane: the
// stamp a warp is about to walk pixel by pixel (first-touch misses were its largest scoreboard stall).  No architectural effect.
BP_D void prefetch_rows(const void* base, size_t pitch_bytes, long long row0, int nrows, size_t b0, size_t b1, int lane) {
    for (int r = lane; r < nrows; r += 32) {
        const char* p = reinterpret_cast<const char*>(base) + (size_t)(row0 + r) * pitch_bytes;
        for (size_t o = b0; o < b1; o += 128) asm volatile("prefetch.global.L1 [%0];" ::"l"(p + o));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(p + b1 - 1));  // the line of the last byte (rows are not 128-byte aligned)
    }
}

template <typename T> struct KeyTraits;
template <> struct KeyTraits<float> {
    BP_D static float inf() { return __int_as_float(0x7f800000); }
};
template <> struct KeyTraits<double> {
    BP_D static double inf() { return __longlong_as_double(0x7ff0000000000000LL); }
};

BP_D double nan64() { return __longlong_as_double(0x7ff8000000000000LL); }

}  // namespace bphot
