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

template <typename T> struct KeyTraits;
template <> struct KeyTraits<float> {
    BP_D static float inf() { return __int_as_float(0x7f800000); }
};
template <> struct KeyTraits<double> {
    BP_D static double inf() { return __longlong_as_double(0x7ff0000000000000LL); }
};

BP_D double nan64() { return __longlong_as_double(0x7ff8000000000000LL); }

}  // namespace bphot
