// salp_kernels.cu -- instantiates the step/reset kernels for ONE arithmetic type.
// Compiled twice by build.py:
//   -DSALP_BUILD_F64=0                 -> launch_step_f32 / launch_reset_f32   (throughput build)
//   -DSALP_BUILD_F64=1 -fmad=false     -> launch_step_f64 / launch_reset_f64   (validation build)
// SALP_STORE / SALP_BLOCK pick the observation-store mode and CTA size that tools/kernel_lab.cu
// measured fastest on B200 (see DESIGN.md "kernel variants").
#include "salp_launch.cuh"

#ifndef SALP_STORE
#define SALP_STORE kStoreWarp
#endif
#ifndef SALP_BLOCK
#define SALP_BLOCK 128
#endif


namespace salp {

#if SALP_BUILD_F64
using real_t = double;
#define SALP_LAUNCH_STEP launch_step_f64
#define SALP_LAUNCH_RESET launch_reset_f64
#else
using real_t = float;
#define SALP_LAUNCH_STEP launch_step_f32
#define SALP_LAUNCH_RESET launch_reset_f32
#endif

// CTAs/SM the register allocator must leave room for (registers per thread <= 65536/(MINB*BLOCK)).
// Tuned with tools/kernel_lab.cu; the double build and the wide sorting networks need more registers.
template <int F> constexpr int min_blocks() {
    constexpr int f32 = F <= 4 ? 8 : (F <= 8 ? 6 : (F <= 12 ? 4 : (F <= 16 ? 3 : 2)));
    return (SALP_BUILD_F64 ? (f32 + 1) / 2 : f32) * 128 / SALP_BLOCK > 0 ? (SALP_BUILD_F64 ? (f32 + 1) / 2 : f32) * 128 / SALP_BLOCK : 1;
}

template <int F>
static cudaError_t launch_step_f(const KernelArgs &a, cudaStream_t s) {
    return a.observed == 3 ? launch_step_variant<real_t, F, 3, SALP_STORE, SALP_BLOCK, min_blocks<F>()>(a, s)
                           : launch_step_variant<real_t, F, 8, SALP_STORE, SALP_BLOCK, min_blocks<F>()>(a, s);
}

cudaError_t SALP_LAUNCH_STEP(const KernelArgs &a, int food_slots, cudaStream_t s) {
    if (a.count <= 0) return cudaSuccess;
    switch (food_slots) {
        case 1: return launch_step_f<1>(a, s);
        case 4: return launch_step_f<4>(a, s);
        case 8: return launch_step_f<8>(a, s);
        case 12: return launch_step_f<12>(a, s);
        case 16: return launch_step_f<16>(a, s);
        case 32: return launch_step_f<32>(a, s);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t SALP_LAUNCH_RESET(const KernelArgs &a, int /*food_slots*/, cudaStream_t s) {
    if (a.count <= 0) return cudaSuccess;
    const unsigned grid = (unsigned)((a.count + 255) / 256);
    salp_reset_kernel<real_t><<<grid, 256, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace salp
