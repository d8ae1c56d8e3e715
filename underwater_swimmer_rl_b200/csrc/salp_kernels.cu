// salp_kernels.cu -- instantiates the step/reset kernels for ONE arithmetic type.
// Compiled twice by build.py:
//   -DSALP_BUILD_F64=0                 -> launch_step_f32 / launch_reset_f32   (throughput build)
//   -DSALP_BUILD_F64=1 -fmad=false     -> launch_step_f64 / launch_reset_f64   (validation build)
#include "salp_launch.cuh"

namespace salp {

#if SALP_BUILD_F64
using real_t = double;
#define SALP_LAUNCH_STEP launch_step_f64
#define SALP_LAUNCH_RESET launch_reset_f64
#define SALP_LAUNCH_OBSERVE launch_observe_f64
#else
using real_t = float;
#define SALP_LAUNCH_STEP launch_step_f32
#define SALP_LAUNCH_RESET launch_reset_f32
#define SALP_LAUNCH_OBSERVE launch_observe_f32
#endif

// Launch configuration per food-slot count, chosen from tools/kernel_lab.cu timings on B200
// (1,048,576 envs, steady state; DESIGN.md "kernel variants"):
//   STORE  how the observation tile leaves the SM      BLOCK  threads per CTA
//   REGS   register budget per thread (f32 build); MINB = 65536 / (REGS * BLOCK) CTAs per SM
template <int F> struct StepCfg {
    static constexpr int store = (F <= 12) ? kStoreTmaBlock : kStoreDirect;
    static constexpr int block = 64;
    static constexpr int prefetch = 0;
    static constexpr int regs_f32 = F <= 4 ? 64 : (F <= 12 ? 80 : (F <= 16 ? 128 : 255));
    // validation build: twice the f32 budget, except that up to four food slots run best at 96 (20 warps/SM, 12-20 bytes of
    // spill: 1 M single-food envs 90.4 us at 128 registers, 80.7 at 96, 84.8 at 80 -- measured on B200)
#ifndef SALP_F64_REGS_SMALLF
#define SALP_F64_REGS_SMALLF 96
#endif
#ifndef SALP_F64_REGS_F12
#define SALP_F64_REGS_F12 160
#endif
    static constexpr int regs = SALP_BUILD_F64 ? (F <= 4 ? SALP_F64_REGS_SMALLF : (F <= 12 ? SALP_F64_REGS_F12 : (2 * regs_f32 > 255 ? 255 : 2 * regs_f32))) : regs_f32;
    static constexpr int minb = (65536 / (regs * block)) > 32 ? 32 : (65536 / (regs * block) < 1 ? 1 : 65536 / (regs * block));
};

template <int F>
static cudaError_t launch_step_f(const KernelArgs &a, cudaStream_t s) {
    using C = StepCfg<F>;
    if (a.rng_mode)
        return a.observed == 3 ? launch_step_variant<real_t, F, 3, C::store, C::block, C::minb, C::prefetch, 1>(a, s)
                               : launch_step_variant<real_t, F, 8, C::store, C::block, C::minb, C::prefetch, 1>(a, s);
    return a.observed == 3 ? launch_step_variant<real_t, F, 3, C::store, C::block, C::minb, C::prefetch, 0>(a, s)
                           : launch_step_variant<real_t, F, 8, C::store, C::block, C::minb, C::prefetch, 0>(a, s);
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

cudaError_t SALP_LAUNCH_OBSERVE(const KernelArgs &a, cudaStream_t s) {
    if (a.count <= 0) return cudaSuccess;
    salp_observe_kernel<real_t><<<(unsigned)((a.count + 127) / 128), 128, 0, s>>>(a);
    return cudaGetLastError();
}

#if !SALP_BUILD_F64
cudaError_t launch_refresh_rng(const KernelArgs &a, cudaStream_t s) {
    if (a.count <= 0) return cudaSuccess;
    salp_refresh_rng_kernel<<<(unsigned)((a.count + 255) / 256), 256, 0, s>>>(a);
    return cudaGetLastError();
}
#endif

}  // namespace salp
