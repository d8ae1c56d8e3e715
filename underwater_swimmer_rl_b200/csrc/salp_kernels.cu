// salp_kernels.cu -- instantiates the step/reset kernels for ONE arithmetic type.
// Compiled twice by build.py:
//   -DSALP_BUILD_F64=0                 -> launch_step_f32 / launch_reset_f32   (throughput build)
//   -DSALP_BUILD_F64=1 -fmad=false     -> launch_step_f64 / launch_reset_f64   (validation build)
#include "salp_step.cuh"

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

template <int F, int MT>
static cudaError_t launch_step_t(const KernelArgs &a, cudaStream_t s) {
    auto kern = salp_step_kernel<real_t, F, MT>;
    const int nbuf = a.K > 1 ? 2 : 1;
    const size_t smem = (size_t)nbuf * kBlockThreads * a.obs_dim * sizeof(float);
    static size_t smem_set = 0;  // per instantiation
    if (smem > 48 * 1024 && smem > smem_set) {
        cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
        smem_set = smem;
    }
    const unsigned grid = (unsigned)((a.count + kBlockThreads - 1) / kBlockThreads);
    kern<<<grid, kBlockThreads, smem, s>>>(a);
    return cudaGetLastError();
}

template <int F>
static cudaError_t launch_step_f(const KernelArgs &a, cudaStream_t s) {
    return a.observed <= 3 ? launch_step_t<F, 3>(a, s) : launch_step_t<F, 8>(a, s);
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
    const unsigned grid = (unsigned)((a.count + kBlockThreads - 1) / kBlockThreads);
    salp_reset_kernel<real_t><<<grid, kBlockThreads, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace salp
