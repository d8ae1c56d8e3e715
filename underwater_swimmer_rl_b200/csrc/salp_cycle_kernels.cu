// salp_cycle_kernels.cu -- instantiates the model (B) cycle kernels for ONE arithmetic type.
// Compiled twice by build.py:
//   -DSALP_BUILD_F64=0                 -> launch_cycle_step_f32 / launch_cycle_reset_f32   (throughput build)
//   -DSALP_BUILD_F64=1 -fmad=false     -> launch_cycle_step_f64 / launch_cycle_reset_f64   (validation build)
#include "salp_cycle.cuh"

namespace salp {

#if SALP_BUILD_F64
using cyc_real_t = double;
#define SALP_LAUNCH_CYCLE_STEP launch_cycle_step_f64
#define SALP_LAUNCH_CYCLE_RESET launch_cycle_reset_f64
#else
using cyc_real_t = float;
#define SALP_LAUNCH_CYCLE_STEP launch_cycle_step_f32
#define SALP_LAUNCH_CYCLE_RESET launch_cycle_reset_f32
#endif

template <int BLOCK, int MINB>
static cudaError_t launch_cycle_variant(const CycleArgs &a, cudaStream_t s) {
    // one resident wave: (CTAs that fit on an SM) x (SMs), fewer when the slab is small
    static int resident = 0;
    if (!resident) {
        int dev = 0, sms = 0, per_sm = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, salp_cycle_step_kernel<cyc_real_t, BLOCK, MINB>, BLOCK, 0);
        if (e != cudaSuccess) return e;
        resident = sms * (per_sm > 0 ? per_sm : 1);
    }
    long long grid = (a.n + 31) / 32;   // at least one 32-env chunk per CTA
    if (grid > resident) grid = resident;
    if (grid < 1) grid = 1;
    salp_cycle_step_kernel<cyc_real_t, BLOCK, MINB><<<(unsigned)grid, BLOCK, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t SALP_LAUNCH_CYCLE_STEP(const CycleArgs &a, cudaStream_t s) {
    if (a.n <= 0) return cudaSuccess;
    // One CTA per SM whose 20 warps share one pool of sorted chunks (float build; 89 registers x 640 threads).  Measured on
    // B200 at 262144 envs: 5 CTAs x 128 threads 1.264 ms; 2 x 320 1.133; 2 x 256 1.142; 1 x 512 1.017; 1 x 576 1.022;
    // 1 x 640 1.017; 1 x 704 / 736 (80 registers) 1.06 (profiles/r01_ncu_robot_cycle.md).  The double build needs 168
    // registers: 128-thread CTAs.
    return launch_cycle_variant<SALP_BUILD_F64 ? 128 : 640, 1>(a, s);
}

cudaError_t SALP_LAUNCH_CYCLE_RESET(const CycleArgs &a, cudaStream_t s) {
    if (a.n <= 0) return cudaSuccess;
    const unsigned grid = (unsigned)((a.n + 255) / 256);
    salp_cycle_reset_kernel<cyc_real_t><<<grid, 256, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace salp
