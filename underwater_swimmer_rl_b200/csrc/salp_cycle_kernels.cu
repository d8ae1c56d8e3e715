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

template <int MINB>
static cudaError_t launch_cycle_variant(const CycleArgs &a, cudaStream_t s) {
    // one resident wave: (CTAs that fit on an SM) x (SMs), fewer when the slab is small
    static int resident = 0;
    if (!resident) {
        int dev = 0, sms = 0, per_sm = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, salp_cycle_step_kernel<cyc_real_t, MINB>, kCycleBlock, 0);
        if (e != cudaSuccess) return e;
        resident = sms * (per_sm > 0 ? per_sm : 1);
    }
    long long grid = (a.n + 2 * kCycleBlock - 1) / (2 * kCycleBlock);   // at least two chunks per warp
    if (grid > resident) grid = resident;
    if (grid < 1) grid = 1;
    salp_cycle_step_kernel<cyc_real_t, MINB><<<(unsigned)grid, kCycleBlock, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t SALP_LAUNCH_CYCLE_STEP(const CycleArgs &a, cudaStream_t s) {
    if (a.n <= 0) return cudaSuccess;
    // register budget 65536 / (128 * MINB): measured on B200 at 262144 envs -- 4: 1.40 ms, 5: 1.39, 6: 1.56 (spills),
    // 7: 1.51, 8: 1.50 -- (profiles/r01_ncu_robot_cycle.md); the double build takes what it needs
    return launch_cycle_variant<SALP_BUILD_F64 ? 1 : 5>(a, s);
}

cudaError_t SALP_LAUNCH_CYCLE_RESET(const CycleArgs &a, cudaStream_t s) {
    if (a.n <= 0) return cudaSuccess;
    const unsigned grid = (unsigned)((a.n + 255) / 256);
    salp_cycle_reset_kernel<cyc_real_t><<<grid, 256, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace salp
