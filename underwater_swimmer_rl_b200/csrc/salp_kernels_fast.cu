// salp_kernels_fast.cu -- instantiates and launches the forced-breathing f32 step kernel (salp_step_fast.cuh).
#include <cstdlib>

#include "salp_step_fast.cuh"

namespace salp {

namespace {

constexpr int kMaxDevices = 64;

template <int F, int RNG>
cudaError_t launch_fast(const KernelArgs &a, cudaStream_t s) {
    using Cfg = FastCfg<F>;
    using L = FastLayout<F>;
    auto kern = salp_step_fast_kernel<F, RNG>;
    constexpr size_t smem = (size_t)L::kWarpBytes * Cfg::warps;
    // per instantiation AND per device (a process may drive several GPUs through several handles)
    static int resident[kMaxDevices] = {0};
    int dev = 0;
    cudaError_t err = cudaGetDevice(&dev);
    if (err != cudaSuccess) return err;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    if (resident[dev] == 0) {
        int sms = 0, occ = 0;
        err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err == cudaSuccess) err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (err == cudaSuccess) err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, Cfg::warps * 32, smem);
        if (err != cudaSuccess) return err;
        resident[dev] = sms * (occ > 0 ? occ : 1);
    }
    const unsigned ntiles = (unsigned)((a.count + 31) / 32);
    const unsigned want = (ntiles + Cfg::warps - 1) / Cfg::warps;
    const unsigned grid = want < (unsigned)resident[dev] ? want : (unsigned)resident[dev];
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(Cfg::warps * 32); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    // programmatic dependent launch: CTAs of the next step may become resident (and run their prologue: barrier init, cycle
    // table -> shared memory) while this step's last CTAs drain; they wait at griddepcontrol.wait before touching state
    cfg.attrs = attr; cfg.numAttrs = (getenv("SALP_NO_PDL") ? 0 : 1);
    return cudaLaunchKernelEx(&cfg, kern, a);
}

}  // namespace

// The specialised kernel covers: f32, forced breathing, max_observed_food == 3, 1 or 12 food slots, and a launch range
// that starts on a 32-env boundary (bulk copies need 16-byte aligned addresses in every per-env array).
cudaError_t launch_step_fast_f32(const KernelArgs &a, int food_slots, cudaStream_t s, bool *handled) {
    *handled = false;
    if (!a.forced || a.observed != 3 || a.act_dim != 1 || a.obs_dim != 24) return cudaSuccess;
    if ((a.first & 31) != 0 || (a.N & 1) != 0 || a.cycle_tab == nullptr) return cudaSuccess;  // food rows [N] float2: 16-byte aligned
    if (a.count <= 0) { *handled = true; return cudaSuccess; }
    cudaError_t e;
    if (food_slots == 1) e = a.rng_mode ? launch_fast<1, 1>(a, s) : launch_fast<1, 0>(a, s);
    else if (food_slots == 12) e = a.rng_mode ? launch_fast<12, 1>(a, s) : launch_fast<12, 0>(a, s);
    else return cudaSuccess;
    *handled = true;
    return e;
}

}  // namespace salp

#ifdef SALP_LAB_CLOCK
extern "C" int salp_lab_read_layout(unsigned long long *dst, int clear) {
    cudaError_t e = cudaMemcpyFromSymbol(dst, salp::g_lab_lay, sizeof(salp::g_lab_lay));
    if (e == cudaSuccess && clear) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, salp::g_lab_lay);
        if (e == cudaSuccess) e = cudaMemset(p, 0, sizeof(salp::g_lab_lay));
    }
    return (int)e;
}
extern "C" int salp_lab_read_clocks(long long *dst, int n, int clear) {
    cudaError_t e = cudaMemcpyFromSymbol(dst, salp::g_lab_clk, (size_t)n * sizeof(long long));
    if (e == cudaSuccess && clear) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, salp::g_lab_clk);
        if (e == cudaSuccess) e = cudaMemset(p, 0, sizeof(salp::g_lab_clk));
    }
    return (int)e;
}
#endif
