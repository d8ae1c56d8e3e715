// salp_launch.cuh -- host-side launch of one instantiation of the step kernel.
#pragma once
#include "salp_step.cuh"

namespace salp {

template <typename T, int F, int MT, int STORE, int BLOCK, int MINB, int PF, int RNG = 0>
cudaError_t launch_step_variant(const KernelArgs &a, cudaStream_t s) {
    static_assert(BLOCK % 32 == 0 && BLOCK >= 32 && BLOCK <= 1024, "BLOCK must be a multiple of the warp size");
    auto kern = salp_step_kernel<T, F, MT, STORE, BLOCK, MINB, PF, RNG>;
    constexpr bool tma = (STORE == kStoreTmaWarp || STORE == kStoreTmaBlock);
    size_t smem;
    if (PF) smem = stage_offset_bytes<STORE, BLOCK>(a.obs_dim) + stage_bytes<T, F, BLOCK>();
    else {
        const int nbuf = STORE == kStoreDirect ? 0 : (tma ? ((a.K > 1 || STORE == kStoreTmaWarp) ? 2 : 1) : 1);
        smem = (size_t)nbuf * BLOCK * a.obs_dim * sizeof(float);
    }
    // per instantiation AND per device: a process may drive several GPUs through several handles
    constexpr int kMaxDev = 64;
    static size_t smem_set[kMaxDev] = {0}, occ_smem[kMaxDev] = {0};
    static int occ_blocks[kMaxDev] = {0}, sms_[kMaxDev] = {0};
    int dev = 0;
    cudaError_t err = cudaGetDevice(&dev);
    if (err != cudaSuccess) return err;
    if (dev < 0 || dev >= kMaxDev) return cudaErrorInvalidDevice;
    if (smem > 48 * 1024 && smem > smem_set[dev]) {
        err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
        smem_set[dev] = smem;
    }
    const unsigned ntiles = (unsigned)((a.count + BLOCK - 1) / BLOCK);
    unsigned grid = ntiles;
    if (occ_blocks[dev] == 0 || occ_smem[dev] != smem) {  // CTAs of this instantiation that fit on the chip at once
        err = cudaDeviceGetAttribute(&sms_[dev], cudaDevAttrMultiProcessorCount, dev);
        if (err == cudaSuccess) err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_blocks[dev], kern, BLOCK, smem);
        if (err != cudaSuccess) return err;
        if (occ_blocks[dev] < 1) occ_blocks[dev] = 1;
        occ_smem[dev] = smem;
    }
    const int sms = sms_[dev];
    const int occ_b = occ_blocks[dev];
    const unsigned resident = (unsigned)(sms * occ_b);
    if (PF) grid = ntiles < resident ? ntiles : resident;  // persistent
    // Programmatic dependent launch for grids of at least one resident wave: the first CTAs of this launch become resident
    // while the last wave of the previous kernel of the stream drains (they block at griddepcontrol.wait before touching
    // global memory), so launch latency and ramp hide behind that tail.  Measured on B200: single_food_1m 52.7 -> 51.0 us,
    // defaults_1m 106.6 -> 104.3 us; grids of a fraction of a wave got slower (65 536 envs, 32-step graph: 7.1 -> 7.6 us)
    // and keep the plain launch.
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(BLOCK); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = grid >= resident ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, a);
}

}  // namespace salp
