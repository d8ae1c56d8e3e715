// salp_launch.cuh -- host-side launch of one instantiation of the step kernel.
#pragma once
#include "salp_step.cuh"

namespace salp {

template <typename T, int F, int MT, int STORE, int BLOCK, int MINB, int PF>
cudaError_t launch_step_variant(const KernelArgs &a, cudaStream_t s) {
    static_assert(BLOCK % 32 == 0 && BLOCK >= 32 && BLOCK <= 1024, "BLOCK must be a multiple of the warp size");
    auto kern = salp_step_kernel<T, F, MT, STORE, BLOCK, MINB, PF>;
    constexpr bool tma = (STORE == kStoreTmaWarp || STORE == kStoreTmaBlock);
    size_t smem;
    if (PF) smem = stage_offset_bytes<STORE, BLOCK>(a.obs_dim) + stage_bytes<T, F, BLOCK>();
    else {
        const int nbuf = STORE == kStoreDirect ? 0 : (tma ? ((a.K > 1 || STORE == kStoreTmaWarp) ? 2 : 1) : 1);
        smem = (size_t)nbuf * BLOCK * a.obs_dim * sizeof(float);
    }
    static size_t smem_set = 0;  // per instantiation
    static size_t occ_smem = ~(size_t)0;
    static int occ_blocks = 0, sms = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
        smem_set = smem;
    }
    const unsigned ntiles = (unsigned)((a.count + BLOCK - 1) / BLOCK);
    unsigned grid = ntiles;
    if (occ_smem != smem) {  // CTAs of this instantiation that fit on the chip at once
        int dev = 0;
        cudaError_t err = cudaGetDevice(&dev);
        if (err == cudaSuccess) err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (err == cudaSuccess) err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_blocks, kern, BLOCK, smem);
        if (err != cudaSuccess) return err;
        occ_smem = smem;
    }
    const unsigned resident = (unsigned)(sms * (occ_blocks > 0 ? occ_blocks : 1));
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
