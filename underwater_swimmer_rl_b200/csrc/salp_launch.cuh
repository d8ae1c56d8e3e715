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
    if (PF) {  // persistent: as many CTAs as fit on the chip at once
        if (occ_smem != smem) {
            int dev = 0;
            cudaError_t err = cudaGetDevice(&dev);
            if (err == cudaSuccess) err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            if (err == cudaSuccess) err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_blocks, kern, BLOCK, smem);
            if (err != cudaSuccess) return err;
            occ_smem = smem;
        }
        const unsigned resident = (unsigned)(sms * (occ_blocks > 0 ? occ_blocks : 1));
        grid = ntiles < resident ? ntiles : resident;
    }
    kern<<<grid, BLOCK, smem, s>>>(a);
    return cudaGetLastError();
}

}  // namespace salp
