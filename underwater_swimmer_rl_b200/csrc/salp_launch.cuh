// salp_launch.cuh -- host-side launch of one instantiation of the step kernel.
#pragma once
#include "salp_step.cuh"

namespace salp {

template <typename T, int F, int MT, int STORE, int BLOCK, int MINB>
cudaError_t launch_step_variant(const KernelArgs &a, cudaStream_t s) {
    static_assert(BLOCK % 32 == 0 && BLOCK >= 32 && BLOCK <= 1024, "BLOCK must be a multiple of the warp size");
    auto kern = salp_step_kernel<T, F, MT, STORE, BLOCK, MINB>;
    // TMA modes double-buffer the tile (the kernel indexes buffers by k&1 / warp*2) -- always 2 for kStoreTmaWarp
    const int nbuf = STORE == kStoreDirect ? 0 : (STORE == kStoreWarp ? 1 : (STORE == kStoreTmaWarp ? 2 : (a.K > 1 ? 2 : 1)));
    const size_t smem = (size_t)nbuf * BLOCK * a.obs_dim * sizeof(float);
    static size_t smem_set = 0;  // per instantiation
    if (smem > 48 * 1024 && smem > smem_set) {
        cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
        smem_set = smem;
    }
    const unsigned grid = (unsigned)((a.count + BLOCK - 1) / BLOCK);
    kern<<<grid, BLOCK, smem, s>>>(a);
    return cudaGetLastError();
}

}  // namespace salp
