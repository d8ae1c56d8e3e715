// Developer probe: which (lane, column) does each thread/register of tcgen05.ld.16x256b.x8 read?
// One warp-group of 4 warps fills 128 lanes x 64 columns of TMEM with lane * 1000 + column through 32x32b stores, then warp w
// reads lanes 32 w + 16 h .. with the 16x256b shape and prints what thread t got in register k.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/_lab/probe_tmem tools/probe_tmem_shapes.cu && tools/_lab/probe_tmem
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void probe(uint32_t *out) {
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem)), "r"(64u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    for (int c = 0; c < 64; c += 4) {
        const uint32_t v = (uint32_t)(warp * 32 + lane) * 1000u + c;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(tmem + lane_base + c), "r"(v), "r"(v + 1), "r"(v + 2), "r"(v + 3) : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (int h = 0; h < 2; ++h) {
        uint32_t r[32];
        asm volatile(
            "tcgen05.ld.sync.aligned.16x256b.x8.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
              "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
              "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
              "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
            : "r"(tmem + lane_base + ((uint32_t)(16 * h) << 16))
            : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int k = 0; k < 32; ++k) out[((warp * 2 + h) * 32 + lane) * 32 + k] = r[k];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64u) : "memory");
}

int main() {
    uint32_t *d, *h = new uint32_t[4 * 2 * 32 * 32];
    cudaMalloc(&d, 4 * 2 * 32 * 32 * 4);
    probe<<<1, 128>>>(d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
    cudaMemcpy(h, d, 4 * 2 * 32 * 32 * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int w = 0; w < 4; ++w)
        for (int hh = 0; hh < 2; ++hh)
            for (int t = 0; t < 32; ++t)
                for (int k = 0; k < 32; ++k) {
                    const uint32_t v = h[((w * 2 + hh) * 32 + t) * 32 + k];
                    const int lane = v / 1000, col = v % 1000;
                    const int j = k >> 2, e = k & 1, up = (k >> 1) & 1;
                    const int want_lane = 32 * w + 16 * hh + t / 4 + 8 * up, want_col = 8 * j + 2 * (t % 4) + e;
                    if (lane != want_lane || col != want_col) {
                        if (bad < 12) printf("warp %d half %d thread %2d reg %2d: lane %3d col %2d (expected lane %3d col %2d)\n", w, hh, t, k, lane, col, want_lane, want_col);
                        ++bad;
                    }
                }
    printf(bad ? "MISMATCHES: %d\n" : "layout as expected: reg 4j+2u+e of thread t = (lane 16h + t/4 + 8u, column 8j + 2(t%%4) + e)%d\n", bad);
    // also dump thread 0..5 of warp 0 half 0, first 8 registers, for the record
    for (int t = 0; t < 6; ++t) {
        printf("w0 h0 t%d:", t);
        for (int k = 0; k < 8; ++k) printf(" (%u,%u)", h[t * 32 + k] / 1000, h[t * 32 + k] % 1000);
        printf("\n");
    }
    return 0;
}
