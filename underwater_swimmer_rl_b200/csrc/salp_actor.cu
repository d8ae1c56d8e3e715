// salp_actor.cu -- the rollout loop's policy step as ONE fused sm_100a kernel:
//   actions = tanh(mean + exp(clamp(log_std, -20, 2)) * noise),  (mean, log_std) = MLP(obs)
// for the reference's SAC actor shape (record_model.py:27-53: obs -> 256 -> 256 -> 2*act, ReLU), which is
// what `agent.select_action(obs)` costs per step in the reference's episode loop
// (src/salp/training/continuous_trainer.py:293) and what BASELINE.json configs[4] puts in the loop with
// the env kernel.  The observation tensor is read in place (f32 rows straight from the env kernel's
// output), the action tensor is written in the env build's real type -- nothing else touches HBM.
//
// Structure (one persistent CTA per SM, 512 threads, tiles of 128 environments, software-pipelined: the next
// tile's observations are prefetched into registers, its GEMM 1 runs behind this tile's GEMM 2 into a second
// TMEM accumulator and overlaps this tile's last epilogue):
//   weights   W1 (K padded 24->32) and W2 live in shared memory for the whole kernel as bf16, pre-packed
//             on the host into the tcgen05 canonical K-major no-swizzle layout (8x8 core matrices)
//   GEMM 1    [128 x 32] x [32 x 256]   2 x tcgen05.mma (M128 N256 K16), accumulator in TMEM (256 cols)
//   epilogue  tcgen05.ld -> +bias, ReLU -> bf16 -> shared memory, canonical A layout for GEMM 2
//   GEMM 2    [128 x 256] x [256 x 256] 16 x tcgen05.mma into the same TMEM columns
//   epilogue  tcgen05.ld -> +bias, ReLU -> dot with the 2*act rows of W3 in registers (f32) -> squash
// In the epilogues warp w reads TMEM lanes 32*(w%4).. (lane == row) and 64 of the 256 columns (w/4); one
// elected thread issues the MMAs and signals completion through mbarriers (tcgen05.commit).
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/salp_b200.h"

namespace salp { int set_error(int code, const char *msg); }

namespace {

constexpr int kTileM = 128, kHid = 256, kK1 = 32, kMaxOut = 8, kMaxAct = 4, kThreads = 512;
constexpr uint32_t kW2Bytes = kHid * kHid * 2, kW1Bytes = kHid * kK1 * 2, kA1Bytes = kTileM * kK1 * 2, kHBytes = kTileM * kHid * 2;
constexpr uint32_t kOffW2 = 0, kOffW1 = kOffW2 + kW2Bytes, kOffA1 = kOffW1 + kW1Bytes, kOffH = kOffA1 + kA1Bytes;
constexpr uint32_t kOffB1 = kOffH + kHBytes, kOffB2 = kOffB1 + kHid * 4, kOffW3 = kOffB2 + kHid * 4;
constexpr uint32_t kOffB3 = kOffW3 + kMaxOut * kHid * 4, kOffBar = kOffB3 + kMaxOut * 4, kOffTmem = kOffBar + 16;
static_assert(kOffTmem + 8 <= 227 * 1024, "shared memory of one CTA");
constexpr uint32_t kSmemBytes = kOffTmem + 8;
static_assert(kSmemBytes <= 232448, "shared memory budget of one sm_100 CTA");
constexpr uint32_t kTmemCols = 512;  // two 128 x 256 f32 accumulators

// instruction descriptor, kind::f16: D f32 (bit 4), A/B bf16 (bits 7, 10), both K-major, N >> 3 at 17, M >> 4 at 24
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kHid >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);

struct ActorArgs {
    const float *obs;      // [n][obs_dim]
    const uint4 *w1p;      // packed bf16 [kK1/8][256][8]
    const uint4 *w2p;      // packed bf16 [256/8][256][8]
    const float *b1, *b2;  // [256]
    const float *w3;       // [2*act][256] f32
    const float *b3;       // [2*act]
    const float *noise;    // [n][act] or null (deterministic)
    void *actions;         // [n][act] float or double
    const void *extra;     // [n][extra_dim] real (dtype): columns appended to obs (GAIL: the action), or null
    unsigned long long *rng;  // device {seed, draw counter, CTA ticket}: in-kernel Philox noise when noise == null
    long long n;
    int obs_dim, act_dim, f64, extra_dim, w3_rows, mode;
    int obs_vec4;          // rows of obs are 16-byte aligned float4 groups (obs_dim % 4 == 0)
    float act_scale[kMaxAct], act_shift[kMaxAct];  // action = tanh(.) * scale + shift (Box rescale; 1, 0 = identity)
};

enum { kModeActor = 0, kModeGailReward = 1 };

// Philox4x32-10 (Salmon et al., SC'11): counter (row, draw), key = seed -> 4 x 32 random bits
__device__ __forceinline__ uint4 philox4x32(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}
// two independent standard normals from 64 random bits (Box-Muller)
__device__ __forceinline__ float2 normal_pair(uint32_t a, uint32_t b) {
    const float u1 = ((float)a + 0.5f) * 2.3283064365386963e-10f, u2 = ((float)b + 0.5f) * 2.3283064365386963e-10f;
    const float r = sqrtf(-2.f * __logf(u1));
    float sn, cs;
    __sincosf(6.283185307179586f * u2, &sn, &cs);
    return make_float2(r * cs, r * sn);
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// shared-memory matrix descriptor, K-major, no swizzle: element (row, k) of an operand tile lives at
//   (k / 8) * lbo + (row / 8) * 128 + (row % 8) * 16 + (k % 8) * 2      [bytes]
// lbo = distance between the two 8-wide K chunks of one K=16 instruction, sbo = 128 = distance between
// 8-row groups; version 1 (bit 46), layout type 0.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) | ((uint64_t)(128u >> 4) << 32) |
           ((uint64_t)1 << 46);
}

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(kIdesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&v);
}

// obs chunk (8 consecutive features) of one row, prefetched into registers one tile ahead
struct ObsChunk { float4 lo, hi; };

__device__ __forceinline__ ObsChunk load_obs_chunk(const ActorArgs &a, long long row, int c) {
    ObsChunk v{make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
    if (row < a.n) {
        if (a.obs_vec4) {
            const float4 *src = reinterpret_cast<const float4 *>(a.obs + row * a.obs_dim);
            if (8 * c < a.obs_dim) v.lo = src[2 * c];
            if (8 * c + 4 < a.obs_dim) v.hi = src[2 * c + 1];
        } else if (8 * c < a.obs_dim) {  // rows that are not float4 groups (model B: 6 features): scalar loads
            float *f = reinterpret_cast<float *>(&v);
    #pragma unroll
            for (int j = 0; j < 8; ++j)
                if (8 * c + j < a.obs_dim) f[j] = a.obs[row * a.obs_dim + 8 * c + j];
        }
        if (a.extra && 8 * c + 8 > a.obs_dim && 8 * c < a.obs_dim + a.extra_dim) {  // appended columns fall in this chunk
            float *f = reinterpret_cast<float *>(&v);
    #pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int col = 8 * c + j - a.obs_dim;
                if (col >= 0 && col < a.extra_dim)
                    f[j] = a.f64 ? (float)reinterpret_cast<const double *>(a.extra)[row * a.extra_dim + col]
                                 : reinterpret_cast<const float *>(a.extra)[row * a.extra_dim + col];
            }
        }
    }
    return v;
}

template <int NOUT>  // rows of W3 padded to 2, 4 or 8
__global__ void __launch_bounds__(kThreads, 1) salp_actor_kernel(const __grid_constant__ ActorArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int quarter = warp & 3;   // TMEM lane quarter this warp may read: rows 32*quarter .. +31
    const int cgroup = warp >> 2;   // which 64 of the 256 hidden columns this warp handles in the epilogues
    const int erow = quarter * 32 + lane;            // row of the tile this thread owns in the epilogues
    const int lrow = tid & (kTileM - 1), lchunk = tid >> 7;  // (row, 8-feature chunk) this thread loads of the obs tile
    constexpr int n_out = NOUT;
    float *s_b1 = reinterpret_cast<float *>(smem + kOffB1), *s_b2 = reinterpret_cast<float *>(smem + kOffB2);
    float *s_w3 = reinterpret_cast<float *>(smem + kOffW3), *s_b3 = reinterpret_cast<float *>(smem + kOffB3);
    float *s_part = reinterpret_cast<float *>(smem + kOffH);  // epilogue-2 partial sums alias H (idle at that point)
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(smem + kOffTmem);
    const uint32_t bar1 = smem_u32(smem + kOffBar), bar2 = bar1 + 8;

    // ---- one-time: weights -> shared memory (already in operand layout), barriers, TMEM columns
    {
        uint4 *d2 = reinterpret_cast<uint4 *>(smem + kOffW2), *d1 = reinterpret_cast<uint4 *>(smem + kOffW1);
        for (int i = tid; i < (int)(kW2Bytes / 16); i += kThreads) d2[i] = a.w2p[i];
        for (int i = tid; i < (int)(kW1Bytes / 16); i += kThreads) d1[i] = a.w1p[i];
        for (int i = tid; i < kHid; i += kThreads) { s_b1[i] = a.b1[i]; s_b2[i] = a.b2[i]; }
        for (int i = tid; i < n_out * kHid; i += kThreads) s_w3[i] = i < a.w3_rows * kHid ? a.w3[i] : 0.f;
        if (tid < n_out) s_b3[tid] = tid < a.w3_rows ? a.b3[tid] : 0.f;
    }
    if (tid == 0) { mbar_init(bar1, 1); mbar_init(bar2, 1); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    const uint32_t sA1 = smem_u32(smem + kOffA1), sH = smem_u32(smem + kOffH), sW1 = smem_u32(smem + kOffW1), sW2 = smem_u32(smem + kOffW2);
    constexpr uint32_t kLboA = kTileM * 16, kLboB = kHid * 16;  // bytes between 8-wide K chunks (rows * 16)
    auto store_obs_chunk = [&](const ObsChunk &v) {  // chunk c of row r at c*2048 + r*16 (canonical A layout)
        *reinterpret_cast<uint4 *>(smem + kOffA1 + lchunk * kLboA + lrow * 16) =
            make_uint4(pack_bf16(v.lo.x, v.lo.y), pack_bf16(v.lo.z, v.lo.w), pack_bf16(v.hi.x, v.hi.y), pack_bf16(v.hi.z, v.hi.w));
    };
    const long long n_tiles = (a.n + kTileM - 1) / kTileM;
    long long tile = blockIdx.x;
    // ---- prologue: first observation tile -> A1, GEMM 1 of it in flight, second tile's obs in registers
    store_obs_chunk(load_obs_chunk(a, tile * kTileM + lrow, lchunk));
    fence_async_smem();  // weights and A1 were written with generic-proxy stores, the MMA reads through the async proxy
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;
    const uint32_t acc2 = tmem, acc1 = tmem + kHid;  // GEMM 2 / GEMM 1 accumulators: 256 f32 columns each
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    auto issue_gemm1 = [&]() {  // acc1[128 x 256] = A1[128 x 32] . W1^T
    #pragma unroll
        for (int kk = 0; kk < kK1 / 16; ++kk)
            umma_bf16(acc1, umma_desc(sA1 + 2 * kk * kLboA, kLboA), umma_desc(sW1 + 2 * kk * kLboB, kLboB), kk > 0);
        umma_commit(bar1);
    };
    if (tid == 0) issue_gemm1();
    ObsChunk nxt = load_obs_chunk(a, (tile + gridDim.x) * kTileM + lrow, lchunk);
    uint32_t phase1 = 0, phase2 = 0;
    const unsigned long long rng_seed = a.rng ? a.rng[0] : 0ull, rng_draw = a.rng ? a.rng[1] : 0ull;

    for (; tile < n_tiles; tile += gridDim.x) {
        const bool has_next = tile + gridDim.x < n_tiles;
        // ---- GEMM 1 of this tile is done: A1 is free again -> next tile's observations go in
        mbar_wait(bar1, phase1);
        phase1 ^= 1;
        tc_fence_after();
        if (has_next) store_obs_chunk(nxt);
        // ---- epilogue 1: h1 = relu(acc1 + b1) -> bf16 -> H (canonical A layout of GEMM 2); 64 columns per thread
        {
            uint32_t r[2][32];
            const int colg = cgroup * 64;
            tmem_ld32_nowait(acc1 + lane_base + colg, r[0]);
            tmem_ld32_nowait(acc1 + lane_base + colg + 32, r[1]);
            tmem_ld_wait();
    #pragma unroll
            for (int i = 0; i < 2; ++i) {
    #pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float4 ba = *reinterpret_cast<const float4 *>(s_b1 + colg + 32 * i + 8 * c);
                    const float4 bb = *reinterpret_cast<const float4 *>(s_b1 + colg + 32 * i + 8 * c + 4);
                    const uint32_t *q = &r[i][8 * c];
                    const float h0 = fmaxf(__uint_as_float(q[0]) + ba.x, 0.f), h1 = fmaxf(__uint_as_float(q[1]) + ba.y, 0.f);
                    const float h2 = fmaxf(__uint_as_float(q[2]) + ba.z, 0.f), h3 = fmaxf(__uint_as_float(q[3]) + ba.w, 0.f);
                    const float h4 = fmaxf(__uint_as_float(q[4]) + bb.x, 0.f), h5 = fmaxf(__uint_as_float(q[5]) + bb.y, 0.f);
                    const float h6 = fmaxf(__uint_as_float(q[6]) + bb.z, 0.f), h7 = fmaxf(__uint_as_float(q[7]) + bb.w, 0.f);
                    *reinterpret_cast<uint4 *>(smem + kOffH + ((colg + 32 * i) / 8 + c) * kLboA + erow * 16) =
                        make_uint4(pack_bf16(h0, h1), pack_bf16(h2, h3), pack_bf16(h4, h5), pack_bf16(h6, h7));
                }
            }
        }
        fence_async_smem();
        tc_fence_before();   // this tile's reads of acc1 and the previous tile's reads of acc2 are complete
        __syncthreads();
        // ---- GEMM 2 of this tile, then GEMM 1 of the next one right behind it (it overlaps epilogue 2)
        if (tid == 0) {
            tc_fence_after();
    #pragma unroll
            for (int kk = 0; kk < kHid / 16; ++kk)
                umma_bf16(acc2, umma_desc(sH + 2 * kk * kLboA, kLboA), umma_desc(sW2 + 2 * kk * kLboB, kLboB), kk > 0);
            umma_commit(bar2);
            if (has_next) issue_gemm1();
        }
        if (has_next) nxt = load_obs_chunk(a, (tile + 2 * (long long)gridDim.x) * kTileM + lrow, lchunk);
        mbar_wait(bar2, phase2);
        phase2 ^= 1;
        tc_fence_after();
        // ---- epilogue 2: h2 = relu(acc2 + b2); partial dot products with the 2*act rows of W3 over 64 columns
        constexpr int kOutRegs = NOUT < 4 ? 4 : NOUT;
        float out[kOutRegs];
    #pragma unroll
        for (int j = 0; j < kOutRegs; ++j) out[j] = 0.f;
        {
            uint32_t r[2][32];
            const int colg = cgroup * 64;
            tmem_ld32_nowait(acc2 + lane_base + colg, r[0]);
            tmem_ld32_nowait(acc2 + lane_base + colg + 32, r[1]);
            tmem_ld_wait();
    #pragma unroll
            for (int i = 0; i < 2; ++i) {
    #pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int col = colg + 32 * i + 4 * c;
                    const float4 b = *reinterpret_cast<const float4 *>(s_b2 + col);
                    const uint32_t *q = &r[i][4 * c];
                    const float h0 = fmaxf(__uint_as_float(q[0]) + b.x, 0.f), h1 = fmaxf(__uint_as_float(q[1]) + b.y, 0.f);
                    const float h2 = fmaxf(__uint_as_float(q[2]) + b.z, 0.f), h3 = fmaxf(__uint_as_float(q[3]) + b.w, 0.f);
    #pragma unroll
                    for (int o = 0; o < NOUT; ++o) {
                        const float4 w = *reinterpret_cast<const float4 *>(s_w3 + o * kHid + col);
                        out[o] = fmaf(h0, w.x, fmaf(h1, w.y, fmaf(h2, w.z, fmaf(h3, w.w, out[o]))));
                    }
                }
            }
        }
        // H is idle (GEMM 2 has completed): the four column groups meet there
        *reinterpret_cast<float4 *>(s_part + (cgroup * kTileM + erow) * 4) = make_float4(out[0], out[1], out[2], out[3]);
        if constexpr (NOUT > 4)  // outputs 4-7 of (column group g, row r) go to H chunk 4 + g of row r
            *reinterpret_cast<float4 *>(s_part + ((4 + cgroup) * kTileM + erow) * 4) = make_float4(out[4], out[5], out[6], out[7]);
        // only the four warps that share this quarter of the rows exchange partial sums: a 128-thread named barrier
        asm volatile("bar.sync %0, %1;" ::"r"(1 + quarter), "r"(kThreads / 4) : "memory");
        const long long row = tile * kTileM + erow;
        if (cgroup == 0 && row < a.n) {
            float4 t = *reinterpret_cast<const float4 *>(s_part + erow * 4);
    #pragma unroll
            for (int g = 1; g < kThreads / kTileM; ++g) {
                const float4 u = *reinterpret_cast<const float4 *>(s_part + (g * kTileM + erow) * 4);
                t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
            }
            float o8[kOutRegs] = {t.x, t.y, t.z, t.w};
            if constexpr (NOUT > 4) {
                float4 t2 = *reinterpret_cast<const float4 *>(s_part + (4 * kTileM + erow) * 4);
    #pragma unroll
                for (int g = 1; g < kThreads / kTileM; ++g) {
                    const float4 u = *reinterpret_cast<const float4 *>(s_part + ((4 + g) * kTileM + erow) * 4);
                    t2.x += u.x; t2.y += u.y; t2.z += u.z; t2.w += u.w;
                }
                o8[4] = t2.x; o8[5] = t2.y; o8[6] = t2.z; o8[7] = t2.w;
            }
            auto pick = [&](int idx) {  // o8[idx] without a local-memory array
                float v = o8[0];
    #pragma unroll
                for (int j = 1; j < kOutRegs; ++j) v = idx == j ? o8[j] : v;
                return v;
            };
            if (a.mode == kModeGailReward) {
                // D = sigmoid(logit); reward = -log(1 - D + 1e-8)   (src/salp/agents/discriminator.py:59-82)
                const float d = 1.f / (1.f + __expf(-(o8[0] + s_b3[0])));
                reinterpret_cast<float *>(a.actions)[row] = -__logf(1.f - d + 1e-8f);
            } else {
                float z[kMaxAct] = {0.f, 0.f, 0.f, 0.f};
                const bool sample = a.noise != nullptr || a.rng != nullptr;
                if (a.noise) {
    #pragma unroll
                    for (int d = 0; d < kMaxAct; ++d)
                        if (d < a.act_dim) z[d] = a.noise[row * a.act_dim + d];
                } else if (a.rng) {
                    const uint4 rnd = philox4x32(make_uint4((uint32_t)row, (uint32_t)(row >> 32), (uint32_t)rng_draw, (uint32_t)(rng_draw >> 32)),
                                                 make_uint2((uint32_t)rng_seed, (uint32_t)(rng_seed >> 32)));
                    const float2 z01 = normal_pair(rnd.x, rnd.y);
                    z[0] = z01.x; z[1] = z01.y;
                    if (a.act_dim > 2) { const float2 z23 = normal_pair(rnd.z, rnd.w); z[2] = z23.x; z[3] = z23.y; }
                }
    #pragma unroll
                for (int d = 0; d < kMaxAct; ++d) {
                    if (d >= a.act_dim || 2 * d >= NOUT) break;
                    const float mean = o8[d] + s_b3[d];
                    float pre = mean;
                    if (sample) {
                        const float log_std = fminf(fmaxf(pick(a.act_dim + d) + s_b3[a.act_dim + d], -20.f), 2.f);
                        pre = fmaf(__expf(log_std), z[d], mean);
                    }
                    const float act = fmaf(tanhf(pre), a.act_scale[d], a.act_shift[d]);
                    if (a.f64) reinterpret_cast<double *>(a.actions)[row * a.act_dim + d] = (double)act;
                    else reinterpret_cast<float *>(a.actions)[row * a.act_dim + d] = act;
                }
            }
        }
        // No barrier here: the partial sums of row r live in H chunks 0-3 of row r, which epilogue 1 of the next tile
        // rewrites from the very thread that has just read them (column group 0 of that row); every other reuse of H,
        // acc1 and acc2 is ordered by the CTA-wide barrier after epilogue 1.
    }
    // ---- teardown: every thread's TMEM reads are done before the columns are released; the last CTA to get
    //      here (every CTA read the draw counter at its start) advances the counter for the next launch
    tc_fence_before();
    __syncthreads();
    if (tid == 0 && a.rng) {
        if (atomicAdd(a.rng + 2, 1ull) == (unsigned long long)gridDim.x - 1) { a.rng[2] = 0ull; a.rng[1] = rng_draw + 1; }
    }
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols) : "memory");
    }
}

}  // namespace

extern "C" int salp_actor_forward(const SalpActorForward *p, void *cuda_stream) {
    if (!p) return salp::set_error(SALP_EINVAL, "null argument");
    if (!p->obs || !p->w1_packed || !p->w2_packed || !p->b1 || !p->b2 || !p->w3 || !p->b3 || !p->actions)
        return salp::set_error(SALP_EINVAL, "salp_actor_forward: a required pointer is null");
    if (p->hidden != kHid) return salp::set_error(SALP_EUNSUPPORTED, "salp_actor_forward: hidden size must be 256");
    if (p->mode != kModeActor && p->mode != kModeGailReward) return salp::set_error(SALP_EINVAL, "salp_actor_forward: bad mode");
    const int extra = p->extra ? p->extra_dim : 0;
    if (p->obs_dim <= 0 || extra < 0 || p->obs_dim + extra > kK1)
        return salp::set_error(SALP_EUNSUPPORTED, "salp_actor_forward: obs_dim + extra_dim must be in 1..32");
    if (p->mode == kModeActor && (p->act_dim < 1 || p->act_dim > kMaxAct))
        return salp::set_error(SALP_EUNSUPPORTED, "salp_actor_forward: act_dim must be in 1..4");
    if (p->n <= 0) return salp::set_error(SALP_EINVAL, "salp_actor_forward: n must be positive");
    if (p->dtype != SALP_F32 && p->dtype != SALP_F64) return salp::set_error(SALP_EINVAL, "salp_actor_forward: bad dtype");
    if (((uintptr_t)p->w1_packed | (uintptr_t)p->w2_packed) % 16 != 0 || (uintptr_t)p->obs % 4 != 0)
        return salp::set_error(SALP_EINVAL, "salp_actor_forward: packed weights must be 16-byte aligned, obs 4-byte aligned");
    ActorArgs a;
    a.obs = p->obs; a.w1p = reinterpret_cast<const uint4 *>(p->w1_packed); a.w2p = reinterpret_cast<const uint4 *>(p->w2_packed);
    a.b1 = p->b1; a.b2 = p->b2; a.w3 = p->w3; a.b3 = p->b3; a.noise = p->noise; a.actions = p->actions;
    a.extra = p->extra; a.extra_dim = extra; a.rng = reinterpret_cast<unsigned long long *>(p->rng_state);
    a.n = p->n; a.obs_dim = p->obs_dim; a.act_dim = p->act_dim; a.f64 = p->dtype == SALP_F64; a.mode = p->mode;
    a.w3_rows = p->mode == kModeGailReward ? 1 : 2 * p->act_dim;
    a.obs_vec4 = (p->obs_dim % 4 == 0) && ((uintptr_t)p->obs % 16 == 0);
    for (int d = 0; d < kMaxAct; ++d) {  // SB3's unscale_action: [-1, 1] -> [low, high] of the env's Box
        const bool on = p->act_rescale && d < p->act_dim;
        a.act_scale[d] = on ? 0.5f * (p->act_high[d] - p->act_low[d]) : 1.f;
        a.act_shift[d] = on ? 0.5f * (p->act_high[d] + p->act_low[d]) : 0.f;
    }
    if (p->mode == kModeGailReward) { a.noise = nullptr; a.rng = nullptr; }
    auto kern = a.w3_rows <= 2 ? salp_actor_kernel<2> : a.w3_rows <= 4 ? salp_actor_kernel<4> : salp_actor_kernel<8>;
    {   // per device and cheap: set on every call rather than caching a per-process flag
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
        if (e != cudaSuccess) return salp::set_error(SALP_ECUDA, cudaGetErrorString(e));
    }
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long n_tiles = (a.n + kTileM - 1) / kTileM;
    const unsigned grid = (unsigned)(n_tiles < sms ? n_tiles : sms);  // persistent: one CTA per SM
    kern<<<grid, kThreads, kSmemBytes, (cudaStream_t)cuda_stream>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return salp::set_error(SALP_ECUDA, cudaGetErrorString(e));
    return SALP_OK;
}
