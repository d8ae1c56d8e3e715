// salp_actor.cu -- the rollout loop's policy step as ONE fused sm_100a kernel:
//   actions = tanh(mean + exp(clamp(log_std, -20, 2)) * noise),  (mean, log_std) = MLP(obs)
// for the reference's SAC actor shape (record_model.py:27-53: obs -> 256 -> 256 -> 2*act, ReLU), which is
// what `agent.select_action(obs)` costs per step in the reference's episode loop
// (src/salp/training/continuous_trainer.py:293) and what BASELINE.json configs[4] puts in the loop with
// the env kernel.  The observation tensor is read in place (f32 rows straight from the env kernel's
// output), the action tensor is written in the env build's real type -- nothing else touches HBM.
//
// Structure (one persistent CTA per SM, 512 threads, tiles of 128 environments):
//   weights   W1 (K padded 24->32) and W2 live in shared memory for the whole kernel as bf16: W1 in the tcgen05
//             K-major no-swizzle layout the host packs (8x8 core matrices), W2 re-laid into 128-byte-swizzled K blocks
//   TMEM      two slots of 128 x 256 f32; tile i keeps BOTH its GEMM outputs in slot i & 1
//   GEMM 1    [128 x 32] x [32 x 256]    2 x tcgen05.mma (M128 N256 K16)
//   epilogue1 tcgen05.ld -> +bias (FADD2), ReLU folded into cvt.bf16x2 -> H in shared memory (swizzled A layout)
//   GEMM 2    [128 x 256] x [256 x 256]  16 x tcgen05.mma over the same slot, issued in two halves by two warps
//   epilogue2 tcgen05.ld -> +bias, ReLU -> FFMA2 dot with the 2*act rows of W3 (f32) -> sums of the 4 column groups
//             meet in the idle A1 buffer (8 output rows: in TMEM) -> squash -> store
//   pipeline  per iteration: epilogue 1 (tile i) -> epilogue 2 (tile i-1), while the tensor core runs GEMM 2 (tile i)
//             and GEMM 1 (tile i+1); hand-offs are bar.arrive / bar.sync pairs and tcgen05.commit -> mbarrier
// In the epilogues warp w reads TMEM lanes 32*(w%4).. (lane == row) and 64 of the 256 columns (w/4).
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/salp_b200.h"

namespace salp { int set_error(int code, const char *msg); }

namespace {

constexpr int kTileM = 128, kHid = 256, kK1 = 32, kMaxOut = 8, kMaxAct = 4;
constexpr int kThreads = 512;  // 16 warps: TMEM lane quarter = warp % 4, column group = warp / 4
// The MMAs are issued by lane 0 of ONE of the 16 warps.  Its issue of a whole GEMM 2 blocks for about the GEMM's duration
// (the issue queue is short), so that warp is the slowest of every iteration; it is taken from column group 3 because
// column group 0 also squashes and stores the actions.  (A 17th warp that only issues was measured slower: 544 threads
// cap the kernel at 96 registers, and its MMAs paced at ~265 instead of ~155 cycles -- profiles/r01_ncu_actor.md.)
constexpr int kIssueWarp = 12, kIssueTid = 32 * kIssueWarp, kIssueQuarter = kIssueWarp & 3, kIssueGroup = kIssueWarp >> 2;
// With at most 4 output rows (kEarlyG1 below) a second warp issues the second half of GEMM 2, so that neither thread sits
// in the issue for the whole GEMM: warp 13, the same column group in the next quarter.
constexpr int kIssueWarpB = kIssueWarp + 1, kIssueTidB = 32 * kIssueWarpB;
static_assert(kIssueGroup != 0 && kIssueQuarter == 0, "the issuing warps: not ones that store the actions; quarters 0 and 1");
constexpr uint32_t kW2Bytes = kHid * kHid * 2, kW1Bytes = kHid * kK1 * 2, kA1Bytes = kTileM * kK1 * 2, kHBytes = kTileM * kHid * 2;
constexpr uint32_t kOffW2 = 0, kOffW1 = kOffW2 + kW2Bytes, kOffA1 = kOffW1 + kW1Bytes, kOffH = kOffA1 + kA1Bytes;
constexpr uint32_t kOffB1 = kOffH + kHBytes, kOffB2 = kOffB1 + kHid * 4, kOffW3 = kOffB2 + kHid * 4;
constexpr uint32_t kOffB3 = kOffW3 + kMaxOut * kHid * 4, kOffBar = kOffB3 + kMaxOut * 4, kOffTmem = kOffBar + 16;
constexpr uint32_t kOffIss = kOffW3 + 4 * kHid * 4;  // kEarlyG1 (W3 has at most 4 rows): the issuing warps' output sums, 2 x 32 rows x 16 bytes
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

#ifdef SALP_ACTOR_TRACE  // developer build (SALP_ACTOR_TRACE=1 python -m ...build): per-phase clocks of CTA 0, warps 0 and kIssueWarp
constexpr int kTraceIters = 12, kTracePoints = 8;
__device__ long long g_trace[2 * kTraceIters * kTracePoints + 8];  // + (clock, globaltimer) at entry / loop start / end
__device__ __forceinline__ long long gtimer() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define TRACE2(k) do { if (blockIdx.x == 0 && threadIdx.x == 0) { g_trace[2 * kTraceIters * kTracePoints + 2 * (k)] = clock64(); \
    g_trace[2 * kTraceIters * kTracePoints + 2 * (k) + 1] = gtimer(); } } while (0)
#define TRACE(pt) do { if (blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == kIssueWarp) && it < kTraceIters) \
    g_trace[((warp != 0) * kTraceIters + it) * kTracePoints + (pt)] = clock64(); } while (0)
#else
#define TRACE(pt) do { } while (0)
#define TRACE2(k) do { } while (0)
#endif

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

// The same for the 128-byte swizzled K-major layout GEMM 2 reads (the layout the tensor core fetches at full rate:
// measured here, the no-swizzle layout above paces a 128 x 256 x 16 MMA at ~300 cycles instead of 128).  An operand
// of R rows is cut into K blocks of 64 elements; a block is R rows of 128 bytes and the 16-byte chunk c of row r sits
// at chunk position c ^ (r % 8).  sbo = 1024 (8 rows), layout type 2 (bits 61-63); the K = 16 steps inside a block
// advance the start address by 32 bytes.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024u >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)2 << 61);
}
__device__ __forceinline__ uint32_t sw128_offset(int rows, int row, int kchunk) {  // byte offset of 16-byte chunk kchunk of a row
    return (uint32_t)((kchunk >> 3) * rows * 128 + row * 128 + (((kchunk & 7) ^ (row & 7)) << 4));
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
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// four consecutive columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_st4(uint32_t taddr, float a, float b, float c, float d) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(__float_as_uint(a)),
                 "r"(__float_as_uint(b)), "r"(__float_as_uint(c)), "r"(__float_as_uint(d))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld4_nowait(uint32_t taddr, float &a, float &b, float &c, float &d) {
    uint32_t x, y, z, w;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(x), "=r"(y), "=r"(z), "=r"(w) : "r"(taddr) : "memory");
    a = __uint_as_float(x); b = __uint_as_float(y); c = __uint_as_float(z); d = __uint_as_float(w);
}
// meeting point on named barrier ID where only the waiting warps block: the others arrive and go on
template <int ID>
__device__ __forceinline__ void cta_meet(bool wait) {
    if (wait) asm volatile("bar.sync %0, %1;" ::"n"(ID), "n"(kThreads) : "memory");
    else asm volatile("bar.arrive %0, %1;" ::"n"(ID), "n"(kThreads) : "memory");
}
// the four warps that own the same 32 rows: named barriers base + quarter (base 1: sums written, base 8: sums consumed)
__device__ __forceinline__ void quarter_meet(int base, int quarter, bool wait) {
    if (wait) asm volatile("bar.sync %0, %1;" ::"r"(base + quarter), "n"(kThreads / 4) : "memory");
    else asm volatile("bar.arrive %0, %1;" ::"r"(base + quarter), "n"(kThreads / 4) : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&v);
}

// bf16x2 of (max(lo, 0), max(hi, 0)): the ReLU rides on the conversion
__device__ __forceinline__ uint32_t pack_relu_bf16(float lo, float hi) {
    uint32_t d;
    asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
}
__device__ __forceinline__ float2 u2f2(uint32_t a, uint32_t b) { return make_float2(__uint_as_float(a), __uint_as_float(b)); }

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
    TRACE2(0);
    float *s_b1 = reinterpret_cast<float *>(smem + kOffB1), *s_b2 = reinterpret_cast<float *>(smem + kOffB2);
    float *s_w3 = reinterpret_cast<float *>(smem + kOffW3), *s_b3 = reinterpret_cast<float *>(smem + kOffB3);
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(smem + kOffTmem);
    const uint32_t bar1 = smem_u32(smem + kOffBar), bar2 = bar1 + 8;

    // ---- one-time: weights -> shared memory (already in operand layout), barriers, TMEM columns
    {
        uint4 *d1 = reinterpret_cast<uint4 *>(smem + kOffW1);
        // all of a thread's loads are issued before its first store (the loops have 16 and 2 trips): one L2 round trip, not 18
        static_assert(kW2Bytes / 16 % kThreads == 0 && kW1Bytes / 16 % kThreads == 0, "weight chunks per thread");
        uint4 w2v[kW2Bytes / 16 / kThreads], w1v[kW1Bytes / 16 / kThreads];
    #pragma unroll
        for (int j = 0; j < (int)(kW2Bytes / 16 / kThreads); ++j) w2v[j] = a.w2p[tid + j * kThreads];
    #pragma unroll
        for (int j = 0; j < (int)(kW1Bytes / 16 / kThreads); ++j) w1v[j] = a.w1p[tid + j * kThreads];
    #pragma unroll
        for (int j = 0; j < (int)(kW2Bytes / 16 / kThreads); ++j) {  // [K/8][256] chunks -> swizzled K blocks
            const int i = tid + j * kThreads;
            *reinterpret_cast<uint4 *>(smem + kOffW2 + sw128_offset(kHid, i & (kHid - 1), i >> 8)) = w2v[j];
        }
    #pragma unroll
        for (int j = 0; j < (int)(kW1Bytes / 16 / kThreads); ++j) d1[tid + j * kThreads] = w1v[j];
        for (int i = tid; i < kHid; i += kThreads) { s_b1[i] = a.b1[i]; s_b2[i] = a.b2[i]; }
        for (int i = tid; i < n_out * kHid; i += kThreads) s_w3[i] = i < a.w3_rows * kHid ? a.w3[i] : 0.f;
        if (tid < n_out) s_b3[tid] = tid < a.w3_rows ? a.b3[tid] : 0.f;
    }
    if (tid == 0) { mbar_init(bar1, 1); mbar_init(bar2, NOUT <= 4 ? 2 : 1); }  // kEarlyG1: GEMM 2 is committed in two halves
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
    // Programmatic dependent launch: everything above reads parameter data only (weights, biases -- never written by the
    // kernels that trigger early, the env step kernels), so a CTA may do it while the previous kernel of the stream -- the
    // env step whose observations it is about to read -- is still draining.  Observations, noise, the RNG state and the
    // action buffer are touched only after the wait.  The kernel after this one (an env step: small CTAs) may start its own
    // prologue as soon as this CTA's resources free up.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    // ---- prologue: first observation tile -> A1, GEMM 1 of it in flight, second tile's obs in registers
    store_obs_chunk(load_obs_chunk(a, tile * kTileM + lrow, lchunk));
    fence_async_smem();  // weights and A1 were written with generic-proxy stores, the MMA reads through the async proxy
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    // Two TMEM slots of 256 f32 columns.  Tile i lives in slot i & 1: GEMM 1 writes its pre-activations there, epilogue 1
    // turns them into H, GEMM 2 then overwrites the same columns (the pre-activations are dead by then).
    auto issue_gemm1 = [&](uint32_t slot) {  // slot[128 x 256] = A1[128 x 32] . W1^T
    #pragma unroll
        for (int kk = 0; kk < kK1 / 16; ++kk)
            umma_bf16(slot, umma_desc(sA1 + 2 * kk * kLboA, kLboA), umma_desc(sW1 + 2 * kk * kLboB, kLboB), kk > 0);
        umma_commit(bar1);
    };
    uint32_t phase1 = 0, phase2 = 0;
    const unsigned long long rng_seed = a.rng ? a.rng[0] : 0ull, rng_draw = a.rng ? a.rng[1] : 0ull;
    constexpr int kOutRegs = NOUT < 4 ? 4 : NOUT;

    auto issue_gemm2 = [&](uint32_t slot, int k0, int k1) {  // slot[128 x 256] (+)= H[:, 16 k0 .. 16 k1) . W2[:, same]^T
    #pragma unroll
        for (int kk = 0; kk < kHid / 16; ++kk)
            if (kk >= k0 && kk < k1)
                umma_bf16(slot, umma_desc_sw128(sH + (kk >> 2) * (kTileM * 128) + (kk & 3) * 32),
                          umma_desc_sw128(sW2 + (kk >> 2) * (kHid * 128) + (kk & 3) * 32), kk > 0);
        umma_commit(bar2);
    };
    // ---- epilogue 2, first part: h2 = relu(slot + b2); partial dot products with the 2*act rows of W3 over this warp's 64
    //      columns; the four warps that own the same 32 rows meet in TMEM (each thread parks its partial sums in the first
    //      columns it has just read -- shared memory has no room left and H is being read by the next tile's GEMM 2);
    //      column group 0 leaves with the row totals in o8.
    // With at most 4 output rows the sums can meet in 16 bytes per (column group, row): they go through the A1 buffer, which
    // frees the other TMEM slot as soon as every warp has LOADED its part of it -- GEMM 1 of the next tile is then issued
    // right behind GEMM 2 instead of after epilogue 2 (kEarlyG1).  8 output rows keep the exchange in TMEM.
    constexpr bool kEarlyG1 = NOUT <= 4;
    // second half of GEMM 2 (kEarlyG1), by lane 0 of warp 13 once warp 12 has issued the first (named barrier 12, 64 threads)
    auto second_half = [&](uint32_t gemm_slot) {
        asm volatile("bar.sync 12, 64;" ::: "memory");
        if (tid == kIssueTidB) { tc_fence_after(); issue_gemm2(gemm_slot, kHid / 32, kHid / 16); }
        __syncwarp();
    };
    // `issue_mid`: warp 13 issues its half of the running GEMM 2 (into gemm_slot) between the two halves of its columns
    // `announce`: 0 nothing, 1 report that this warp's part of the slot is in registers (named barrier 7), 2 (warp 12) wait for
    // all of those reports and issue GEMM 1 of the next tile into the slot (`g1_next`)
    auto epilogue2_math = [&](uint32_t slot, float (&out)[kOutRegs], int announce, bool issue_mid, uint32_t gemm_slot, bool g1_next) {
    #pragma unroll
        for (int j = 0; j < kOutRegs; ++j) out[j] = 0.f;
        const int colg = cgroup * 64;
        uint32_t r[2][32];
        tmem_ld32_nowait(slot + lane_base + colg, r[0]);
        tmem_ld32_nowait(slot + lane_base + colg + 32, r[1]);
        tmem_ld_wait();
        if (announce == 1) { tc_fence_before(); cta_meet<7>(false); }
        if (announce == 2) {
            tc_fence_before();
            cta_meet<7>(true);
            if (tid == kIssueTid && g1_next) { tc_fence_after(); issue_gemm1(slot); }
            __syncwarp();
        }
        float2 acc[NOUT];  // packed f32x2 accumulators (even, odd columns), folded at the end
    #pragma unroll
        for (int o = 0; o < NOUT; ++o) acc[o] = make_float2(0.f, 0.f);
    #pragma unroll
        for (int i = 0; i < 2; ++i) {
            if (i == 1 && issue_mid) second_half(gemm_slot);
    #pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int col = colg + 32 * i + 4 * c;
                const float4 b = *reinterpret_cast<const float4 *>(s_b2 + col);
                const uint32_t *q = &r[i][4 * c];
                float2 h01 = __fadd2_rn(u2f2(q[0], q[1]), make_float2(b.x, b.y)), h23 = __fadd2_rn(u2f2(q[2], q[3]), make_float2(b.z, b.w));
                h01.x = fmaxf(h01.x, 0.f); h01.y = fmaxf(h01.y, 0.f); h23.x = fmaxf(h23.x, 0.f); h23.y = fmaxf(h23.y, 0.f);
    #pragma unroll
                for (int o = 0; o < NOUT; ++o) {
                    const float4 w = *reinterpret_cast<const float4 *>(s_w3 + o * kHid + col);
                    acc[o] = __ffma2_rn(h01, make_float2(w.x, w.y), acc[o]);
                    acc[o] = __ffma2_rn(h23, make_float2(w.z, w.w), acc[o]);
                }
            }
        }
    #pragma unroll
        for (int o = 0; o < NOUT; ++o) out[o] = acc[o].x + acc[o].y;
    };
    // the four column groups of a row meet: column group 0 leaves with the row totals in o8
    auto exchange_tmem = [&](uint32_t slot, const float (&out)[kOutRegs], float (&o8)[kOutRegs]) {
        const int colg = cgroup * 64;
        if (cgroup != 0) {  // parked in the first columns this thread has just read
            tmem_st4(slot + lane_base + colg, out[0], out[1], out[2], out[3]);
            if constexpr (NOUT > 4) tmem_st4(slot + lane_base + colg + 4, out[4], out[5], out[6], out[7]);
            tmem_st_wait();
        }
        tc_fence_before();
        quarter_meet(1, quarter, cgroup == 0);  // only column group 0 reads the others' sums: the others arrive and go on
        tc_fence_after();
    #pragma unroll
        for (int j = 0; j < kOutRegs; ++j) o8[j] = out[j];
        if (cgroup == 0) {
            float u[3][kOutRegs];
    #pragma unroll
            for (int g = 1; g < kThreads / kTileM; ++g) {
                tmem_ld4_nowait(slot + lane_base + 64 * g, u[g - 1][0], u[g - 1][1], u[g - 1][2], u[g - 1][3]);
                if constexpr (NOUT > 4) tmem_ld4_nowait(slot + lane_base + 64 * g + 4, u[g - 1][4], u[g - 1][5], u[g - 1][6], u[g - 1][7]);
            }
            tmem_ld_wait();
    #pragma unroll
            for (int g = 0; g < 3; ++g)
    #pragma unroll
                for (int j = 0; j < kOutRegs; ++j) o8[j] += u[g][j];
        }
    };
    // kEarlyG1: through A1 (idle between GEMM 1 of the next tile and the observation store of the one after): the sums of
    // (column group g, row r) take the 16 bytes of chunk g of row r -- exactly what the same thread stores there next.
    // `more` = another exchange or observation store follows: column group 0 then reports when it has read the sums.
    // The issuing warp has a 512-byte buffer of its own: it computes its sums a whole phase earlier than the others and
    // must not wait for A1 (that is, for GEMM 1 of the next tile) on its way back into the MMA issue.
    float4 *s_iss = reinterpret_cast<float4 *>(smem + kOffIss);
    auto issuer_publish = [&](const float (&out)[kOutRegs]) {
        s_iss[quarter * 32 + lane] = make_float4(out[0], out[1], out[2], out[3]);  // quarter 0 (warp 12) or 1 (warp 13)
        __threadfence_block();
        quarter_meet(1, quarter, false);
    };
    auto exchange_smem = [&](const float (&out)[kOutRegs], float (&o8)[kOutRegs], bool more) {
        float4 *sc = reinterpret_cast<float4 *>(smem + kOffA1);
        if (cgroup != 0) { sc[cgroup * kTileM + erow] = make_float4(out[0], out[1], out[2], out[3]); __threadfence_block(); }
        quarter_meet(1, quarter, cgroup == 0);
    #pragma unroll
        for (int j = 0; j < kOutRegs; ++j) o8[j] = out[j];
        if (cgroup == 0) {
    #pragma unroll
            for (int g = 1; g < kThreads / kTileM; ++g) {
                const bool from_issuer = g == kIssueGroup && (quarter == kIssueQuarter || quarter == kIssueQuarter + 1);
                const float4 u = from_issuer ? s_iss[quarter * 32 + lane] : sc[g * kTileM + erow];
                o8[0] += u.x; o8[1] += u.y; o8[2] += u.z; o8[3] += u.w;
            }
            if (more) quarter_meet(8, quarter, false);
        }
    };
    // the row's standard-normal draws, fetched (or drawn) before epilogue 2 so that their latency hides behind it
    auto draw_noise = [&](long long row, float (&z)[kMaxAct]) {
    #pragma unroll
        for (int d = 0; d < kMaxAct; ++d) z[d] = 0.f;
        if (a.mode != kModeActor || row >= a.n) return;
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
    };
    auto epilogue2_store = [&](long long row, const float (&o8)[kOutRegs], const float (&z)[kMaxAct]) {
        if (row >= a.n) return;
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
            const bool sample = a.noise != nullptr || a.rng != nullptr;
    #pragma unroll
            for (int d = 0; d < kMaxAct; ++d) {
                if (d >= a.act_dim || 2 * d >= NOUT) break;
                const float mean = o8[d] + s_b3[d];
                float pre = mean;
                if (sample) {
                    const float log_std = fminf(fmaxf(pick(a.act_dim + d) + s_b3[a.act_dim + d], -20.f), 2.f);
                    pre = fmaf(__expf(log_std), z[d], mean);
                }
                // tanh(x) = 1 - 2 / (exp(2x) + 1): absolute error ~2e-7, no branches (exp overflow -> 1, underflow -> -1)
                const float th = 1.f - __fdividef(2.f, __expf(2.f * pre) + 1.f);
                const float act = fmaf(th, a.act_scale[d], a.act_shift[d]);
                if (a.f64) reinterpret_cast<double *>(a.actions)[row * a.act_dim + d] = (double)act;
                else reinterpret_cast<float *>(a.actions)[row * a.act_dim + d] = act;
            }
        }
    };

    // Per iteration the warps run  epilogue 1 (tile i)  ->  epilogue 2 (tile i-1)  while the tensor core runs GEMM 2 (tile i)
    // behind epilogue 1 and GEMM 1 (tile i+1) behind epilogue 2: epilogue 2, the longer of the two, overlaps GEMM 2 of the
    // following tile, and one H buffer is enough because epilogue 1 (tile i+1) starts only when GEMM 2 (tile i) has finished
    // reading it.  The issuing warp does its share of epilogue 2 FIRST (alone, while the others are in epilogue 1), because
    // after the meeting point its lane 0 sits in the MMA issue for the length of GEMM 2.  The two CTA-wide meeting points are
    // arrive/sync pairs on named barriers: only the issuing warp waits, the others arrive and go on to their next epilogue.
    long long prev_tile = -1;
    uint32_t it = 0;
    bool sums_in_a1 = false;  // kEarlyG1: the last exchange's sums may still be unread in A1
    TRACE2(1);
    if (tid == kIssueTid) issue_gemm1(tmem);
    ObsChunk nxt = load_obs_chunk(a, (tile + gridDim.x) * kTileM + lrow, lchunk);
    for (; tile < n_tiles; tile += gridDim.x, ++it) {
        const bool has_next = tile + gridDim.x < n_tiles, has_prev = prev_tile >= 0;
        const uint32_t slot = tmem + (it & 1) * kHid, slot_prev = tmem + ((it & 1) ^ 1) * kHid;
        float out[kOutRegs], o8[kOutRegs], z[kMaxAct];
        // ---- GEMM 1 of this tile is done (A1 is free: the next tile's observations go in; with kEarlyG1 that was awaited
        //      at the end of the previous iteration); GEMM 2 of the previous tile is done (H is free, its output waits in
        //      the other slot)
        TRACE(0);
        if (!kEarlyG1 || it == 0 || warp == kIssueWarp) { mbar_wait(bar1, phase1); phase1 ^= 1; }  // (warp 12 issues it: waits here)
        if (has_prev) { mbar_wait(bar2, phase2); phase2 ^= 1; }
        tc_fence_after();
        TRACE(1);
        if constexpr (!kEarlyG1) {  // the one issuing warp does its epilogue 2 first: it will sit in the issue of all of GEMM 2
            if (has_prev && warp == kIssueWarp) { epilogue2_math(slot_prev, out, 0, false, 0u, false); exchange_tmem(slot_prev, out, o8); }
        }
        TRACE(2);
        if (kEarlyG1 && sums_in_a1 && cgroup != 0) quarter_meet(8, quarter, true);  // column group 0 has read the last sums
        sums_in_a1 = false;
        if (has_next) store_obs_chunk(nxt);
        // ---- epilogue 1: h1 = relu(slot + b1) -> bf16 -> H (swizzled A layout of GEMM 2); 64 columns per thread
        {
            uint32_t r[2][32];
            const int colg = cgroup * 64;
            tmem_ld32_nowait(slot + lane_base + colg, r[0]);
            tmem_ld32_nowait(slot + lane_base + colg + 32, r[1]);
            tmem_ld_wait();
    #pragma unroll
            for (int i = 0; i < 2; ++i) {
    #pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float4 ba = *reinterpret_cast<const float4 *>(s_b1 + colg + 32 * i + 8 * c);
                    const float4 bb = *reinterpret_cast<const float4 *>(s_b1 + colg + 32 * i + 8 * c + 4);
                    const uint32_t *q = &r[i][8 * c];
                    // packed f32x2 adds (FADD2) and the ReLU folded into the bf16x2 conversion: one issue slot per element
                    const float2 x01 = __fadd2_rn(u2f2(q[0], q[1]), make_float2(ba.x, ba.y)), x23 = __fadd2_rn(u2f2(q[2], q[3]), make_float2(ba.z, ba.w));
                    const float2 x45 = __fadd2_rn(u2f2(q[4], q[5]), make_float2(bb.x, bb.y)), x67 = __fadd2_rn(u2f2(q[6], q[7]), make_float2(bb.z, bb.w));
                    *reinterpret_cast<uint4 *>(smem + kOffH + sw128_offset(kTileM, erow, (colg + 32 * i) / 8 + c)) =
                        make_uint4(pack_relu_bf16(x01.x, x01.y), pack_relu_bf16(x23.x, x23.y), pack_relu_bf16(x45.x, x45.y), pack_relu_bf16(x67.x, x67.y));
                }
            }
        }
        fence_async_smem();
        tc_fence_before();   // this tile's reads of its slot are complete
        TRACE(3);
        cta_meet<5>(warp == kIssueWarp);
        // GEMM 2 of this tile over its own slot: all of it, or (kEarlyG1) its first half -- warp 13 issues the second one
        if (tid == kIssueTid) { tc_fence_after(); issue_gemm2(slot, 0, kEarlyG1 ? kHid / 32 : kHid / 16); }
        __syncwarp();
        TRACE(4);
        if constexpr (kEarlyG1) {
            if (warp == kIssueWarp) {
                asm volatile("bar.arrive 12, 64;" ::: "memory");
                if (has_next) nxt = load_obs_chunk(a, (tile + 2 * (long long)gridDim.x) * kTileM + lrow, lchunk);
                // its part of the other slot into registers, then (every warp has done the same) GEMM 1 of the next tile
                // into that slot, then the arithmetic while the second half of GEMM 2 and that GEMM 1 run
                if (has_prev) epilogue2_math(slot_prev, out, 2, false, 0u, has_next);
                else {
                    cta_meet<7>(true);
                    if (tid == kIssueTid && has_next) { tc_fence_after(); issue_gemm1(slot_prev); }
                    __syncwarp();
                }
            } else {
                if (has_next) nxt = load_obs_chunk(a, (tile + 2 * (long long)gridDim.x) * kTileM + lrow, lchunk);
                if (has_prev && cgroup == 0) draw_noise(prev_tile * kTileM + erow, z);
                // ---- epilogue 2 of the previous tile (other slot) while GEMM 2 of this one runs; warp 13 issues its half
                //      of that GEMM between the two halves of its columns (warp 12 is through with the first by then)
                if (has_prev) {
                    epilogue2_math(slot_prev, out, 1, warp == kIssueWarpB, slot, false);
                } else {
                    tc_fence_before(); cta_meet<7>(false);
                    if (warp == kIssueWarpB) second_half(slot);
                }
            }
            TRACE(5);
            if ((warp == kIssueWarp || warp == kIssueWarpB) && has_prev) issuer_publish(out);  // buffers of their own: no wait for A1
            if (warp != kIssueWarp) {
                if (has_next) { mbar_wait(bar1, phase1); phase1 ^= 1; }  // GEMM 1 of the next tile has read A1: the sums may go in
                if (has_prev && warp != kIssueWarpB) {
                    exchange_smem(out, o8, true);
                    if (cgroup == 0) epilogue2_store(prev_tile * kTileM + erow, o8, z);
                }
            }
            sums_in_a1 = has_prev;
        } else {
            if (has_next) nxt = load_obs_chunk(a, (tile + 2 * (long long)gridDim.x) * kTileM + lrow, lchunk);
            if (has_prev && cgroup == 0) draw_noise(prev_tile * kTileM + erow, z);
            if (has_prev && warp != kIssueWarp) { epilogue2_math(slot_prev, out, 0, false, 0u, false); exchange_tmem(slot_prev, out, o8); }
            tc_fence_before();   // every read of the other slot is complete
            TRACE(5);
            cta_meet<6>(warp == kIssueWarp);
            if (tid == kIssueTid && has_next) { tc_fence_after(); issue_gemm1(slot_prev); }  // GEMM 1 of the next tile goes there
            __syncwarp();
            if (has_prev && cgroup == 0) epilogue2_store(prev_tile * kTileM + erow, o8, z);
        }
        TRACE(6);
        prev_tile = tile;
    }
    // ---- drain: epilogue 2 of this CTA's last tile
    {
        mbar_wait(bar2, phase2);
        tc_fence_after();
        float out[kOutRegs], o8[kOutRegs], z[kMaxAct];
        const uint32_t slot_last = tmem + ((it & 1) ^ 1) * kHid;
        if (cgroup == 0) draw_noise(prev_tile * kTileM + erow, z);
        epilogue2_math(slot_last, out, 0, false, 0u, false);
        if constexpr (kEarlyG1) {
            if (sums_in_a1 && cgroup != 0) quarter_meet(8, quarter, true);
            if (warp == kIssueWarp || warp == kIssueWarpB) issuer_publish(out);
            else exchange_smem(out, o8, false);
        } else exchange_tmem(slot_last, out, o8);
        if (cgroup == 0) epilogue2_store(prev_tile * kTileM + erow, o8, z);
    }
    TRACE2(2);
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

#ifdef SALP_ACTOR_TRACE
extern "C" int salp_actor_trace_read(long long *host_out) {
    return (int)cudaMemcpyFromSymbol(host_out, g_trace, sizeof(long long) * (2 * kTraceIters * kTracePoints + 8));
}
#endif

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
    // per device, once: the shared-memory opt-in of this instantiation and the SM count (the eager launch path is host-bound
    // for a kernel this short; racing first calls do the same idempotent work)
    constexpr int kMaxDev = 64;
    static bool attr_set[3][kMaxDev];
    static int sm_count[kMaxDev];
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) != cudaSuccess) return salp::set_error(SALP_ECUDA, "cudaGetDevice failed");
    const int ki = a.w3_rows <= 2 ? 0 : a.w3_rows <= 4 ? 1 : 2;
    const bool cached = dev >= 0 && dev < kMaxDev;
    if (!cached || !attr_set[ki][dev]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
        if (e != cudaSuccess) return salp::set_error(SALP_ECUDA, cudaGetErrorString(e));
        if (cached) attr_set[ki][dev] = true;
    }
    if (cached && sm_count[dev] > 0) sms = sm_count[dev];
    else if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && cached) sm_count[dev] = sms;
    const long long n_tiles = (a.n + kTileM - 1) / kTileM;
    const unsigned grid = (unsigned)(n_tiles < sms ? n_tiles : sms);  // persistent: one CTA per SM
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = kSmemBytes; cfg.stream = (cudaStream_t)cuda_stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = getenv("SALP_NO_PDL") ? 0 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, a);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return salp::set_error(SALP_ECUDA, cudaGetErrorString(e));
    return SALP_OK;
}
