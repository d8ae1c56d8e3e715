// salp_step_fast.cuh -- the f32 throughput kernel for forced breathing (every BASELINE config): persistent warps,
// TMA bulk loads / stores through shared memory, breathing subsystem read from the 273-step cycle table.
//
// What it replaces (reference tree, Python): the same path as salp_step.cuh --
//   R = scripts/utilities/salp_robot.py             SalpRobotEnv.step            R:119-156
//   S = src/salp/environments/salp_snake_env.py     SalpSnakeEnv.step / reset    S:133-202
// for `forced_breathing=True` (S:31,66; R:158-167), max_observed_food = 3 (S:32) and 1 or 12 food slots, i.e. the four
// reference YAMLs.  Everything else (manual breathing, other food counts, the fp64 validation build) runs the generic
// kernel of salp_step.cuh; both are checked against the same oracle.
//
// Why a second kernel.  The generic one-thread-per-env kernel loads its state straight into registers: a warp sits on
// its first dependent instruction for a DRAM round trip, then runs ~760 dependent instructions, and the only thing that
// hides either is occupancy (ncu round 1: 45 % of stall samples on the state loads, 13 % on the CTA barrier before the TMA
// store, issue slots 45 % busy).  Here
//   * every WARP is its own pipeline over 32-env tiles: lane 0 issues cp.async.bulk (TMA) loads of the tile's SoA
//     records -- each record array of a tile is 512 contiguous bytes -- into a two-stage shared-memory ring, completion
//     on an mbarrier; the loads of tile i+1 are in flight while tile i is integrated, so HBM latency is off the
//     critical path at any occupancy;
//   * results leave the same way: the new state is written over the tile's input slots, the 32 observation rows
//     (3 KB, contiguous in HBM), rewards and flags into an output buffer, and lane 0 issues bulk stores -- no per-thread
//     global address arithmetic, no CTA-wide barrier (warps never synchronise with each other);
//   * forced breathing makes phase, timer, ellipse axes, water volume, thrust window (and whether a wall touch at that
//     radius registers as a contact in the reference's doubles) functions of the position in the 273-step
//     cycle: two 128-bit loads from a host-built table (CycleEntry, L1-resident) replace the FSM, the morph, the
//     divisions and the water-volume recurrence.
#pragma once
#include "salp_step.cuh"

namespace salp {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n"
        "SALP_MBAR_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra SALP_MBAR_WAIT_%=;\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 1-D TMA: global -> this CTA's shared memory, completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_load(void *sdst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(sdst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_store(void *gdst, const void *ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
}
// same, with an L2 eviction-priority hint (createpolicy): evict_last keeps the state records of the slab in the 126 MB L2
// between consecutive steps, evict_first lets the observation tiles (written once, never read by this kernel) leave first
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_normal() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ float ld_hint_f32(const float *p, uint64_t pol) {
    float v;
    asm volatile("ld.global.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(v) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ void st_hint_f32(float *p, float v, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.f32 [%0], %1, %2;" ::"l"(p), "f"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_hint_u8(uint8_t *p, uint32_t v, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.u8 [%0], %1, %2;" ::"l"(p), "r"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void bulk_load_hint(void *sdst, const void *gsrc, uint32_t bytes, uint64_t *bar, uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(sdst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
                 : "memory");
}
__device__ __forceinline__ void bulk_store_hint(void *gdst, const void *ssrc, uint32_t bytes, uint64_t pol) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_hint_v4(void *p, uint32_t x, uint32_t y, uint32_t z, uint32_t w, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "r"(x), "r"(y), "r"(z), "r"(w), "l"(pol) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// One lane of a converged warp.  With elect.sync (instead of `lane == 0`) and operands derived from warp-uniform values
// the compiler issues the bulk copies from the uniform datapath; with per-thread operands it wraps every copy in a
// ~12-instruction "waterfall" loop (measured: half of all warp instructions of the first version of this kernel).
__device__ __forceinline__ bool elect_one_sync() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred != 0;
}

// Per-warp shared-memory layout (bytes).  The "core" stage holds the 32-env tile's four 128-bit state records; it is
// single-buffered: a lane reads its records into registers first thing, after which the stage is refilled with the NEXT
// tile (a whole tile of arithmetic ahead of its use).  One food record per env travels with the core stage; twelve are a
// separate, double-buffered block (the kernel keeps using it as (dx, dy) scratch until the observation is assembled).
template <int F> struct FastLayout {
    static constexpr bool kFoodInCore = (F == 1);
    static constexpr int kC0 = 0, kC1 = 512, kB0 = 1024, kB1 = 1536, kFoodC = 2048;
    static constexpr int kCore = 2048 + (kFoodInCore ? 256 : 0);
    static constexpr int kFood = kFoodInCore ? 0 : F * 256;     // [F][32] float2, two stages
    static constexpr int kObs = 32 * 24 * 4;                    // 32 rows of 24 floats
    static constexpr int kOffCore = 0;
    static constexpr int kOffFood = kCore;
    static constexpr int kOffObs = kOffFood + 2 * kFood;
    static constexpr int kOffBar = kOffObs + kObs;              // core, food[2]: 8 bytes each
    static constexpr int kWarpBytes = (kOffBar + 24 + 127) & ~127;
};

// Launch shape per food-slot count: warps per CTA, CTAs per SM (registers: 65536 / (32 * warps * minb)).
template <int F> struct FastCfg;
#ifndef SALP_FAST_F1_MINB
#define SALP_FAST_F1_MINB 7
#endif
template <> struct FastCfg<1> { static constexpr int warps = 4, minb = SALP_FAST_F1_MINB; };    // 28 warps/SM, 72 registers
template <> struct FastCfg<12> { static constexpr int warps = 8, minb = 2; };                    // 16 warps/SM, 128 registers

#ifdef SALP_LAB_CLOCK  // developer builds: per-warp phase timestamps of the first tile (tools/_lab)
__device__ long long g_lab_clk[24 * 8192];
#define SALP_LAB_T(i) do { if (it == 0 && lane == 0 && gw < 8192u) g_lab_clk[gw * 24 + (i)] = clock64(); } while (0)
#define SALP_LAB_V(i, v) do { if (it == 0 && lane == 0 && gw < 8192u) g_lab_clk[gw * 24 + (i)] = (long long)(v); } while (0)
#else
#define SALP_LAB_T(i) do { } while (0)
#define SALP_LAB_V(i, v) do { } while (0)
#endif

template <int F, int RNG>
__global__ void __launch_bounds__(FastCfg<F>::warps * 32, FastCfg<F>::minb) salp_step_fast_kernel(const __grid_constant__ KernelArgs a) {
    using Cfg = FastCfg<F>;
    using L = FastLayout<F>;
    constexpr int OBS = 24;
    constexpr int MU = F < 3 ? F : 3;
    constexpr unsigned kSlotMask = F <= 1 ? 0u : (F <= 2 ? 1u : (F <= 4 ? 3u : (F <= 8 ? 7u : (F <= 16 ? 15u : 31u))));
    constexpr unsigned kFull = 0xffffffffu;
    extern __shared__ __align__(128) unsigned char s_raw[];
    __shared__ __align__(16) CycleEntry s_tab[kCycleLen];  // the forced-breathing cycle, 4.4 KB per CTA
    const int lane = threadIdx.x & 31;
    const int warp = __shfl_sync(kFull, (int)(threadIdx.x >> 5), 0);  // warp-uniform for the compiler
    unsigned char *ws = s_raw + (size_t)warp * L::kWarpBytes;
    uint64_t *bar_core = reinterpret_cast<uint64_t *>(ws + L::kOffBar);
    uint64_t *bar_food = bar_core + 1;  // [2]
    float *s_obs = reinterpret_cast<float *>(ws + L::kOffObs);

    const unsigned N = (unsigned)a.N;
    const unsigned first = (unsigned)a.first, end = first + (unsigned)a.count;
    const unsigned ntiles = ((unsigned)a.count + 31u) / 32u;
    const unsigned gw = blockIdx.x * Cfg::warps + warp, nw = gridDim.x * Cfg::warps;

    const float4 *g_body = reinterpret_cast<const float4 *>(a.body);
    const float2 *g_food = reinterpret_cast<const float2 *>(a.food);
    float *g_epret = reinterpret_cast<float *>(a.ep_return);
    const float *g_act = reinterpret_cast<const float *>(a.actions);
    float *g_reward = reinterpret_cast<float *>(a.reward);
    const float W = a.fW, H = a.fH, invW = a.f_inv_w, invH = a.f_inv_h, invDiag = a.f_inv_diag;
    // L2 residency (a.l2_mode, chosen per slab by the host): the per-env state that is read and rewritten every step stays
    // in the 126 MB L2 between consecutive steps (evict_last) when it fits, and what is written once and never read by
    // this kernel (observation tiles, rewards, flags) or read once (actions) leaves first (evict_first)
    const uint64_t pol_state = (a.l2_mode & 1) ? l2_policy_evict_last() : l2_policy_evict_normal();
    const uint64_t pol_out = (a.l2_mode & 2) ? l2_policy_evict_first() : l2_policy_evict_normal();
    const uint64_t pol_food = (a.l2_mode & 4) ? l2_policy_evict_last() : l2_policy_evict_normal();

#ifdef SALP_LAB_CLOCK
    { const unsigned it = 0; SALP_LAB_T(0); SALP_LAB_V(15, blockIdx.x); unsigned smid; asm("mov.u32 %0, %%smid;" : "=r"(smid)); SALP_LAB_V(14, smid); }
#endif
    if (elect_one_sync()) {
        mbar_init(bar_core, 1);
        mbar_init(bar_food + 0, 1);
        mbar_init(bar_food + 1, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // the cycle table is parameter data of the handle (never written by a kernel): it may be read before the previous
    // kernel of the stream has completed
    for (unsigned i = threadIdx.x; i < (unsigned)kCycleLen; i += Cfg::warps * 32)
        reinterpret_cast<uint4 *>(s_tab)[i] = __ldg(reinterpret_cast<const uint4 *>(a.cycle_tab) + i);
    __syncthreads();
    // programmatic dependent launch (salp_launch.cuh): the next kernel of the stream may become resident as CTAs of this
    // one retire; no state is touched before the previous kernel has completed
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    // ---- tile loads.  Full tiles: TMA, one elected lane, every operand warp-uniform.  The last, partial tile of a slab
    //      whose size is not a multiple of 32: each lane copies its own records (bulk copies need 16-byte multiples).
    auto rows_of = [&](unsigned tile) -> unsigned { const unsigned e0 = first + tile * 32u; return end - e0 < 32u ? end - e0 : 32u; };
    auto load_core = [&](unsigned tile) {
        unsigned char *sc = ws + L::kOffCore;
        const unsigned e0 = first + tile * 32u;
        if (rows_of(tile) == 32u) {
            if (elect_one_sync()) {
                mbar_expect_tx(bar_core, L::kCore);
                bulk_load_hint(sc + L::kC0, a.ctrl + e0, 512, bar_core, pol_state);
                bulk_load_hint(sc + L::kC1, a.ctrl + (size_t)N + e0, 512, bar_core, pol_state);
                bulk_load_hint(sc + L::kB0, g_body + e0, 512, bar_core, pol_state);
                bulk_load_hint(sc + L::kB1, g_body + (size_t)N + e0, 512, bar_core, pol_state);
                if constexpr (L::kFoodInCore) bulk_load_hint(sc + L::kFoodC, g_food + e0, 256, bar_core, pol_food);
            }
        } else {
            const unsigned e = e0 + lane;
            const bool v = e < end;
            reinterpret_cast<int4 *>(sc + L::kC0)[lane] = v ? a.ctrl[e] : make_int4(0, 0, 0, 0);
            reinterpret_cast<int4 *>(sc + L::kC1)[lane] = v ? a.ctrl[(size_t)N + e] : make_int4(0, 0, 0, 0);
            reinterpret_cast<float4 *>(sc + L::kB0)[lane] = v ? g_body[e] : make_float4(0.f, 0.f, 0.f, 0.f);
            reinterpret_cast<float4 *>(sc + L::kB1)[lane] = v ? g_body[(size_t)N + e] : make_float4(0.f, 0.f, 0.f, 0.f);
            if constexpr (L::kFoodInCore) reinterpret_cast<float2 *>(sc + L::kFoodC)[lane] = v ? g_food[e] : make_float2(0.f, 0.f);
            __syncwarp();
        }
    };
    auto load_food = [&](unsigned tile, int st) {  // F > 1: one of the two food blocks, its own mbarrier
        unsigned char *sf = ws + L::kOffFood + st * L::kFood;
        const unsigned e0 = first + tile * 32u;
        if (rows_of(tile) == 32u) {
            if (elect_one_sync()) {
                mbar_expect_tx(bar_food + st, L::kFood);
#pragma unroll
                for (int f = 0; f < F; ++f) bulk_load_hint(sf + f * 256, g_food + (size_t)f * N + e0, 256, bar_food + st, pol_food);
            }
        } else {
            const unsigned e = e0 + lane;
#pragma unroll
            for (int f = 0; f < F; ++f)
                reinterpret_cast<float2 *>(sf)[f * 32 + lane] = e < end ? g_food[(size_t)f * N + e] : make_float2(0.f, 0.f);
            __syncwarp();
        }
    };

#ifdef SALP_LAB_CLOCK
    { const unsigned it = 0; SALP_LAB_T(1); }
#endif
    unsigned ph_core = 0u, ph_food = 0u;  // mbarrier phase parities (bit st of ph_food)
    // the two scalars of the state that are not 128-bit records travel in registers, fetched one tile ahead
    float act_nx = 0.f, ep_nx = 0.f;
    if (gw < ntiles) {
        load_core(gw);
        if constexpr (!L::kFoodInCore) load_food(gw, 0);
        const unsigned e = first + gw * 32u + lane;
        if (e < end) { act_nx = ld_hint_f32(g_act + e, pol_out); ep_nx = ld_hint_f32(g_epret + e, pol_state); }
    }

    unsigned it = 0;
    for (unsigned tile = gw; tile < ntiles; tile += nw, ++it) {
        const int st = (int)(it & 1u);
        const unsigned e0 = first + tile * 32u;
        const unsigned e = e0 + lane;
        const unsigned rows = rows_of(tile);
        const bool full = rows == 32u;
        const bool valid = (unsigned)lane < rows;
        const unsigned nxt = tile + nw;
        unsigned char *sc = ws + L::kOffCore;
        // F == 1: the tile's food position, private copy (the core stage is refilled below); F > 1: this tile's block
        float2 *sfood = reinterpret_cast<float2 *>(ws + L::kOffFood + (L::kFoodInCore ? 0 : st * L::kFood));

        float act = act_nx, ep_ret = ep_nx;
        if (nxt < ntiles) {
            const unsigned en = first + nxt * 32u + lane;
            if (en < end) { act_nx = ld_hint_f32(g_act + en, pol_out); ep_nx = ld_hint_f32(g_epret + en, pol_state); }
        }
        SALP_LAB_T(2);
        if (full) { mbar_wait(bar_core, ph_core & 1u); ph_core ^= 1u; }
        SALP_LAB_T(3);
        // ---- state of this lane's env
        const int4 c0 = reinterpret_cast<const int4 *>(sc + L::kC0)[lane];
        const int4 c1 = reinterpret_cast<const int4 *>(sc + L::kC1)[lane];
        const float4 b0 = reinterpret_cast<const float4 *>(sc + L::kB0)[lane];
        const float4 b1 = reinterpret_cast<const float4 *>(sc + L::kB1)[lane];
        float2 food1 = make_float2(0.f, 0.f);
        if constexpr (L::kFoodInCore) food1 = reinterpret_cast<const float2 *>(sc + L::kFoodC)[lane];
        // ---- every lane holds its records: refill the stage with the next tile, a whole tile of arithmetic ahead of use
        __syncwarp();
        if (nxt < ntiles) {
            load_core(nxt);
            if constexpr (!L::kFoodInCore) load_food(nxt, st ^ 1);
        }

        float px = b0.x, py = b0.y, vx = b0.z, vy = b0.w;
        float angle = b1.x, omega = b1.y, nozzle = b1.z;
        int fsm = c0.x;
        int steps_since_food = c0.y, food_collected = c0.z;
        unsigned mask = (unsigned)c0.w;
        unsigned cur_food = (unsigned)c1.x, cur_jit = (unsigned)c1.y;
        int ep_len = c1.z & kEpLenMask, n_food = (int)((unsigned)c1.z >> 26);
        unsigned jcache = (unsigned)c1.w;
        float wv_now = b1.w;
        // position in the breathing cycle: 0 fresh reset, 1..121 inhaling (timer 0..120), 122..272 exhaling
        // (timer 0..150), 273 the one rest step
        int q;
        {
            const int phase = fsm & 3, timer = (fsm >> 2) & 0xff;
            q = phase == kPhaseInhale ? 1 + timer : (phase == kPhaseExhale ? 122 + timer : (((fsm >> 18) & 3) == kShapeReset ? 0 : 273));
        }
        const uint32_t *jit_row = nullptr;
        if constexpr (RNG == 0) jit_row = a.jit_words + (size_t)((a.rows == N) ? e : e % a.rows) * a.jit_len;

        for (int k = 0; k < a.K; ++k) {
            if (k > 0) act = valid ? g_act[(size_t)k * N + e] : 0.f;  // K fused substeps: later actions come straight from HBM
            float *row = s_obs + lane * OBS;
            q = q >= 273 ? 1 : q + 1;
            const uint4 t0 = *reinterpret_cast<const uint4 *>(s_tab + q);
            const float radius = __uint_as_float(t0.x), tm = __uint_as_float(t0.y);
            const bool thrust = tm != 0.f;
            // ---- thrust jitter draw (R:311): table mode -- the word for THIS draw came with the state (ctrl rec1.w), fetch
            //      the one for the next draw now, it is needed only at write-back; Philox mode -- computed, no memory
            uint32_t jword = jcache;
            if (thrust) {
                if constexpr (RNG == 0) jcache = ld_evict_last_u1(jit_row + ((cur_jit + 2u) & a.jit_len_mask));
#ifdef SALP_FAST_NOJIT
                else jword = 0x80000000u;
#else
                else jword = philox_jitter_pair(a, e, cur_jit).x;
#endif
            }
            // ---- nozzle slew (R:169-182)
            {
                const float target = act * (float)kc::MAX_NOZZLE;
                const float diff = target - nozzle;
                if (fabsf(diff) > (float)kc::NOZZLE_RATE) nozzle = (diff > 0.f) ? nozzle + (float)kc::NOZZLE_RATE : nozzle - (float)kc::NOZZLE_RATE;
                else nozzle = target;
                nozzle = pymax<float>((float)(-kc::MAX_NOZZLE), pymin<float>((float)kc::MAX_NOZZLE, nozzle));
            }
            // ---- jet thrust through the steerable nozzle (R:261-314); Tm = 100 * water_volume * 0.4 from the table
            if (thrust) {
                const float wvb = tm * 0.025f;  // water volume before this step's decay
                const float th = angle - nozzle;
                float s0, cs0;
                __sincosf(th, &s0, &cs0);
                vx += (cs0 * tm) * 0.012f;
                vy += (s0 * tm) * 0.012f;
                const float primary = -nozzle * tm * 0.0002f;
                const float perp = tm * __sinf(-nozzle);
                const float moment = perp * (radius * 0.7f) * 0.00005f;
                const float shape = -nozzle * tm * wvb * 0.00003f;
                omega += primary + moment + shape;
                const float sm = tm * fabsf(nozzle) * 0.3f;
                vx += (-s0 * sm) * 0.008f;  // cos(x + pi/2) = -sin x
                vy += (cs0 * sm) * 0.008f;  // sin(x + pi/2) =  cos x
                cur_jit += 2;
                const float u = (float)(jword >> 5) * (1.0f / 134217728.0f);
                const float dn = (u - 0.5f) * 0.05f;
                const float nf = tm * 0.04f;
                const float d2 = dn * dn;  // sin/cos(th + dn) by angle addition, |dn| <= 0.025
                const float sd = dn * (1.0f - d2 * ((1.0f / 6.0f) - d2 * (1.0f / 120.0f)));
                const float cd = 1.0f - d2 * (0.5f - d2 * (1.0f / 24.0f));
                vx += (cs0 * cd - s0 * sd) * nf * 0.002f;
                vy += (s0 * cd + cs0 * sd) * nf * 0.002f;
            }
            wv_now = __uint_as_float(t0.z);
            fsm = (int)(t0.w & 0x0fffffffu);
            // ---- drag, integrate, wrap, wall bounce (R:316-352)
            vx *= 0.98f; vy *= 0.98f; omega *= 0.95f;
            px += vx; py += vy;
            angle = wrap_pi<float>(angle + omega);
            unsigned clamped = 0u;
            {
                const float margin = (float)kc::TANK_MARGIN + radius;
                if (px < margin) { px = margin; vx = fabsf(vx) * 0.4f; omega *= 0.7f; clamped |= 1u; }
                else if (px > W - margin) { px = W - margin; vx = -fabsf(vx) * 0.4f; omega *= 0.7f; clamped |= 2u; }
                if (py < margin) { py = margin; vy = fabsf(vy) * 0.4f; omega *= 0.7f; clamped |= 1u; }
                else if (py > H - margin) { py = H - margin; vy = -fabsf(vy) * 0.4f; omega *= 0.7f; clamped |= 4u; }
            }
            // ---- wall contact (S:219-230): a bounce leaves the robot exactly on the clamp line; whether that registers is
            //      the rounding of (50 + r) - r <= 50 in the reference's doubles -- a table bit per cycle position
            const bool collision = (clamped & (t0.w >> 28)) != 0u;

            // ---- food: offsets, distances, first live food within reach is eaten (S:204-217), nearest MU selected
            bool hit = false;
            float dx0 = 0.f, dy0 = 0.f, dd0 = 0.f;  // F == 1
            unsigned key[MU > 0 ? MU : 1];
            float dist_sum = 0.f;
            const float key_scale = invDiag * (4294967296.0f * 0.9999f);
            auto make_key = [&](float d, int f) -> unsigned { return (__float2uint_rz(d * key_scale) & ~kSlotMask) | (unsigned)f; };
            auto insert_key = [&](unsigned c) {
#pragma unroll
                for (int i = 0; i < MU; ++i) { const unsigned lo = min(key[i], c); c = max(key[i], c); key[i] = lo; }
            };
            auto reselect = [&]() {  // rare: a food was eaten / the env was reset -- redo the selection from the scratch
                dist_sum = 0.f;
#pragma unroll
                for (int i = 0; i < MU; ++i) key[i] = 0xffffffffu;
#pragma unroll 1
                for (int f = 0; f < F; ++f) {
                    const float2 v = sfood[f * 32 + lane];
                    const float d = sqrt_(v.x * v.x + v.y * v.y);
                    const bool alive = (mask >> f) & 1u;
                    if (alive) dist_sum += d;
                    insert_key(alive ? make_key(d, f) : (~kSlotMask | (unsigned)f));
                }
            };
            const float reach = radius + (float)kc::FOOD_R;
            if constexpr (F == 1) {
                dx0 = food1.x - px; dy0 = food1.y - py;
                dd0 = sqrt_(dx0 * dx0 + dy0 * dy0);
                if ((mask & 1u) && dd0 < reach) { hit = true; mask &= ~1u; }
            } else {
                SALP_LAB_T(4);
                if (k == 0 && full) { mbar_wait(bar_food + st, (ph_food >> st) & 1u); ph_food ^= 1u << st; }
                SALP_LAB_T(5);
                unsigned hitbits = 0u;
#pragma unroll
                for (int i = 0; i < MU; ++i) key[i] = 0xffffffffu;
#pragma unroll
                for (int f = 0; f < F; ++f) {
                    const float2 fv = sfood[f * 32 + lane];
                    const float ox = fv.x - px, oy = fv.y - py;  // the buffer holds positions on entry, offsets afterwards
                    const float d = sqrt_(ox * ox + oy * oy);
                    sfood[f * 32 + lane] = make_float2(ox, oy);
                    const bool alive = (mask >> f) & 1u;
                    if (d < reach) hitbits |= 1u << f;
                    if (alive) dist_sum += d;
                    insert_key(alive ? make_key(d, f) : (~kSlotMask | (unsigned)f));
                }
                hitbits &= mask;
                if (hitbits != 0u) { hit = true; mask &= ~(hitbits & (0u - hitbits)); reselect(); }
            }
            SALP_LAB_T(6);
            // ---- reward (S:278-327)
            float reward = 0.f;
            if (hit) {
                reward += a.f_food_reward;
                if (a.f_efficiency_bonus > 0.f) reward += a.f_efficiency_bonus * (float)(a.max_steps - steps_since_food);
            }
            if (collision) reward += a.f_collision_penalty;
            bool any_food = mask != 0u;
            float near_x, near_y, near_rel = 0.f;
            if constexpr (F == 1) { near_x = dx0; near_y = dy0; }
            else { const float2 nv = sfood[(key[0] & kSlotMask) * 32 + lane]; near_x = nv.x; near_y = nv.y; }
            if (any_food) near_rel = wrap_pi<float>(atan2_(near_y, near_x) - angle);
            if (a.f_proximity_w > 0.f && any_food) reward += a.f_proximity_w * __cosf(near_rel);
            reward += a.f_time_penalty;
            // ---- counters, respawn (S:171-180)
            ++steps_since_food;
            if (hit) { ++food_collected; steps_since_food = 0; }
            if (a.respawn) {  // the whole warp finds the new position of each eaten food in turn (respawn_warp)
                for (unsigned m = __ballot_sync(kFull, hit); m != 0u; m &= m - 1u) {
                    const int src = __ffs((int)m) - 1;
                    const RespawnResult<float> r = respawn_warp<float, F>(a, __shfl_sync(kFull, e, src), __shfl_sync(kFull, px, src),
                        __shfl_sync(kFull, py, src), __shfl_sync(kFull, mask, src), __shfl_sync(kFull, n_food, src),
                        __shfl_sync(kFull, cur_food, src), lane);
                    if (lane == src) {
                        mask = r.mask; cur_food = r.cur;
                        if (r.slot >= 0) {
                            const float ox = r.x - px, oy = r.y - py;
                            const float d = sqrt_(ox * ox + oy * oy);
                            if constexpr (F == 1) { dx0 = ox; dy0 = oy; dd0 = d; food1 = make_float2(r.x, r.y); }
                            else { sfood[r.slot * 32 + lane] = make_float2(ox, oy); dist_sum += d; insert_key(make_key(d, r.slot)); }
                        }
                    }
                }
            }
            // ---- termination (S:182-189): collision > timeout > all collected
            bool terminated = false, truncated = false;
            if (collision) terminated = true;
            else if (steps_since_food > a.max_steps) truncated = true;
            else if (!a.respawn && mask == 0u) terminated = true;
            if (!valid) { terminated = false; truncated = false; }
            ep_ret += reward;
            ++ep_len;
            // ---- per-step scalars: one coalesced store each (a warp's 32 rewards are 128 contiguous bytes, its flags 32)
            const bool done = terminated || truncated;
            if (valid) {
                const size_t oi = (size_t)k * N + e;
                st_hint_f32(g_reward + oi, reward, pol_out);
                st_hint_u8(a.terminated + oi, terminated ? 1u : 0u, pol_out);
                st_hint_u8(a.truncated + oi, truncated ? 1u : 0u, pol_out);
                if (a.events) st_hint_u8(a.events + oi, (terminated ? 1u : 0u) | (truncated ? 2u : 0u) | (collision ? 4u : 0u) | (hit ? 8u : 0u), pol_out);
            }

            float obs6 = radius * (float)(1.0 / kc::BASE_R), obs7 = (float)(t0.w & 3u) * 0.5f;
            // the observation tile buffer is single: the bulk store of the previous tile / substep must have read it
            // (it was issued a tile of arithmetic ago)
            SALP_LAB_T(7);
            if (elect_one_sync()) bulk_wait_read<0>();
            __syncwarp();
            SALP_LAB_V(12, 0);
#pragma unroll 1
            for (int pass = 0;; ++pass) {
                // ---- observation (R:377-388 + S:366-428) assembled in registers, written as 128-bit pieces of the row
                const int n_live = __popc(mask);
                auto food_entry = [&](int i) -> float4 {
                    float4 qv = make_float4(0.f, 0.f, 1.f, 0.f);  // S:412 padding
                    if (i < MU && i < n_live) {
                        float sx, sy, sd;
                        if constexpr (F == 1) { sx = dx0; sy = dy0; sd = dd0; }
                        else {
                            const float2 v = sfood[(key[i < MU ? i : 0] & kSlotMask) * 32 + lane];
                            sx = v.x; sy = v.y; sd = sqrt_(v.x * v.x + v.y * v.y);
                        }
                        float rel;
                        if (i == 0) rel = hit ? wrap_pi<float>(atan2_(sy, sx) - angle) : near_rel;
                        else rel = wrap_pi<float>(atan2_(sy, sx) - angle);
                        qv = make_float4(sx * invW, sy * invH, sd * invDiag, rel * (float)(1.0 / kc::PI));
                    }
                    return qv;
                };
                const float cnt = fminf((float)n_live * 0.1f, 1.0f);
                float avg = 1.f;
                if constexpr (F == 1) { if (n_live > 0) avg = dd0 * invDiag; }
                else if (n_live > 0) avg = __fdividef(dist_sum, (float)n_live) * invDiag;
                const float4 e0v = food_entry(0), e1v = food_entry(1), e2v = food_entry(2);
                float4 *r4 = reinterpret_cast<float4 *>(row);
                r4[0] = make_float4(px * invW, py * invH, vx * 0.2f, vy * 0.2f);
                r4[1] = make_float4(angle * (float)(1.0 / kc::PI), omega * 10.0f, obs6, obs7);
                r4[2] = make_float4(wv_now, nozzle * (float)(1.0 / kc::MAX_NOZZLE), e0v.x, e0v.y);
                r4[3] = make_float4(e0v.z, e0v.w, e1v.x, e1v.y);
                r4[4] = make_float4(e1v.z, e1v.w, e2v.x, e2v.y);
                r4[5] = make_float4(e2v.z, e2v.w, cnt, avg);
                if (pass) break;
                // ---- episode statistics: warp-shuffle reduction, only in warps where an episode ended
                const unsigned done_mask = __ballot_sync(kFull, done);
                SALP_LAB_T(8);
                if (done_mask == 0u) break;
                SALP_LAB_V(12, __popc(done_mask));
                if (a.stats) {
                    // ---- episode statistics: the (rare) finished lanes add their own episode -- reductions without a
                    //      return value, nothing on the warp's dependent chain (a shuffle tree of doubles was ~1 300 cycles)
                    if (done) {
                        atomicAdd(a.stats + 0, 1.0);
                        atomicAdd(a.stats + 1, (double)ep_ret);
                        atomicAdd(a.stats + 2, (double)ep_len);
                        atomicAdd(a.stats + 3, (double)food_collected);
                        if (collision) atomicAdd(a.stats + 4, 1.0);
                        if (truncated) atomicAdd(a.stats + 5, 1.0);
                        if (!(isfinite(px) && isfinite(py))) atomicAdd(a.stats + 6, 1.0);
                    }
                }
                SALP_LAB_T(16);
                if (done && a.final_info)
                    reinterpret_cast<float4 *>(a.final_info)[(size_t)k * N + e] =
                        make_float4(ep_ret, (float)ep_len, (float)food_collected, (float)steps_since_food);
                if (!a.autoreset) break;
                if (done && a.final_obs) {  // terminal observation (gymnasium "final_observation")
                    float4 *dst = reinterpret_cast<float4 *>(a.final_obs + ((size_t)k * N + e) * OBS);
#pragma unroll
                    for (int j = 0; j < OBS / 4; ++j) dst[j] = r4[j];
                }
                SALP_LAB_T(17);
                if constexpr (F < 8) {
                    // few foods: the in-thread reset writes the reset observation itself.  (Measured on B200 against the
                    // warp-cooperative layout below, which needs the whole warp's state live across the call: 72-register
                    // budget -> spills in the hot path, 41.8 -> 46.2 us at 1 M envs; with 80 registers no gain at any size.)
                    if (done) {
                        const ResetResult rr = F == 1 ? reset_single_f32(a, e, cur_food, row) : reset_slow<float>(a, e, cur_food, row);
                        cur_food = rr.cur; mask = rr.mask; n_food = rr.n_food;
                        px = W * 0.5f; py = H * 0.5f; vx = 0.f; vy = 0.f; angle = 0.f; omega = 0.f; nozzle = 0.f; wv_now = 0.f;
                        fsm = kPhaseRest | (kShapeReset << 18); q = 0;
                        steps_since_food = 0; food_collected = 0; ep_ret = 0.f; ep_len = 0;
                        if (a.K > 1) food1 = g_food[e];  // the next substep needs the new food position
                    }
                    SALP_LAB_T(19);
                    break;
                } else {
                    // ---- auto-reset (S:133-155): the whole warp lays out the food of each finished env in turn
                    for (unsigned m = done_mask; m != 0u; m &= m - 1u) {
                        const int src = __ffs((int)m) - 1;
                        const unsigned e_src = __shfl_sync(kFull, e, src);
                        const LayoutResult<float> Lr = layout_warp<float>(a, e_src, __shfl_sync(kFull, cur_food, src), lane);
                        if (lane < Lr.n_food) {
                            reinterpret_cast<float2 *>(a.food)[(size_t)lane * N + e_src] = make_float2(Lr.x, Lr.y);
                            // food `lane` of env `src`, as the offset from the reset position, straight into that env's
                            // column of the scratch (slots >= n_food keep stale offsets; their mask bits are clear)
                            sfood[lane * 32 + src] = make_float2(Lr.x - W * 0.5f, Lr.y - H * 0.5f);
                        }
                        // distances and selection keys of the new foods, one per lane: the MU nearest by MU warp-wide minima,
                        // the distance sum in slot order (the order `reselect` adds them in)
                        const float ox = Lr.x - W * 0.5f, oy = Lr.y - H * 0.5f;
                        const float dj = sqrt_(ox * ox + oy * oy);
                        unsigned kj = lane < Lr.n_food ? make_key(dj, lane) : (lane < F ? (~kSlotMask | (unsigned)lane) : 0xffffffffu);
                        unsigned kmin[MU];
#pragma unroll
                        for (int i = 0; i < MU; ++i) { kmin[i] = __reduce_min_sync(kFull, kj); if (kj == kmin[i]) kj = 0xffffffffu; }
                        float ds = 0.f;
#pragma unroll
                        for (int f = 0; f < F; ++f) { const float df = __shfl_sync(kFull, dj, f); if (f < Lr.n_food) ds += df; }
                        if (lane == src) {
                            cur_food = Lr.cur; n_food = Lr.n_food;
                            mask = Lr.n_food >= 32 ? 0xffffffffu : ((1u << Lr.n_food) - 1u);
#pragma unroll
                            for (int i = 0; i < MU; ++i) key[i] = kmin[i];
                            dist_sum = ds;
                        }
                    }
                    __syncwarp();
                    SALP_LAB_T(18);
                    if (done) {
                        px = W * 0.5f; py = H * 0.5f; vx = 0.f; vy = 0.f; angle = 0.f; omega = 0.f; nozzle = 0.f; wv_now = 0.f;
                        fsm = kPhaseRest | (kShapeReset << 18); q = 0;
                        steps_since_food = 0; food_collected = 0; ep_ret = 0.f; ep_len = 0;
                        obs6 = 1.0f; obs7 = 0.f;
                        hit = true;  // the bearing computed for the reward belongs to the finished episode
                    }
                    SALP_LAB_T(19);
                }
            }  // pass
            SALP_LAB_T(9);
            if constexpr (F > 1) {
                // K > 1: the next substep expects positions in the buffer again; re-read them from HBM (L1/L2 hits;
                // foods change only through the slow paths, which write HBM)
                if (k + 1 < a.K) {
#pragma unroll
                    for (int f = 0; f < F; ++f) sfood[f * 32 + lane] = valid ? g_food[(size_t)f * N + e] : make_float2(0.f, 0.f);
                }
            }
            // ---- publish this substep's observation tile (3 KB, contiguous in HBM)
            if (full) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (elect_one_sync()) {
                    bulk_store_hint(a.obs + ((size_t)k * N + e0) * OBS, s_obs, L::kObs, pol_out);
                    bulk_commit();
                }
            } else if (valid) {
                float4 *dst = reinterpret_cast<float4 *>(a.obs + ((size_t)k * N + e) * OBS);
                const float4 *src = reinterpret_cast<const float4 *>(row);
#pragma unroll
                for (int j = 0; j < OBS / 4; ++j) dst[j] = src[j];
            }
        }  // substeps

        // ---- write the state back: four coalesced 128-bit stores (a warp writes 512 contiguous bytes each) + ep_return;
        //      food records are only written by the slow paths
        if (valid) {
            st_hint_v4(a.ctrl + e, (uint32_t)fsm, (uint32_t)steps_since_food, (uint32_t)food_collected, mask, pol_state);
            st_hint_v4(a.ctrl + (size_t)N + e, cur_food, cur_jit, (uint32_t)((ep_len < kEpLenMask ? ep_len : kEpLenMask) | (n_food << 26)), jcache, pol_state);
            st_hint_v4(reinterpret_cast<float4 *>(a.body) + e, __float_as_uint(px), __float_as_uint(py), __float_as_uint(vx), __float_as_uint(vy), pol_state);
            st_hint_v4(reinterpret_cast<float4 *>(a.body) + (size_t)N + e, __float_as_uint(angle), __float_as_uint(omega), __float_as_uint(nozzle), __float_as_uint(wv_now), pol_state);
            st_hint_f32(g_epret + e, ep_ret, pol_state);
        }
        SALP_LAB_T(10);
    }  // tiles
#ifdef SALP_LAB_CLOCK
    { const unsigned it = 0; SALP_LAB_T(11); }
#endif
    if (blockIdx.x == 0 && threadIdx.x == 0 && a.stats) atomicAdd(a.stats + 7, (double)a.count * (double)a.K);
    if (elect_one_sync()) bulk_wait_read<0>();  // shared memory must stay allocated until the last bulk store has read it
}

}  // namespace salp
