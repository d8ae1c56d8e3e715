// salp_internal.h -- shared between the C-ABI host code (salp_capi.cu) and the per-dtype
// kernel translation units (salp_kernels.cu compiled twice: float, and double with -fmad=false).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace salp {

constexpr int kHostChunkAlign = 256;  // host-path chunks start on multiples of this many envs (covers every CTA size)
constexpr int kStatsLen = 8;
constexpr int kMaxFood = 32;
constexpr int kMaxObserved = 8;

// fsm word layout (see include/salp_b200.h)
constexpr int kPhaseRest = 0, kPhaseInhale = 1, kPhaseExhale = 2;
constexpr int kEpLenMask = (1 << 26) - 1;  // ctrl rec1.z = episode_len (26 bits, saturating) | n_food << 26
constexpr int kShapeReset = 0, kShapeRest = 1, kShapeInhale = 2, kShapeExhale = 3;

// One entry per position of the forced-breathing cycle: index 0 = fresh reset (R:95-117), 1..273 = state after the
// 1 rest + 121 inhale + 151 exhale steps of a cycle.  Built on the host in double with the reference's
// expressions (R:184-259) and rounded once; 16 bytes = one 128-bit shared-memory load.
struct CycleEntry {
    float radius;        // max(ellipse_a, ellipse_b) after this step's FSM update (R:190-191,207-213,240-246)
    float tm;            // jet thrust 100 * water_volume * 0.4 of this step (R:268), 0 outside the thrust window
    float wv;            // water_volume after this step (R:215,253,258)
    uint32_t fsm_touch;  // bits 0..27: fsm word after this step; bits 28..30: does a clamp onto the low / right / bottom
                         // wall register as a wall contact -- (50+r)-r <= 50, (W-(50+r))+r >= W-50, same with H -- evaluated
                         // in DOUBLE exactly as the reference does (R:335-352 + S:225-228)
};
constexpr int kCycleLen = 274;

// Passed by value as a __grid_constant__ kernel parameter (lives in the constant bank).
struct KernelArgs {
    void *body;        // real [2][N][4]
    int4 *ctrl;        // i32  [2][N][4]
    void *ep_return;   // real [N]
    void *food;        // real [F][N][2]
    const void *actions;
    float *obs;
    void *reward;
    uint8_t *terminated, *truncated, *events;
    float *final_obs;
    void *final_info;  // real [K][N][4]
    double *stats;
    const uint32_t *food_words, *jit_words;
    const uint8_t *reset_mask;  // reset kernel only (device pointer or null)
    long long N;                // array stride == num_envs of the handle
    long long first, count;     // env range [first, first+count) this launch covers
    unsigned rows, food_len_mask, jit_len_mask, food_len, jit_len;
    int K, obs_dim, observed, act_dim;
    int forced, random_food_count, respawn, autoreset, max_steps, num_food, rewind_rng;
    double W, H, food_reward, collision_penalty, time_penalty, efficiency_bonus, proximity_w;
    double diag, inv_w, inv_h, inv_diag;  // sqrt(W^2+H^2) and reciprocals, computed once on the host
    uint8_t exhale_lut[128];  // int(150 * max(t/120, 0.3)) for inhale timer t (R:223)
    // ---- f32 throughput build: the same constants rounded once on the host (a per-thread F2F.F32.F64 is an XU-pipe op)
    float fW, fH, f_inv_w, f_inv_h, f_inv_diag, f_diag;
    float f_food_reward, f_collision_penalty, f_time_penalty, f_efficiency_bonus, f_proximity_w;
    // ---- forced breathing: everything the breathing subsystem contributes to a step is a function of the position in
    //      the 273-step cycle; table of kCycleLen entries in device memory owned by the handle
    const CycleEntry *cycle_tab;
    // ---- RNG word source: 0 = host-seeded MT19937 tables (parity with the reference), 1 = Philox4x32-10 computed in the
    //      kernel, keyed by rng_seed, counter (word index / 4, stream, global env id); see include/salp_b200.h
    int rng_mode;
    unsigned long long rng_seed;
    long long env_id0;          // global id of env 0 of this handle
    int l2_mode;                // forced-breathing kernel: 1 state evict_last, 2 outputs / actions evict_first, 4 food evict_last
};

cudaError_t launch_step_fast_f32(const KernelArgs &a, int food_slots, cudaStream_t s, bool *handled);

// food_slots must be one of 1, 2, 4, 8, 12, 16, 32
cudaError_t launch_step_f32(const KernelArgs &a, int food_slots, cudaStream_t s);
cudaError_t launch_step_f64(const KernelArgs &a, int food_slots, cudaStream_t s);
cudaError_t launch_reset_f32(const KernelArgs &a, int food_slots, cudaStream_t s);
cudaError_t launch_reset_f64(const KernelArgs &a, int food_slots, cudaStream_t s);
cudaError_t launch_observe_f32(const KernelArgs &a, cudaStream_t s);
cudaError_t launch_observe_f64(const KernelArgs &a, cudaStream_t s);
cudaError_t launch_refresh_rng(const KernelArgs &a, cudaStream_t s);  // defined in the f32 translation unit

}  // namespace salp
