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
};

// food_slots must be one of 1, 2, 4, 8, 12, 16, 32
cudaError_t launch_step_f32(const KernelArgs &a, int food_slots, cudaStream_t s);
cudaError_t launch_step_f64(const KernelArgs &a, int food_slots, cudaStream_t s);
cudaError_t launch_reset_f32(const KernelArgs &a, int food_slots, cudaStream_t s);
cudaError_t launch_reset_f64(const KernelArgs &a, int food_slots, cudaStream_t s);
cudaError_t launch_refresh_rng(const KernelArgs &a, cudaStream_t s);  // defined in the f32 translation unit

}  // namespace salp
