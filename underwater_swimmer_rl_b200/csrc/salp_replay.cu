// salp_replay.cu -- rollout -> replay: append one env step of N transitions to a device ring buffer.
//
// What it replaces (reference tree, Python): the per-step `replay_buffer.add(obs, action, reward,
// next_obs, done or truncated)` of the episode loops
//   src/salp/training/continuous_trainer.py:296-303   (and trainer.py:203-216)
// for N environments at once, without leaving the GPU.  The reference never auto-resets, so its
// `next_obs` on a finishing step is the TERMINAL observation; under the vector env's auto-reset the
// step's `obs` row already belongs to the next episode and the terminal row is in `final_obs`, so the
// kernel takes next_obs from final_obs for finished lanes.  Rewards may be mixed on the fly with a
// second (imitation) reward stream: r = w_env * r_env + w_aux * r_aux (base_config.py:77-78).
//
// Layout of the ring (caller-owned, f32): obs [C][obs_dim], act [C][act_dim], rew [C], next [C][obs_dim],
// done u8 [C].  Row of env e in this call = (*head + e) mod C; `head` lives on the device so the call can
// be captured in a CUDA graph; the last CTA to finish advances it (ticket counter), or, without a ticket
// buffer, a one-thread kernel does afterwards.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/salp_b200.h"

namespace {

struct ReplayArgs {
    const float *obs_prev;   // [N][obs_dim]  observation the action was computed from
    const void *actions;     // [N][act_dim]  real
    const void *reward;      // [N]           real
    const float *reward_aux; // [N] f32 or null
    const uint8_t *terminated, *truncated;  // [N]
    const float *obs_next;   // [N][obs_dim]  obs returned by the step (next episode's first obs for finished lanes)
    const float *final_obs;  // [N][obs_dim]  terminal obs of finished lanes, or null (no auto-reset)
    float *rb_obs, *rb_act, *rb_rew, *rb_next;
    uint8_t *rb_done;
    long long *head;         // device scalar: next free row
    long long *size;         // device scalar: rows filled (saturates at capacity), may be null
    unsigned long long *ticket;  // device scalar (0 between calls): lets the last CTA advance head/size itself
    long long n, capacity;
    int obs_dim, act_dim, f64;
    float w_env, w_aux;
};

// One thread per 16-byte chunk of an observation row (obs_dim is a multiple of 4); the thread holding
// chunk 0 also moves the env's action, reward and done flag.  Reads and writes are fully coalesced:
// consecutive threads touch consecutive 16-byte chunks of both the source rows and the ring rows.
__global__ void __launch_bounds__(256) replay_append_kernel(const __grid_constant__ ReplayArgs a) {
    const int chunks = a.obs_dim >> 2;
    const long long total = a.n * chunks;
    const long long head = *a.head;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long e = i / chunks;
        const int c = (int)(i - e * chunks);
        long long row = head + e;
        row -= (row >= a.capacity) ? a.capacity : 0;  // n <= capacity and head < capacity
        const bool done = (a.terminated[e] | a.truncated[e]) != 0;
        const float4 o = reinterpret_cast<const float4 *>(a.obs_prev)[i];
        const float4 *nsrc = reinterpret_cast<const float4 *>((done && a.final_obs) ? a.final_obs : a.obs_next);
        const float4 nx = nsrc[i];
        reinterpret_cast<float4 *>(a.rb_obs)[row * chunks + c] = o;
        reinterpret_cast<float4 *>(a.rb_next)[row * chunks + c] = nx;
        if (c == 0) {
            float r;
            if (a.f64) {
                r = (float)reinterpret_cast<const double *>(a.reward)[e];
                for (int j = 0; j < a.act_dim; ++j)
                    a.rb_act[row * a.act_dim + j] = (float)reinterpret_cast<const double *>(a.actions)[e * a.act_dim + j];
            } else {
                r = reinterpret_cast<const float *>(a.reward)[e];
                for (int j = 0; j < a.act_dim; ++j)
                    a.rb_act[row * a.act_dim + j] = reinterpret_cast<const float *>(a.actions)[e * a.act_dim + j];
            }
            if (a.reward_aux) r = a.w_env * r + a.w_aux * a.reward_aux[e];
            a.rb_rew[row] = r;
            a.rb_done[row] = done ? 1 : 0;
        }
    }
    // every CTA read *head when it started; the last one to finish moves it (no second launch)
    if (a.ticket) {
        __syncthreads();
        if (threadIdx.x == 0 && atomicAdd(a.ticket, 1ull) == (unsigned long long)gridDim.x - 1) {
            *a.ticket = 0ull;
            const long long h = head + a.n;
            *a.head = h >= a.capacity ? h - a.capacity : h;
            if (a.size) { const long long sz = *a.size + a.n; *a.size = sz > a.capacity ? a.capacity : sz; }
        }
    }
}

__global__ void replay_advance_kernel(long long *head, long long *size, long long n, long long capacity) {
    long long h = *head + n;
    *head = h >= capacity ? h - capacity : h;
    if (size) { const long long s = *size + n; *size = s > capacity ? capacity : s; }
}

}  // namespace

namespace salp { int set_error(int code, const char *msg); }  // salp_capi.cu owns the thread-local error text
#define salp_set_error_ salp::set_error

extern "C" int salp_replay_append(const SalpReplayAppend *p, void *cuda_stream) {
    if (!p) return salp_set_error_(SALP_EINVAL, "null argument");
    if (!p->obs_prev || !p->actions || !p->reward || !p->terminated || !p->truncated || !p->obs_next || !p->rb_obs ||
        !p->rb_act || !p->rb_rew || !p->rb_next || !p->rb_done || !p->head)
        return salp_set_error_(SALP_EINVAL, "salp_replay_append: a required pointer is null");
    if (p->n <= 0 || p->capacity <= 0 || p->n > p->capacity)
        return salp_set_error_(SALP_EINVAL, "salp_replay_append: need 0 < n <= capacity");
    if (p->obs_dim <= 0 || (p->obs_dim & 3) || p->act_dim <= 0)
        return salp_set_error_(SALP_EINVAL, "salp_replay_append: obs_dim must be a positive multiple of 4, act_dim > 0");
    if (p->dtype != SALP_F32 && p->dtype != SALP_F64) return salp_set_error_(SALP_EINVAL, "salp_replay_append: bad dtype");
    const uintptr_t al[] = {(uintptr_t)p->obs_prev, (uintptr_t)p->obs_next, (uintptr_t)p->final_obs, (uintptr_t)p->rb_obs,
                            (uintptr_t)p->rb_next};
    for (uintptr_t q : al)
        if (q % 16 != 0) return salp_set_error_(SALP_EINVAL, "salp_replay_append: observation buffers must be 16-byte aligned");
    ReplayArgs a;
    a.obs_prev = p->obs_prev; a.actions = p->actions; a.reward = p->reward; a.reward_aux = p->reward_aux;
    a.terminated = p->terminated; a.truncated = p->truncated; a.obs_next = p->obs_next; a.final_obs = p->final_obs;
    a.rb_obs = p->rb_obs; a.rb_act = p->rb_act; a.rb_rew = p->rb_rew; a.rb_next = p->rb_next; a.rb_done = p->rb_done;
    a.head = reinterpret_cast<long long *>(p->head); a.size = reinterpret_cast<long long *>(p->size);
    a.ticket = reinterpret_cast<unsigned long long *>(p->ticket);
    a.n = p->n; a.capacity = p->capacity; a.obs_dim = p->obs_dim; a.act_dim = p->act_dim; a.f64 = p->dtype == SALP_F64;
    a.w_env = p->w_env; a.w_aux = p->w_aux;
    const long long total = a.n * (a.obs_dim >> 2);
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;  // grid-stride beyond 32 CTAs per SM
    cudaStream_t s = (cudaStream_t)cuda_stream;
    replay_append_kernel<<<(unsigned)blocks, 256, 0, s>>>(a);
    if (!a.ticket) replay_advance_kernel<<<1, 1, 0, s>>>(a.head, a.size, a.n, a.capacity);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return salp_set_error_(SALP_ECUDA, cudaGetErrorString(e));
    return SALP_OK;
}
