// kernel_lab.cu -- A/B timing of step-kernel variants on a B200 (developer tool, not shipped).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -DLAB_F=1 -I underwater_swimmer_rl_b200/csrc \
//        tools/kernel_lab.cu -o tools/_lab/lab_f1
// Primes N envs to steady state once with a baseline variant, snapshots the state, then for every
// variant restores the snapshot, warms up and times STEPS launches with CUDA events.
// RNG tables hold splitmix64 noise (performance only; parity is tested elsewhere).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "salp_launch.cuh"

using namespace salp;

#ifndef LAB_F
#define LAB_F 1
#endif
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

static uint64_t sm64(uint64_t &s) { uint64_t z = (s += 0x9e3779b97f4a7c15ull); z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull; z = (z ^ (z >> 27)) * 0x94d049bb133111ebull; return z ^ (z >> 31); }

struct Variant { const char *name; cudaError_t (*fn)(const KernelArgs &, cudaStream_t); };

#define V(STORE, BLOCK, MINB) {#STORE "/b" #BLOCK "/minb" #MINB, &launch_step_variant<float, LAB_F, 3, STORE, BLOCK, MINB, 0>}
#define P(STORE, BLOCK, MINB) {#STORE "/b" #BLOCK "/minb" #MINB "/PF", &launch_step_variant<float, LAB_F, 3, STORE, BLOCK, MINB, 1>}

int main(int argc, char **argv) {
    const long long N = argc > 1 ? atoll(argv[1]) : (1 << 20);
    const int STEPS = argc > 2 ? atoi(argv[2]) : 100;
    const int PRIME = argc > 3 ? atoi(argv[3]) : 2048;
    const int F = LAB_F, OBS = 24, NACT = 8;
    const unsigned FL = 128, JL = 512;
    KernelArgs a;
    memset(&a, 0, sizeof(a));
    size_t body_b = 2 * N * 4 * sizeof(float), ctrl_b = 2 * N * 16, food_b = (size_t)F * N * 8;
    CK(cudaMalloc(&a.body, body_b)); CK(cudaMalloc((void **)&a.ctrl, ctrl_b)); CK(cudaMalloc(&a.ep_return, N * 4));
    CK(cudaMalloc(&a.food, food_b)); CK(cudaMalloc((void **)&a.obs, N * OBS * 4)); CK(cudaMalloc(&a.reward, N * 4));
    CK(cudaMalloc((void **)&a.terminated, N)); CK(cudaMalloc((void **)&a.truncated, N)); CK(cudaMalloc((void **)&a.events, N));
    CK(cudaMalloc((void **)&a.final_obs, N * OBS * 4)); CK(cudaMalloc(&a.final_info, N * 16)); CK(cudaMalloc((void **)&a.stats, 64));
    CK(cudaMemset(a.stats, 0, 64)); CK(cudaMemset(a.ctrl, 0, ctrl_b));
    std::vector<uint32_t> hw((size_t)N * JL);
    uint64_t seed = 42;
    for (auto &w : hw) w = (uint32_t)sm64(seed);
    uint32_t *fw, *jw;
    CK(cudaMalloc((void **)&fw, (size_t)N * FL * 4)); CK(cudaMalloc((void **)&jw, (size_t)N * JL * 4));
    CK(cudaMemcpy(jw, hw.data(), (size_t)N * JL * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(fw, hw.data(), (size_t)N * FL * 4, cudaMemcpyHostToDevice));
    a.food_words = fw; a.jit_words = jw;
    a.N = N; a.first = 0; a.count = N; a.rows = (unsigned)N; a.food_len = FL; a.jit_len = JL;
    a.food_len_mask = FL - 1; a.jit_len_mask = JL - 1;
    a.K = 1; a.obs_dim = OBS; a.observed = 3; a.act_dim = 1; a.forced = 1; a.respawn = 1; a.autoreset = argc > 5 ? atoi(argv[5]) : 1; a.stats = (argc > 6 && atoi(argv[6]) == 0) ? nullptr : a.stats; if (argc > 7 && atoi(argv[7]) == 0) { a.events = nullptr; a.final_obs = nullptr; a.final_info = nullptr; }
    a.W = 800; a.H = 600; a.diag = 1000.0; a.inv_w = 1.0 / 800; a.inv_h = 1.0 / 600; a.inv_diag = 1e-3;
    if (F == 1) { a.max_steps = 1500; a.num_food = 1; a.food_reward = 1000; a.collision_penalty = -10; a.time_penalty = -0.1; a.proximity_w = 5.0; }
    else { a.max_steps = 1500; a.num_food = 12; a.food_reward = 15; a.collision_penalty = -30; a.time_penalty = -0.05; a.efficiency_bonus = 2.0; }
    for (int t = 0; t < 128; ++t) { double wv = (t < 120 ? t : 120) / 120.0; a.exhale_lut[t] = (uint8_t)(int)(150.0 * (wv > 0.3 ? wv : 0.3)); }
    float *acts[NACT];
    std::vector<float> ha(N);
    for (int i = 0; i < NACT; ++i) {
        for (auto &x : ha) x = (float)((sm64(seed) >> 40) * (2.0 / 16777216.0) - 1.0);
        CK(cudaMalloc((void **)&acts[i], N * 4));
        CK(cudaMemcpy(acts[i], ha.data(), N * 4, cudaMemcpyHostToDevice));
    }
    CK(launch_reset_f32(a, F, 0));
    Variant base = V(kStoreWarp, 128, 8);
    for (int i = 0; i < PRIME; ++i) { a.actions = acts[i % NACT]; CK(base.fn(a, 0)); }
    CK(cudaDeviceSynchronize());
    void *snap_body, *snap_ctrl, *snap_ep, *snap_food;
    CK(cudaMalloc(&snap_body, body_b)); CK(cudaMalloc(&snap_ctrl, ctrl_b)); CK(cudaMalloc(&snap_ep, N * 4)); CK(cudaMalloc(&snap_food, food_b));
    CK(cudaMemcpy(snap_body, a.body, body_b, cudaMemcpyDeviceToDevice)); CK(cudaMemcpy(snap_ctrl, a.ctrl, ctrl_b, cudaMemcpyDeviceToDevice));
    CK(cudaMemcpy(snap_ep, a.ep_return, N * 4, cudaMemcpyDeviceToDevice)); CK(cudaMemcpy(snap_food, a.food, food_b, cudaMemcpyDeviceToDevice));

    std::vector<Variant> vs = {
#if LAB_F == 1
        V(kStoreWarp, 64, 16), V(kStoreWarp, 128, 8), V(kStoreTmaBlock, 64, 16), V(kStoreTmaBlock, 32, 32),
        V(kStoreTmaWarp, 64, 16), V(kStoreTmaWarp, 128, 8), V(kStoreTmaWarp, 256, 4),
        V(kStoreTmaBlock, 64, 17), V(kStoreTmaBlock, 64, 18), V(kStoreTmaBlock, 64, 20), V(kStoreTmaBlock, 32, 36), V(kStoreTmaBlock, 32, 40), V(kStoreTmaBlock, 128, 8), V(kStoreTmaBlock, 128, 9),
        P(kStoreWarp, 64, 16), P(kStoreWarp, 128, 8), P(kStoreWarp, 256, 4), P(kStoreWarp, 64, 12), P(kStoreWarp, 128, 6),
        P(kStoreTmaBlock, 64, 16), P(kStoreTmaBlock, 128, 8), P(kStoreTmaBlock, 256, 4), P(kStoreTmaBlock, 64, 12), P(kStoreTmaBlock, 32, 32),
        P(kStoreDirect, 64, 16), P(kStoreDirect, 128, 8), P(kStoreDirect, 64, 12),
#else
        V(kStoreDirect, 64, 14), V(kStoreDirect, 64, 12), V(kStoreDirect, 64, 10), V(kStoreWarp, 64, 14), V(kStoreWarp, 64, 12), V(kStoreTmaBlock, 64, 14), V(kStoreTmaBlock, 64, 12),
        V(kStoreDirect, 128, 7), V(kStoreDirect, 128, 6), V(kStoreWarp, 128, 7), V(kStoreDirect, 32, 28), V(kStoreTmaBlock, 32, 28),
        V(kStoreTmaBlock, 64, 16), V(kStoreTmaBlock, 32, 32), V(kStoreTmaBlock, 128, 7), V(kStoreTmaBlock, 64, 10), V(kStoreTmaBlock, 64, 8),
        P(kStoreTmaBlock, 64, 14), P(kStoreTmaBlock, 64, 12), P(kStoreTmaBlock, 128, 7), P(kStoreDirect, 64, 14), P(kStoreTmaBlock, 32, 28),
        V(kStoreTmaWarp, 64, 14), V(kStoreTmaWarp, 64, 12), V(kStoreTmaWarp, 128, 7), V(kStoreTmaWarp, 32, 28),
#endif
    };
    if (argc > 4) {  // keep only variants whose name contains argv[4]
        std::vector<Variant> keep;
        for (auto &v : vs) if (strstr(v.name, argv[4])) keep.push_back(v);
        vs = keep;
    }
    const double bpe = F == 1 ? 250.0 : 426.0;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    printf("N=%lld F=%d steps=%d prime=%d\n", N, F, STEPS, PRIME);
    for (int rep = 0; rep < 2; ++rep)
        for (auto &v : vs) {
            CK(cudaMemcpy(a.body, snap_body, body_b, cudaMemcpyDeviceToDevice)); CK(cudaMemcpy(a.ctrl, snap_ctrl, ctrl_b, cudaMemcpyDeviceToDevice));
            CK(cudaMemcpy(a.ep_return, snap_ep, N * 4, cudaMemcpyDeviceToDevice)); CK(cudaMemcpy(a.food, snap_food, food_b, cudaMemcpyDeviceToDevice));
            for (int i = 0; i < 5; ++i) { a.actions = acts[i % NACT]; CK(v.fn(a, 0)); }
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0));
            for (int i = 0; i < STEPS; ++i) { a.actions = acts[i % NACT]; CK(v.fn(a, 0)); }
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            const double us = 1e3 * ms / STEPS;
            // all variants run the same arithmetic from the same snapshot: state/obs hashes must agree
            static std::vector<uint32_t> hb;
            uint64_t h = 1469598103934665603ull;
            auto mix = [&](const void *dptr, size_t bytes) {
                hb.resize(bytes / 4);
                CK(cudaMemcpy(hb.data(), dptr, bytes, cudaMemcpyDeviceToHost));
                for (uint32_t w : hb) { h ^= w; h *= 1099511628211ull; }
            };
            mix(a.body, body_b); mix(a.ctrl, ctrl_b); mix(a.obs, (size_t)N * OBS * 4); mix(a.reward, (size_t)N * 4);
            printf("rep%d %-32s %8.2f us/step  %7.1f GB/s(alg)  %6.2f Gsteps/s  hash %016llx\n", rep, v.name, us, bpe * N / us / 1e3, N / us / 1e3, (unsigned long long)h);
            fflush(stdout);
        }
    return 0;
}
