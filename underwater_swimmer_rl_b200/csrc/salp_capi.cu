// salp_capi.cu -- the extern "C" boundary declared in include/salp_b200.h.
// Host-side only: parameter validation, buffer binding, kernel dispatch, the host-buffer
// (pinned memory) step path, and the MT19937 word-table generator that reproduces the
// reference's two global random streams.  No CPU implementation of the env step exists here:
// without a CUDA device every compute entry point returns SALP_ECUDA.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/salp_b200.h"
#include "salp_internal.h"
#include "salp_shape_host.h"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int cuda_fail(cudaError_t e, const char *what) {
    return fail(SALP_ECUDA, "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
}

#define SALP_CUDA(expr)                                   \
    do {                                                  \
        cudaError_t err__ = (expr);                       \
        if (err__ != cudaSuccess) return cuda_fail(err__, #expr); \
    } while (0)

constexpr int kHostChunks = 16;  // streams available to salp_step_host; it uses host_chunks() of them

// chunks of one salp_step_host call: enough that the first chunk's copy-in + kernel (the only part of the step that does not
// overlap a device->host copy) is short, few enough that the per-copy overhead stays small.  SALP_HOST_CHUNKS overrides.
int host_chunks() {
    static int n = 0;
    if (n == 0) {
        int v = 8;
        if (const char *ev = getenv("SALP_HOST_CHUNKS")) v = atoi(ev);
        n = v < 1 ? 1 : (v > kHostChunks ? kHostChunks : v);
    }
    return n;
}

int pick_food_slots(int n) {
    const int opts[] = {1, 4, 8, 12, 16, 32};
    for (int o : opts)
        if (n <= o) return o;
    return -1;
}

bool is_pow2(int64_t v) { return v > 0 && (v & (v - 1)) == 0; }

}  // namespace

namespace salp {
int set_error(int code, const char *msg) { return fail(code, "%s", msg); }  // for the other translation units
}  // namespace salp

namespace salp {

// Forced breathing (R:158-167 `(breathing_timer % 330) < 120`): the FSM, the body shape, the water volume and the thrust
// window are functions of the position in the 273-step cycle.  Run the reference's update (R:184-259) once on
// the host, in double, for two cycles from the reset state and keep the second one (its fsm words carry the steady-state
// exhale duration of 150, which is also what R:233's getattr default gives before the first exhale).
// out[0] = fresh reset; out[i], i = 1..273 = after the i-th step of a cycle.  wv64 / a64 / b64 (optional) receive the
// doubles the entries were rounded from (salp_forced_cycle: tests check them bit-for-bit against the fp64 kernel).
void build_forced_cycle(double W, double H, const uint8_t *exhale_lut, CycleEntry *out, double *wv64, double *a64, double *b64) {
    int phase = kPhaseRest, timer = 0, exd = 0, skind = kShapeReset, stimer = 0;
    double wv = 0.0;
    auto emit = [&](int i, double tm) {
        double ea, eb;
        shape_axes_f64(skind, stimer, exd, ea, eb);
        const double r = eb > ea ? eb : ea;
        CycleEntry &c = out[i];
        c.radius = (float)r;
        c.tm = (float)tm;
        c.wv = (float)wv;
        const uint32_t fsm = (uint32_t)(phase | (timer << 2) | (exd << 10) | (skind << 18) | (stimer << 20));
        c.fsm_touch = fsm | (wall_touch_bits_r(r, W, H) << 28);
        if (wv64) wv64[i] = wv;
        if (a64) a64[i] = ea;
        if (b64) b64[i] = eb;
    };
    emit(0, 0.0);
    for (int pass = 0; pass < 2; ++pass)
        for (int i = 1; i < kCycleLen; ++i) {
            const bool inhaling = timer < kc::INHALE_DUR;  // timer <= 151 < 330: the modulo is the identity
            const int t1 = timer + 1;
            const bool thrust = (phase == kPhaseExhale) && (t1 <= exd) && (10 * t1 >= exd) && (2 * t1 <= exd);
            int wv_mode = 0;
            if (phase == kPhaseRest) {
                skind = kShapeRest;
                if (inhaling) { phase = kPhaseInhale; timer = 0; }
            } else if (phase == kPhaseInhale) {
                if (inhaling && timer < kc::INHALE_DUR) { ++timer; skind = kShapeInhale; stimer = timer; wv_mode = 1; }
                else if (wv > 0.05) { exd = exhale_lut[timer]; phase = kPhaseExhale; timer = 0; }
                else { phase = kPhaseRest; timer = 0; wv_mode = 3; }
            } else {
                ++timer;
                if (timer <= exd) { skind = kShapeExhale; stimer = timer; wv_mode = 2; }
                else { phase = kPhaseRest; timer = 0; wv_mode = 3; }
            }
            const double prog = (skind == kShapeInhale) ? (double)stimer / (double)kc::INHALE_DUR
                                                        : (skind == kShapeExhale ? (double)stimer / (double)exd : 0.0);
            if (wv_mode == 1) wv = prog;
            const double tm = thrust ? 100.0 * wv * 0.4 : 0.0;  // R:268, uses water_volume before this step's decay
            if (wv_mode == 2) { const double w2 = wv * (1.0 - prog); wv = w2 > 0.0 ? w2 : 0.0; }
            else if (wv_mode == 3) wv = 0.0;
            if (pass == 1) emit(i, tm);
        }
}

}  // namespace salp

struct SalpHandle {
    SalpParams params;
    SalpLayout layout;
    SalpBuffers buf;
    salp::KernelArgs base;  // everything except per-launch fields
    int64_t num_envs, env_id0;
    int device, dtype;
    bool bound;
    int64_t launches;
    cudaStream_t host_streams[kHostChunks];
    cudaEvent_t host_event;
    cudaEvent_t chunk_done[kHostChunks];  // step_host: "the step kernel of chunk c has run"
    bool host_streams_ready;
    salp::CycleEntry *cycle_tab;  // device copy of the forced-breathing cycle table (handle-owned parameter data)
};

extern "C" {

int salp_abi_version(void) { return SALP_ABI_VERSION; }

const char *salp_last_error(void) { return g_err; }

int salp_query_layout(const SalpParams *p, int dtype, SalpLayout *out) {
    if (!p || !out) return fail(SALP_EINVAL, "null argument");
    if (dtype != SALP_F32 && dtype != SALP_F64) return fail(SALP_EINVAL, "dtype must be SALP_F32 or SALP_F64");
    if (!(p->width > 0) || !(p->height > 0)) return fail(SALP_EINVAL, "width/height must be positive");
    if (p->num_food_items < 0 || p->num_food_items > SALP_MAX_FOOD)
        return fail(SALP_EUNSUPPORTED, "num_food_items=%d outside 0..%d", p->num_food_items, SALP_MAX_FOOD);
    if (p->max_observed_food < 0 || p->max_observed_food > SALP_MAX_OBSERVED_FOOD)
        return fail(SALP_EUNSUPPORTED, "max_observed_food=%d outside 0..%d", p->max_observed_food, SALP_MAX_OBSERVED_FOOD);
    if (p->max_steps_without_food < 0) return fail(SALP_EINVAL, "max_steps_without_food must be >= 0");
    memset(out, 0, sizeof(*out));
    out->real_bytes = dtype == SALP_F64 ? 8 : 4;
    out->obs_dim = 10 + 4 * p->max_observed_food + 2;
    out->act_dim = p->forced_breathing ? 1 : 2;
    out->food_slots = pick_food_slots(std::max(1, p->num_food_items));
    out->body_reals = 8;
    out->ctrl_ints = 8;
    out->stats_len = salp::kStatsLen;
    return SALP_OK;
}

int salp_create(const SalpParams *p, int64_t num_envs, int64_t env_id0, int device, int dtype, SalpHandle **out) {
    if (!out) return fail(SALP_EINVAL, "null out pointer");
    *out = nullptr;
    SalpLayout lay;
    int rc = salp_query_layout(p, dtype, &lay);
    if (rc != SALP_OK) return rc;
    if (num_envs <= 0 || num_envs > (int64_t)1 << 30) return fail(SALP_EINVAL, "num_envs=%lld outside 1..2^30", (long long)num_envs);
    if (env_id0 < 0) return fail(SALP_EINVAL, "env_id0 must be >= 0");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(SALP_ECUDA, "no CUDA device available (%s); this library has no CPU fallback",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(SALP_EINVAL, "device %d outside 0..%d", device, ndev - 1);
    cudaDeviceProp prop;
    SALP_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(SALP_EUNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);

    SalpHandle *h = new SalpHandle();
    memset(h, 0, sizeof(*h));
    h->params = *p;
    h->layout = lay;
    h->num_envs = num_envs;
    h->env_id0 = env_id0;
    h->device = device;
    h->dtype = dtype;
    salp::KernelArgs &a = h->base;
    a.N = num_envs;
    a.first = 0;
    a.count = num_envs;
    a.K = 1;
    a.obs_dim = lay.obs_dim;
    a.observed = p->max_observed_food;
    a.act_dim = lay.act_dim;
    a.forced = p->forced_breathing ? 1 : 0;
    a.random_food_count = p->random_food_count ? 1 : 0;
    a.respawn = p->respawn_food ? 1 : 0;
    a.autoreset = (p->flags & SALP_AUTORESET) ? 1 : 0;
    a.max_steps = p->max_steps_without_food;
    a.num_food = p->num_food_items;
    a.W = p->width;
    a.H = p->height;
    a.diag = std::sqrt(p->width * p->width + p->height * p->height);  // S:396
    a.inv_w = 1.0 / a.W;
    a.inv_h = 1.0 / a.H;
    a.inv_diag = 1.0 / a.diag;
    a.food_reward = p->food_reward;
    a.collision_penalty = p->collision_penalty;
    a.time_penalty = p->time_penalty;
    a.efficiency_bonus = p->efficiency_bonus;
    a.proximity_w = p->proximity_reward_weight;
    for (int t = 0; t < 128; ++t) {  // R:223 int(self.exhale_duration * max(self.water_volume, 0.3)), wv = t/120
        const double wv = (double)std::min(t, 120) / 120.0;
        a.exhale_lut[t] = (uint8_t)(int)(150.0 * std::max(wv, 0.3));
    }
    a.fW = (float)a.W; a.fH = (float)a.H; a.f_diag = (float)a.diag;
    a.f_inv_w = (float)a.inv_w; a.f_inv_h = (float)a.inv_h; a.f_inv_diag = (float)a.inv_diag;
    a.f_food_reward = (float)p->food_reward; a.f_collision_penalty = (float)p->collision_penalty;
    a.f_time_penalty = (float)p->time_penalty; a.f_efficiency_bonus = (float)p->efficiency_bonus;
    a.f_proximity_w = (float)p->proximity_reward_weight;
    a.rng_mode = (p->flags & SALP_RNG_PHILOX) ? 1 : 0;
    a.rng_seed = 0;
    a.env_id0 = env_id0;
    {   // L2 plan of the forced-breathing kernel (salp_step_fast.cuh), bit mask as KernelArgs::l2_mode.  Default 2: what is
        // written once and never read by the kernel (observation tiles, rewards, flags) and what is read once (actions)
        // leaves L2 first, so the state records -- read and rewritten every step -- survive in the 126 MB L2 from one step
        // to the next on their own.  Measured on B200, 1 M single-food envs: 50.5 us without hints, 43.1 us with mode 2,
        // 45.0 / 44.6 us when the state is additionally pinned evict_last (modes 3 / 7).  SALP_L2_MODE overrides.
        int mode = 2;
        if (const char *ev = getenv("SALP_L2_MODE")) mode = atoi(ev);
        a.l2_mode = mode;
    }
    {   // forced-breathing cycle table: parameter data of the handle, like the exhale LUT, but too large for the constant bank
        salp::CycleEntry tab[salp::kCycleLen];
        salp::build_forced_cycle(a.W, a.H, a.exhale_lut, tab, nullptr, nullptr, nullptr);
        cudaError_t ce = cudaSetDevice(device);
        if (ce == cudaSuccess) ce = cudaMalloc(&h->cycle_tab, sizeof(tab));
        if (ce == cudaSuccess) ce = cudaMemcpy(h->cycle_tab, tab, sizeof(tab), cudaMemcpyHostToDevice);
        if (ce != cudaSuccess) { if (h->cycle_tab) cudaFree(h->cycle_tab); delete h; return cuda_fail(ce, "forced-breathing cycle table"); }
        a.cycle_tab = h->cycle_tab;
    }
    *out = h;
    return SALP_OK;
}

int salp_forced_cycle(const SalpParams *p, double *wv, double *ellipse_a, double *ellipse_b, int32_t *fsm, int32_t *touch) {
    if (!p) return fail(SALP_EINVAL, "null argument");
    uint8_t lut[128];
    for (int t = 0; t < 128; ++t) lut[t] = (uint8_t)(int)(150.0 * std::max((double)std::min(t, 120) / 120.0, 0.3));
    salp::CycleEntry tab[salp::kCycleLen];
    salp::build_forced_cycle(p->width, p->height, lut, tab, wv, ellipse_a, ellipse_b);
    for (int i = 0; i < salp::kCycleLen; ++i) {
        if (fsm) fsm[i] = (int32_t)(tab[i].fsm_touch & 0x0fffffffu);
        if (touch) touch[i] = (int32_t)(tab[i].fsm_touch >> 28);
    }
    return SALP_OK;
}

int salp_set_rng_seed(SalpHandle *h, uint64_t seed) {
    if (!h) return fail(SALP_EINVAL, "null handle");
    h->base.rng_seed = seed;
    return SALP_OK;
}

int salp_destroy(SalpHandle *h) {
    if (!h) return SALP_OK;
    if (h->cycle_tab) { cudaSetDevice(h->device); cudaFree(h->cycle_tab); }
    if (h->host_streams_ready) {
        cudaSetDevice(h->device);
        for (auto &s : h->host_streams) cudaStreamDestroy(s);
        cudaEventDestroy(h->host_event);
        for (auto &ev : h->chunk_done) cudaEventDestroy(ev);
    }
    delete h;
    return SALP_OK;
}

int salp_bind(SalpHandle *h, const SalpBuffers *b) {
    if (!h || !b) return fail(SALP_EINVAL, "null argument");
    if (!b->body || !b->ctrl || !b->ep_return || !b->food || !b->obs || !b->reward || !b->terminated || !b->truncated)
        return fail(SALP_EINVAL, "body/ctrl/ep_return/food/obs/reward/terminated/truncated must be non-null");
    const bool philox = (h->params.flags & SALP_RNG_PHILOX) != 0;
    if (!philox) {
        if (!b->food_words || !b->jit_words) return fail(SALP_EINVAL, "RNG word tables must be non-null (or create the handle with SALP_RNG_PHILOX)");
        if (b->table_rows <= 0 || b->table_rows > h->num_envs) return fail(SALP_EINVAL, "table_rows must be in 1..num_envs");
        if (!is_pow2(b->food_len) || !is_pow2(b->jit_len) || b->jit_len < 2 || b->food_len < 2)
            return fail(SALP_EINVAL, "food_len/jit_len must be powers of two >= 2");
        if (b->food_len > ((int64_t)1 << 30) || b->jit_len > ((int64_t)1 << 30)) return fail(SALP_EINVAL, "table rows too long");
    }
    if (b->max_substeps < 1) return fail(SALP_EINVAL, "max_substeps must be >= 1");
    const uintptr_t align16[] = {(uintptr_t)b->body, (uintptr_t)b->ctrl, (uintptr_t)b->food, (uintptr_t)b->obs,
                                 (uintptr_t)b->final_obs, (uintptr_t)b->final_info, (uintptr_t)b->jit_words};
    for (uintptr_t p : align16)
        if (p % 32 != 0) return fail(SALP_EINVAL, "device buffers must be 32-byte aligned");
    h->buf = *b;
    salp::KernelArgs &a = h->base;
    a.body = b->body;
    a.ctrl = reinterpret_cast<int4 *>(b->ctrl);
    a.ep_return = b->ep_return;
    a.food = b->food;
    a.obs = b->obs;
    a.reward = b->reward;
    a.terminated = b->terminated;
    a.truncated = b->truncated;
    a.events = b->events;
    a.final_obs = b->final_obs;
    a.final_info = b->final_info;
    a.stats = b->stats;
    a.food_words = b->food_words;
    a.jit_words = b->jit_words;
    a.rows = philox ? 1u : (unsigned)b->table_rows;
    a.food_len = philox ? 2u : (unsigned)b->food_len;
    a.jit_len = philox ? 2u : (unsigned)b->jit_len;
    a.food_len_mask = a.food_len - 1;
    a.jit_len_mask = a.jit_len - 1;
    h->bound = true;
    return SALP_OK;
}

static int launch_step(SalpHandle *h, const salp::KernelArgs &a, cudaStream_t s) {
    cudaError_t e;
    bool handled = false;
    if (h->dtype == SALP_F32 && !(h->params.flags & SALP_GENERIC_KERNEL)) {
        e = salp::launch_step_fast_f32(a, h->layout.food_slots, s, &handled);  // forced breathing, 1 / 12 foods, 3 observed
        if (e != cudaSuccess) return cuda_fail(e, "salp_step (forced-cycle kernel) launch");
    }
    if (!handled)
        e = h->dtype == SALP_F64 ? salp::launch_step_f64(a, h->layout.food_slots, s)
                                 : salp::launch_step_f32(a, h->layout.food_slots, s);
    if (e != cudaSuccess) return cuda_fail(e, "salp_step kernel launch");
    ++h->launches;
    return SALP_OK;
}

int salp_reset(SalpHandle *h, const uint8_t *mask, int rewind_rng, void *stream) {
    if (!h) return fail(SALP_EINVAL, "null handle");
    if (!h->bound) return fail(SALP_EUNBOUND, "salp_bind has not been called");
    SALP_CUDA(cudaSetDevice(h->device));
    salp::KernelArgs a = h->base;
    a.reset_mask = mask;
    a.rewind_rng = rewind_rng ? 1 : 0;
    cudaError_t e = h->dtype == SALP_F64 ? salp::launch_reset_f64(a, h->layout.food_slots, (cudaStream_t)stream)
                                         : salp::launch_reset_f32(a, h->layout.food_slots, (cudaStream_t)stream);
    if (e != cudaSuccess) return cuda_fail(e, "salp_reset kernel launch");
    ++h->launches;
    return SALP_OK;
}

int salp_refresh_rng(SalpHandle *h, void *stream) {
    if (!h) return fail(SALP_EINVAL, "null handle");
    if (!h->bound) return fail(SALP_EUNBOUND, "salp_bind has not been called");
    SALP_CUDA(cudaSetDevice(h->device));
    cudaError_t e = salp::launch_refresh_rng(h->base, (cudaStream_t)stream);
    if (e != cudaSuccess) return cuda_fail(e, "salp_refresh_rng kernel launch");
    ++h->launches;
    return SALP_OK;
}

int salp_observe(SalpHandle *h, int slot, void *stream) {
    if (!h) return fail(SALP_EINVAL, "null handle");
    if (!h->bound) return fail(SALP_EUNBOUND, "salp_bind has not been called");
    if (slot < 0 || slot >= h->buf.max_substeps) return fail(SALP_EINVAL, "slot=%d outside 0..max_substeps-1=%d", slot, h->buf.max_substeps - 1);
    SALP_CUDA(cudaSetDevice(h->device));
    salp::KernelArgs a = h->base;
    a.obs += (size_t)h->num_envs * (size_t)slot * (size_t)h->layout.obs_dim;
    cudaError_t e = h->dtype == SALP_F64 ? salp::launch_observe_f64(a, (cudaStream_t)stream) : salp::launch_observe_f32(a, (cudaStream_t)stream);
    if (e != cudaSuccess) return cuda_fail(e, "salp_observe kernel launch");
    ++h->launches;
    return SALP_OK;
}

int salp_set_num_food_items(SalpHandle *h, int32_t num_food_items) {
    if (!h) return fail(SALP_EINVAL, "null handle");
    if (num_food_items < 0) num_food_items = 0;  // S:36 max(0, num_food_items)
    if (num_food_items > h->layout.food_slots)
        return fail(SALP_EUNSUPPORTED, "num_food_items=%d exceeds the %d food slots this handle was created with; create a new handle",
                    (int)num_food_items, (int)h->layout.food_slots);
    h->params.num_food_items = num_food_items;
    h->base.num_food = num_food_items;
    return SALP_OK;
}

int salp_step(SalpHandle *h, int K, const void *actions, void *stream) {
    if (!h) return fail(SALP_EINVAL, "null handle");
    if (!h->bound) return fail(SALP_EUNBOUND, "salp_bind has not been called");
    if (!actions) return fail(SALP_EINVAL, "actions pointer is null");
    if ((uintptr_t)actions % (size_t)(h->layout.real_bytes * h->layout.act_dim) != 0)
        return fail(SALP_EINVAL, "actions pointer is not aligned to one action record");
    if (K < 1 || K > h->buf.max_substeps) return fail(SALP_EINVAL, "K=%d outside 1..max_substeps=%d", K, h->buf.max_substeps);
    SALP_CUDA(cudaSetDevice(h->device));
    salp::KernelArgs a = h->base;
    a.K = K;
    a.actions = actions;
    return launch_step(h, a, (cudaStream_t)stream);
}

int salp_step_slot(SalpHandle *h, int slot, const void *actions, void *stream) {
    if (!h) return fail(SALP_EINVAL, "null handle");
    if (!h->bound) return fail(SALP_EUNBOUND, "salp_bind has not been called");
    if (!actions) return fail(SALP_EINVAL, "actions pointer is null");
    if (slot < 0 || slot >= h->buf.max_substeps) return fail(SALP_EINVAL, "slot=%d outside 0..max_substeps-1=%d", slot, h->buf.max_substeps - 1);
    SALP_CUDA(cudaSetDevice(h->device));
    salp::KernelArgs a = h->base;
    const size_t n = (size_t)h->num_envs * (size_t)slot, rb = (size_t)h->layout.real_bytes, od = (size_t)h->layout.obs_dim;
    a.K = 1;
    a.actions = actions;
    a.obs += n * od;
    a.reward = (char *)a.reward + n * rb;
    a.terminated += n;
    a.truncated += n;
    if (a.events) a.events += n;
    if (a.final_obs) a.final_obs += n * od;
    if (a.final_info) a.final_info = (char *)a.final_info + n * 4 * rb;
    return launch_step(h, a, (cudaStream_t)stream);
}

int salp_step_host(SalpHandle *h, void *actions_staging, const void *actions_host, float *obs_host,
                   void *reward_host, uint8_t *terminated_host, uint8_t *truncated_host, void *caller_stream) {
    if (!h) return fail(SALP_EINVAL, "null handle");
    if (!h->bound) return fail(SALP_EUNBOUND, "salp_bind has not been called");
    if (!actions_staging || !actions_host) return fail(SALP_EINVAL, "actions (device staging and host) must be non-null");
    SALP_CUDA(cudaSetDevice(h->device));
    if (!h->host_streams_ready) {
        for (auto &s : h->host_streams) SALP_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        SALP_CUDA(cudaEventCreateWithFlags(&h->host_event, cudaEventDisableTiming));
        for (auto &ev : h->chunk_done) SALP_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        h->host_streams_ready = true;
    }
    // the chunk streams start after everything already queued on the caller's stream (e.g. the reset, or set_attr copies)
    SALP_CUDA(cudaEventRecord(h->host_event, (cudaStream_t)caller_stream));
    const int64_t N = h->num_envs;
    const size_t rb = (size_t)h->layout.real_bytes, ad = (size_t)h->layout.act_dim, od = (size_t)h->layout.obs_dim;
    // chunk boundaries on CTA / tile granularity so every chunk but the last is a whole number of tiles
    const int64_t tiles = (N + salp::kHostChunkAlign - 1) / salp::kHostChunkAlign;
    const int chunks = (int)std::min<int64_t>(host_chunks(), tiles);
    int rc = SALP_OK;
    for (int c = 0; c < chunks && rc == SALP_OK; ++c) {
        const int64_t t0 = tiles * c / chunks, t1 = tiles * (c + 1) / chunks;
        const int64_t lo = t0 * salp::kHostChunkAlign, hi = std::min<int64_t>(N, t1 * salp::kHostChunkAlign);
        const size_t n = (size_t)(hi - lo);
        cudaStream_t s = h->host_streams[c];
        SALP_CUDA(cudaStreamWaitEvent(s, h->host_event, 0));
        SALP_CUDA(cudaMemcpyAsync((char *)actions_staging + lo * ad * rb, (const char *)actions_host + lo * ad * rb,
                                  n * ad * rb, cudaMemcpyHostToDevice, s));
        salp::KernelArgs a = h->base;
        a.K = 1;
        a.actions = actions_staging;
        a.first = lo;
        a.count = hi - lo;
        rc = launch_step(h, a, s);
        if (rc != SALP_OK) break;
        SALP_CUDA(cudaEventRecord(h->chunk_done[c], s));
        // the observation rows (94 % of the bytes) leave chunk by chunk, overlapping the later chunks' kernels
        if (obs_host)
            SALP_CUDA(cudaMemcpyAsync(obs_host + lo * od, h->buf.obs + lo * od, n * od * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    if (rc == SALP_OK && chunks > 0) {
        // rewards and flags (6 bytes per env) as ONE copy per array once every chunk's kernel has run: 3 copies at full link
        // rate instead of 24 short ones (a 128 KB copy costs ~5 us of link time whatever its size)
        cudaStream_t s = h->host_streams[chunks - 1];
        for (int c = 0; c + 1 < chunks; ++c) SALP_CUDA(cudaStreamWaitEvent(s, h->chunk_done[c], 0));
        if (reward_host) SALP_CUDA(cudaMemcpyAsync(reward_host, h->buf.reward, (size_t)N * rb, cudaMemcpyDeviceToHost, s));
        if (terminated_host) SALP_CUDA(cudaMemcpyAsync(terminated_host, h->buf.terminated, (size_t)N, cudaMemcpyDeviceToHost, s));
        if (truncated_host) SALP_CUDA(cudaMemcpyAsync(truncated_host, h->buf.truncated, (size_t)N, cudaMemcpyDeviceToHost, s));
    }
    for (int c = 0; c < chunks; ++c) {
        cudaError_t e = cudaStreamSynchronize(h->host_streams[c]);
        if (e != cudaSuccess && rc == SALP_OK) rc = cuda_fail(e, "salp_step_host synchronize");
    }
    return rc;
}

int64_t salp_launch_count(const SalpHandle *h) { return h ? h->launches : 0; }

int salp_stats(SalpHandle *h, double *out8_host, int reset, void *stream) {
    if (!h || !out8_host) return fail(SALP_EINVAL, "null argument");
    if (!h->bound) return fail(SALP_EUNBOUND, "salp_bind has not been called");
    if (!h->buf.stats) return fail(SALP_EUNBOUND, "no stats buffer is bound (SalpBuffers.stats is NULL)");
    SALP_CUDA(cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)stream;
    SALP_CUDA(cudaMemcpyAsync(out8_host, h->buf.stats, salp::kStatsLen * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (reset) SALP_CUDA(cudaMemsetAsync(h->buf.stats, 0, salp::kStatsLen * sizeof(double), s));
    SALP_CUDA(cudaStreamSynchronize(s));
    return SALP_OK;
}

// ---------------------------------------------------------------------------------------------
// MT19937 word streams (host).  Reproduces CPython's random.seed(int) and numpy's legacy
// np.random.seed(int) so the tables hold exactly the words the reference would consume.
// ---------------------------------------------------------------------------------------------
namespace {
struct MT {
    uint32_t mt[624];
    int idx;
    void init_genrand(uint32_t s) {
        mt[0] = s;
        for (int i = 1; i < 624; ++i) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (uint32_t)i;
        idx = 624;
    }
    void init_by_array(const uint32_t *key, int len) {
        init_genrand(19650218u);
        int i = 1, j = 0;
        for (int k = std::max(624, len); k; --k) {
            mt[i] = (mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1664525u)) + key[j] + (uint32_t)j;
            if (++i >= 624) { mt[0] = mt[623]; i = 1; }
            if (++j >= len) j = 0;
        }
        for (int k = 623; k; --k) {
            mt[i] = (mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1566083941u)) - (uint32_t)i;
            if (++i >= 624) { mt[0] = mt[623]; i = 1; }
        }
        mt[0] = 0x80000000u;
        idx = 624;
    }
    void twist() {
        for (int k = 0; k < 624; ++k) {
            const uint32_t y = (mt[k] & 0x80000000u) | (mt[(k + 1) % 624] & 0x7fffffffu);
            mt[k] = mt[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        idx = 0;
    }
    uint32_t next() {
        if (idx >= 624) twist();
        uint32_t y = mt[idx++];
        y ^= y >> 11;
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= y >> 18;
        return y;
    }
};
}  // namespace

int salp_fill_rng_words(uint32_t *dst, int64_t rows, int64_t len, uint64_t seed0, int64_t first_row, int kind,
                        int num_threads) {
    if (!dst || rows < 0 || len <= 0 || first_row < 0) return fail(SALP_EINVAL, "bad arguments");
    if (kind != 0 && kind != 1) return fail(SALP_EINVAL, "kind must be 0 (python random) or 1 (numpy legacy)");
    if (seed0 + (uint64_t)first_row + (uint64_t)rows > 0xffffffffull) return fail(SALP_EINVAL, "seeds must stay below 2^32");
    int nt = num_threads > 0 ? num_threads : (int)std::thread::hardware_concurrency();
    nt = (int)std::max<int64_t>(1, std::min<int64_t>(nt, rows));
    auto work = [=](int t) {
        MT g;
        for (int64_t r = rows * t / nt; r < rows * (t + 1) / nt; ++r) {
            const uint32_t s = (uint32_t)(seed0 + (uint64_t)first_row + (uint64_t)r);
            if (kind == 0) g.init_by_array(&s, 1);
            else g.init_genrand(s);
            uint32_t *out = dst + r * len;
            for (int64_t i = 0; i < len; ++i) out[i] = g.next();
        }
    };
    if (nt == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < nt; ++t) th.emplace_back(work, t);
        for (auto &t : th) t.join();
    }
    return SALP_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// salp_step_host_noalloc -- the host-buffer step for callers that cannot allocate device memory.
//
// salp_step_host needs a caller-owned DEVICE staging buffer for the actions the step kernel reads
// (actions_staging) and creates eight streams and an event on its first call.  This variant takes that
// scratch from a file-scope device array instead and works on the caller's stream alone: it calls no
// cudaMalloc / cudaFree / cudaStreamCreate / cudaEventCreate, so the caller needs no allocator at all.
// Every GPU that loads the library holds SALP_NOALLOC_STAGING_BYTES (1 MiB) for it until exit.
//
// In host code `g_noalloc_staging` is only a host-side stand-in for the device array, NOT a device
// address: the address the copy engine and the kernel need comes from cudaGetSymbolAddress, asked once
// per device with that device current (the first call of the entry point on a device is the set-up).
// ---------------------------------------------------------------------------------------------
#include <mutex>

static __device__ __align__(256) unsigned char g_noalloc_staging[SALP_NOALLOC_STAGING_BYTES];

namespace {

constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
void *g_noalloc_ptr[kNoallocMaxDevices] = {};

// device address of g_noalloc_staging on `device`, which must be the current device
int noalloc_staging(int device, void **out) {
    if (device < 0 || device >= kNoallocMaxDevices) return fail(SALP_EUNSUPPORTED, "device ordinal %d out of range", device);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_staging);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(static action staging)");
        if (!p || (uintptr_t)p % 256 != 0) return fail(SALP_ECUDA, "static action staging resolved to a bad device address %p", p);
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return SALP_OK;
}

}  // namespace

extern "C" int salp_step_host_noalloc(SalpHandle *h, const void *actions_host, float *obs_host, void *reward_host,
                                      uint8_t *terminated_host, uint8_t *truncated_host, void *caller_stream) {
    if (!h) return fail(SALP_EINVAL, "null handle");
    if (!h->bound) return fail(SALP_EUNBOUND, "salp_bind has not been called");
    if (!actions_host) return fail(SALP_EINVAL, "actions_host must be non-null");
    const int64_t N = h->num_envs;
    const size_t rb = (size_t)h->layout.real_bytes, ad = (size_t)h->layout.act_dim, od = (size_t)h->layout.obs_dim;
    const size_t need = (size_t)N * ad * rb;
    if (need > (size_t)SALP_NOALLOC_STAGING_BYTES)
        return fail(SALP_EUNSUPPORTED,
                    "salp_step_host_noalloc: %lld envs x %zu-byte action records = %zu bytes exceed the %d-byte static staging "
                    "array (SALP_NOALLOC_STAGING_BYTES); use salp_step_host with a caller-owned staging buffer",
                    (long long)N, ad * rb, need, (int)SALP_NOALLOC_STAGING_BYTES);
    SALP_CUDA(cudaSetDevice(h->device));
    void *staging = nullptr;
    if (int rc = noalloc_staging(h->device, &staging)) return rc;
    cudaStream_t s = (cudaStream_t)caller_stream;
    // Nothing in the staging array has to start at zero: the kernel reads exactly the N action records written here.
    SALP_CUDA(cudaMemcpyAsync(staging, actions_host, need, cudaMemcpyHostToDevice, s));
    salp::KernelArgs a = h->base;
    a.K = 1;
    a.actions = staging;
    if (int rc = launch_step(h, a, s)) return rc;
    const size_t n = (size_t)N;
    if (obs_host) SALP_CUDA(cudaMemcpyAsync(obs_host, h->buf.obs, n * od * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (reward_host) SALP_CUDA(cudaMemcpyAsync(reward_host, h->buf.reward, n * rb, cudaMemcpyDeviceToHost, s));
    if (terminated_host) SALP_CUDA(cudaMemcpyAsync(terminated_host, h->buf.terminated, n, cudaMemcpyDeviceToHost, s));
    if (truncated_host) SALP_CUDA(cudaMemcpyAsync(truncated_host, h->buf.truncated, n, cudaMemcpyDeviceToHost, s));
    cudaError_t e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return cuda_fail(e, "salp_step_host_noalloc synchronize");
    return SALP_OK;
}
