// salp_cycle_capi.cu -- extern "C" boundary of the model (B) cycle env (include/salp_b200.h: salp_cycle_*).
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include <mutex>

#include "salp_cycle.cuh"

namespace salp { int set_error(int code, const char *msg); }  // salp_capi.cu owns the thread-local error text

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_mu;
double *g_ct[kMaxDevices] = {};

// ct[k] = the value of Robot.cycle_time after k executions of `self.cycle_time += self.dt` from 0.0
// (robot.py:351,391): one table per device, built on first use (a synchronous copy -- make the first
// salp_cycle_reset / salp_cycle_step call outside any stream capture).
int time_table(const double **out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return salp::set_error(SALP_ECUDA, cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return salp::set_error(SALP_EUNSUPPORTED, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_ct[dev]) {
        static double host[SALP_CYCLE_MAX_SUBSTEPS + 1];
        double t = 0.0;
        for (int k = 0; k <= SALP_CYCLE_MAX_SUBSTEPS; ++k) { host[k] = t; t += 0.01; }
        double *d = nullptr;
        e = cudaMalloc(&d, sizeof host);
        if (e == cudaSuccess) e = cudaMemcpy(d, host, sizeof host, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { if (d) cudaFree(d); return salp::set_error(SALP_ECUDA, cudaGetErrorString(e)); }
        g_ct[dev] = d;
    }
    *out = g_ct[dev];
    return SALP_OK;
}

int fill_args(const SalpCycleParams *p, const SalpCycleBuffers *b, salp::CycleArgs &a) {
    if (!p || !b) return salp::set_error(SALP_EINVAL, "null argument");
    if (!b->state || !b->ctrl || !b->obs || !b->reward || !b->terminated || !b->truncated || !b->words)
        return salp::set_error(SALP_EINVAL, "salp_cycle: a required buffer pointer is null");
    if (b->n <= 0 || b->table_rows <= 0 || b->words_len < 4 || (b->words_len & (b->words_len - 1)))
        return salp::set_error(SALP_EINVAL, "salp_cycle: need n > 0, table_rows > 0, words_len a power of two >= 4");
    if (b->dtype != SALP_F32 && b->dtype != SALP_F64) return salp::set_error(SALP_EINVAL, "salp_cycle: bad dtype");
    if (!(p->init_length > 0) || !(p->init_width > 0) || !(p->density > 0) || !(p->nozzle_area > 0) || !(p->dry_mass >= 0) ||
        !(p->max_contraction > 0) || !(p->max_contraction < p->init_length))
        return salp::set_error(SALP_EINVAL, "salp_cycle: robot parameters out of range");
    if (p->max_cycles <= 0 || p->max_cycles > 0xFFFF) return salp::set_error(SALP_EINVAL, "salp_cycle: max_cycles must be in 1..65535");
    if (p->max_substeps < 0 || p->max_substeps > SALP_CYCLE_MAX_SUBSTEPS)
        return salp::set_error(SALP_EINVAL, "salp_cycle: max_substeps must be in 0..SALP_CYCLE_MAX_SUBSTEPS");
    const uintptr_t align = b->dtype == SALP_F64 ? 32 : 16;
    if ((uintptr_t)b->state % align || (uintptr_t)b->ctrl % 16 || (uintptr_t)b->obs % 8 || (b->final_obs && (uintptr_t)b->final_obs % 8))
        return salp::set_error(SALP_EINVAL, "salp_cycle: state must be 16/32-byte, ctrl 16-byte, obs 8-byte aligned");
    a.state = b->state; a.ctrl = reinterpret_cast<int4 *>(b->ctrl); a.actions = nullptr; a.obs = b->obs; a.reward = b->reward;
    a.terminated = b->terminated; a.truncated = b->truncated; a.n_sub = b->n_sub; a.final_obs = b->final_obs;
    a.stats = b->stats; a.words = b->words; a.mask = nullptr;
    a.n = b->n; a.env_id0 = b->env_id0; a.rows = b->table_rows; a.len_mask = (unsigned)(b->words_len - 1);
    a.max_cycles = p->max_cycles; a.max_substeps = p->max_substeps ? p->max_substeps : SALP_CYCLE_MAX_SUBSTEPS;
    a.autoreset = (p->flags & SALP_CYCLE_AUTORESET) ? 1 : 0; a.rewind_rng = 0;
    a.sort_buckets = 0;  // sized per slice by the kernel
    a.dry_mass = p->dry_mass; a.L0 = p->init_length; a.W0 = p->init_width; a.density = p->density;
    a.noz_l12 = p->nozzle_length1 + p->nozzle_length2; a.noz_area = p->nozzle_area; a.noz_mass = p->nozzle_mass;
    // generate_target_point("random") bounds (salp_robot_env.py:266-277)
    const double scale = 200.0, margin = 50.0;
    a.x_min = (-p->width / 2 + margin) / scale; a.x_max = (p->width / 2 - margin) / scale;
    a.y_min = (-p->height / 2 + margin) / scale; a.y_max = (p->height / 2 - margin) / scale;
    // _get_drag_coefficient normalisation (robot.py:638-641)
    a.init_aspect = p->init_length / p->init_width;
    const double cl = p->init_length - p->max_contraction;
    const double cw = p->init_length - cl + p->init_width;
    a.min_aspect = cl / cw;
    a.f_min_aspect = (float)a.min_aspect; a.f_inv_aspect_span = (float)(1.0 / (a.init_aspect - a.min_aspect));
    a.f_density = (float)p->density; a.f_inv_noz_area = (float)(1.0 / p->nozzle_area);
    a.f_L0 = (float)p->init_length; a.f_W0 = (float)p->init_width; a.f_l12 = (float)a.noz_l12;
    a.f_noz_mass = (float)p->nozzle_mass; a.f_m_fixed = (float)(p->dry_mass + p->nozzle_mass);
    // Robot.reset geometry (robot.py:303-311), reference operation order
    const double kPi = 3.141592653589793;
    const double hw = p->init_width / 2, hl = p->init_length / 2;
    a.rest_volume = 4.0 / 3.0 * kPi * hl * (hw * hw);
    const double mass = p->dry_mass + p->density * a.rest_volume + p->nozzle_mass;
    const double rx = -a.noz_l12 + (-p->init_length / 2);
    const double nrm = sqrt(rx * rx + 0.0 + 0.0);
    const double In = p->nozzle_mass * (nrm * nrm);
    a.rest_I[0] = 0.2 * mass * (hw * hw + hw * hw) + In * 0.0;
    a.rest_I[1] = 0.2 * mass * (hl * hl + hw * hw) + In;
    a.rest_I[2] = 0.2 * mass * (hw * hw + hl * hl) + In;
    return time_table(&a.ct);
}

}  // namespace

extern "C" int salp_cycle_reset(const SalpCycleParams *p, const SalpCycleBuffers *b, const uint8_t *mask, int rewind_rng,
                                void *cuda_stream) {
    salp::CycleArgs a;
    if (int rc = fill_args(p, b, a)) return rc;
    a.mask = mask;
    a.rewind_rng = rewind_rng ? 1 : 0;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    cudaError_t e = b->dtype == SALP_F64 ? salp::launch_cycle_reset_f64(a, s) : salp::launch_cycle_reset_f32(a, s);
    if (e != cudaSuccess) return salp::set_error(SALP_ECUDA, cudaGetErrorString(e));
    return SALP_OK;
}

extern "C" int salp_cycle_step(const SalpCycleParams *p, const SalpCycleBuffers *b, const void *actions, void *cuda_stream) {
    salp::CycleArgs a;
    if (int rc = fill_args(p, b, a)) return rc;
    if (!actions) return salp::set_error(SALP_EINVAL, "salp_cycle_step: actions is null");
    a.actions = actions;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    cudaError_t e = b->dtype == SALP_F64 ? salp::launch_cycle_step_f64(a, s) : salp::launch_cycle_step_f32(a, s);
    if (e != cudaSuccess) return salp::set_error(SALP_ECUDA, cudaGetErrorString(e));
    return SALP_OK;
}
