// salp_cycle.cuh -- model (B): one env step = one breathing cycle of the 3-D Newton-Euler robot.
//
// What it replaces (reference tree, Python/numpy):
//   SalpRobotEnv.step / reset / _calculate_reward / _get_observation   src/salp/environments/salp_robot_env.py:100-242,400-420
//   Nozzle.solve_angles, rotation matrices, jet direction              src/salp/environments/robot.py:55-83,115-130,153-178
//   Robot.set_control, step_through_cycle, step, Newton / Euler equations, geometry   robot.py:335-777
//
// One thread owns one environment for the whole cycle (0 .. 1450 integrator substeps, dt = 0.01 s), state in
// registers.  The work per env step is ~10^5 flops against ~250 bytes of state traffic: this kernel is bound
// by the FP32 / SFU pipes, not by HBM.  What the design does about that:
//   * substep counts and phase boundaries are integers found once per cycle from the table of the reference's
//     accumulated cycle times (cycle_time += 0.01 from 0.0 is the same double sequence for every env), so the
//     loop carries no double accumulator and the discrete trajectory is exact in both builds;
//   * body geometry (length, width, area, water volume, mass, drag coefficient, inertia and their rates) only
//     changes while the body contracts or releases: that block is skipped for the coast / rest substeps (the
//     majority) and hands the dynamics a dozen coefficients.  The double build evaluates it with the reference's
//     expressions; the float build folds every shape factor into the coefficients and gets the volume / inertia
//     rates from exact difference expansions in float (cyc_shape_f), no double arithmetic;
//   * sin / cos of the Euler angles are carried from the end of one substep to the start of the next (the
//     reference re-evaluates them on the same values);
//   * lanes of a warp would get cycles of different lengths (the coast time is an action): one CTA per SM sorts
//     its env range by substep count in shared memory and its 20 warps pull 32-env chunks of nearly equal length,
//     longest first, from a shared counter (salp_cycle_step_kernel).
// The double build keeps the reference's operation order (compiled with -fmad=false); the float build uses
// the closed form of the nozzle inverse kinematics (the reference's arcsin argument is analytically
// -sign(yaw): its result carries ~2e-8 of rounding noise, far below float resolution) and SFU sin/cos/rsqrt/rcp.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/salp_b200.h"

namespace salp {

constexpr int kCycleStats = 8;

struct CycleArgs {
    void *state;       // real [6][N][4]
    int4 *ctrl;        // i32  [N][4]
    const void *actions;
    float *obs;
    void *reward;
    uint8_t *terminated, *truncated;
    int32_t *n_sub;
    float *final_obs;
    double *stats;
    const uint32_t *words;
    const uint8_t *mask;   // reset kernel only
    const double *ct;      // ct[k] = k-fold accumulated 0.01 (device table, SALP_CYCLE_MAX_SUBSTEPS + 1 entries)
    long long n, env_id0, rows;
    unsigned len_mask;
    int max_cycles, max_substeps, autoreset, rewind_rng;
    int sort_buckets;  // breathing-length groups in front of the substep count in the sort key (0 = by slice size)
    // robot / nozzle / tank parameters
    double dry_mass, L0, W0, density, noz_l12, noz_area, noz_mass;
    double x_min, x_max, y_min, y_max;
    double min_aspect, init_aspect;       // drag-coefficient normalisation (robot.py:638-644), host-computed
    float f_min_aspect, f_inv_aspect_span, f_density, f_inv_noz_area;  // float build copies
    float f_L0, f_W0, f_l12, f_noz_mass, f_m_fixed;                    // m_fixed = dry mass + nozzle mass
    // rest-state geometry, host-computed in the reference's operation order
    double rest_volume, rest_I[3];
};

__device__ __forceinline__ double cyc_clip(double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); }

__device__ __forceinline__ double cyc_next_double(const CycleArgs &a, long long row, unsigned &cur) {
    const uint32_t *r = a.words + row * (long long)(a.len_mask + 1u);
    const uint32_t hi = __ldg(r + (cur & a.len_mask)) >> 5;
    const uint32_t lo = __ldg(r + ((cur + 1u) & a.len_mask)) >> 6;
    cur += 2u;
    return (hi * 67108864.0 + lo) / 9007199254740992.0;
}

// largest k in [0, K] with ct[k] <= x (ct[0] = 0; returns 0 when x < 0 or x is NaN)
__device__ __forceinline__ int cyc_count_le(const double *ct, int K, double x) {
    int k = (int)fmin(fmax(x * 100.0, 0.0), (double)K);
    while (k < K && __ldg(ct + k + 1) <= x) ++k;
    while (k > 0 && __ldg(ct + k) > x) --k;
    return k;
}

template <typename T> struct CycleEnvRegs {
    T pos[3], vel[3], w[3], eu[3], vw[3];
    T prev_I[3];
    T volume, prev_dist, ep_return, length, width;
    float tx, ty;
    int cycle, phase;
    unsigned cur;
};

template <typename T> struct Rec4 { T x, y, z, w; };
template <> struct __align__(16) Rec4<float> { float x, y, z, w; };
template <> struct __align__(32) Rec4<double> { double x, y, z, w; };

template <typename T>
__device__ __forceinline__ void cyc_load(const CycleArgs &a, long long e, CycleEnvRegs<T> &s) {
    const Rec4<T> *st = reinterpret_cast<const Rec4<T> *>(a.state);
    const Rec4<T> r0 = st[0 * a.n + e], r1 = st[1 * a.n + e], r2 = st[2 * a.n + e], r3 = st[3 * a.n + e],
                  r4 = st[4 * a.n + e], r5 = st[5 * a.n + e];
    s.pos[0] = r0.x; s.pos[1] = r0.y; s.pos[2] = r0.z; s.prev_dist = r0.w;
    s.vel[0] = r1.x; s.vel[1] = r1.y; s.vel[2] = r1.z; s.volume = r1.w;
    s.w[0] = r2.x; s.w[1] = r2.y; s.w[2] = r2.z; s.ep_return = r2.w;
    s.eu[0] = r3.x; s.eu[1] = r3.y; s.eu[2] = r3.z; s.length = r3.w;
    s.prev_I[0] = r4.x; s.prev_I[1] = r4.y; s.prev_I[2] = r4.z; s.width = r4.w;
    s.vw[0] = r5.x; s.vw[1] = r5.y; s.vw[2] = r5.z;
    const int4 c = a.ctrl[e];
    s.cycle = c.x & 0xFFFF; s.phase = c.x >> 16; s.cur = (unsigned)c.y;
    s.tx = __int_as_float(c.z); s.ty = __int_as_float(c.w);
}

template <typename T>
__device__ __forceinline__ void cyc_store(const CycleArgs &a, long long e, const CycleEnvRegs<T> &s) {
    Rec4<T> *st = reinterpret_cast<Rec4<T> *>(a.state);
    st[0 * a.n + e] = Rec4<T>{s.pos[0], s.pos[1], s.pos[2], s.prev_dist};
    st[1 * a.n + e] = Rec4<T>{s.vel[0], s.vel[1], s.vel[2], s.volume};
    st[2 * a.n + e] = Rec4<T>{s.w[0], s.w[1], s.w[2], s.ep_return};
    st[3 * a.n + e] = Rec4<T>{s.eu[0], s.eu[1], s.eu[2], s.length};
    st[4 * a.n + e] = Rec4<T>{s.prev_I[0], s.prev_I[1], s.prev_I[2], s.width};
    st[5 * a.n + e] = Rec4<T>{s.vw[0], s.vw[1], s.vw[2], (T)0};
    a.ctrl[e] = make_int4(s.cycle | (s.phase << 16), (int)s.cur, __float_as_int(s.tx), __float_as_int(s.ty));
}

template <typename T>
__device__ __forceinline__ void cyc_write_obs(const CycleEnvRegs<T> &s, float *o) {  // _get_observation :413-420
    float2 *p = reinterpret_cast<float2 *>(o);
    p[0] = make_float2((float)(s.pos[0] - (T)s.tx), (float)(s.pos[1] - (T)s.ty));
    p[1] = make_float2((float)s.vel[0], (float)s.vel[1]);
    p[2] = make_float2((float)s.eu[2], (float)s.w[2]);
}

// SalpRobotEnv.reset (:100-127): new target from the env's np.random row, Robot.reset (robot.py:287-312)
template <typename T>
__device__ __forceinline__ void cyc_reset_env(const CycleArgs &a, long long e, CycleEnvRegs<T> &s) {
    const long long row = (a.env_id0 + e) % a.rows;
    const double tx = a.x_min + (a.x_max - a.x_min) * cyc_next_double(a, row, s.cur);
    const double ty = a.y_min + (a.y_max - a.y_min) * cyc_next_double(a, row, s.cur);
    s.tx = (float)cyc_clip(tx, a.x_min, a.x_max);
    s.ty = (float)cyc_clip(ty, a.y_min, a.y_max);
#pragma unroll
    for (int i = 0; i < 3; ++i) { s.pos[i] = 0; s.vel[i] = 0; s.w[i] = 0; s.eu[i] = 0; s.vw[i] = 0; s.prev_I[i] = (T)a.rest_I[i]; }
    s.volume = (T)a.rest_volume;
    s.length = (T)a.L0; s.width = (T)a.W0;
    s.cycle = 0; s.phase = 3;
    const T dx = (T)0 - (T)s.tx, dy = (T)0 - (T)s.ty;
    s.prev_dist = sqrt(dx * dx + dy * dy);
    s.ep_return = 0;
}

// Everything a cycle needs that does not change from substep to substep.
template <typename T> struct CyclePlan {
    double contraction, refill;
    int n_total, kA, kB, kC;  // substeps k in 1..kA REFILL, ..kB JET, ..kC COAST, beyond REST (robot.py:360-373)
    T dir[3];                 // jet direction in the body frame
};

constexpr double kCycPi = 3.141592653589793;
constexpr double kContractRate = 0.06 / 3, kReleaseRate = 0.06 / 1.5;  // robot.py:767,776

// _rescale_action (:129-137), Nozzle.solve_angles + rotation chain (robot.py:55-83,115-130,153-178),
// set_control timing (robot.py:335-358, 761-777) and the integer phase boundaries.
template <typename T>
__device__ __forceinline__ void cyc_plan(const CycleArgs &a, double a0, double a1, double a2, CyclePlan<T> &p) {
    p.contraction = a0 * 0.06;
    const double coast = a1 * 10.0;
    const double yaw = a2 * (kCycPi / 2);
    if constexpr (sizeof(T) == 8) {
        // reference operation order: target = R_br^T @ -(cos yaw, sin yaw, 0) = (-0, -sin yaw, cos yaw)
        const double t1 = -sin(yaw), t2 = cos(yaw);
        double angle2 = acos(cyc_clip(2 * t2 - 1, -1.0, 1.0));
        double angle1;
        if (angle2 == 0) {
            angle1 = 0.0;
        } else {
            const double qa = 0.5 * (cos(angle2) - 1);
            const double qb = sqrt(2.0) * sin(angle2) / 2;
            const double val1 = cyc_clip(t1 / sqrt(qa * qa + qb * qb), -1.0, 1.0);
            angle1 = asin(val1) - atan2(qb, qa);
        }
        if (angle1 <= -kCycPi) angle1 += 2 * kCycPi;
        else if (angle1 > kCycPi) angle1 -= 2 * kCycPi;
        const double cg = cos(kCycPi / 4), sg = sin(kCycPi / 4);
        const double c1 = cos(angle1), s1 = sin(angle1), c2 = cos(angle2), s2 = sin(angle2);
        const double n00 = cg * c2, n02 = -sg, n10 = s2, n12 = 0.0, n20 = sg * c2, n22 = cg;
        const double b00 = -n20, b02 = -n22;
        const double b10 = s1 * n00 + c1 * n10, b12 = s1 * n02 + c1 * n12;
        const double b20 = c1 * n00 + -s1 * n10, b22 = c1 * n02 + -s1 * n12;
        p.dir[0] = (T)(b00 * cg + b02 * sg);
        p.dir[1] = (T)(b10 * cg + b12 * sg);
        p.dir[2] = (T)(b20 * cg + b22 * sg);
    } else {
        // closed form of the same chain: the IK aims the jet at -(cos yaw, sin yaw, 0) in the body frame
        float sy, cy;
        sincosf((float)yaw, &sy, &cy);
        p.dir[0] = (T)(-cy); p.dir[1] = (T)(-sy); p.dir[2] = (T)0;
    }
    p.refill = p.contraction / kContractRate;
    const double jet = p.contraction / kReleaseRate;
    const double rj = p.refill + jet;
    const double total = rj + coast;
    const int K = a.max_substeps;
    const int c = cyc_count_le(a.ct, K, total);
    int n_total = (__ldg(a.ct + c) < total) ? c + 1 : c;   // while cycle_time < total (robot.py:435)
    p.n_total = n_total > K ? K : n_total;
    p.kA = cyc_count_le(a.ct, K, p.refill);
    p.kB = cyc_count_le(a.ct, K, rj);
    p.kC = cyc_count_le(a.ct, K, total);
    if (p.kB < p.kA) p.kB = p.kA;
    if (p.kC < p.kB) p.kC = p.kB;
}

// Coefficients of the dynamics for one body shape (everything in update_properties, robot.py:375-385, plus
// the shape-only factors of the force / torque laws :494-522, :534-575, :627-674).  Recomputed every substep
// while the body contracts or releases, constant through coast / rest.
template <typename T> struct CycleCoef;
template <> struct CycleCoef<double> {
    double mass, inv_m, kd, kt0, rx, cm, jet_speed;
    double I[3], inv_I[3], dI_rate[3];
};

// double build: the reference's expressions in the reference's operation order
__device__ __forceinline__ void cyc_shape(const CycleArgs &a, double length, double width, bool jet_phase,
                                          double &prev_vol, double prev_I[3], CycleCoef<double> &c) {
    using T = double;
    const double hw = width / 2, hl = length / 2;
    const double area = kCycPi * hl * hw;
    const double volume = 4.0 / 3.0 * kCycPi * hl * (hw * hw);
    const double water_mass = a.density * volume;
    const double mass = a.dry_mass + water_mass + a.noz_mass;
    double nr = (length / width - a.min_aspect) / (a.init_aspect - a.min_aspect);
    nr = cyc_clip(nr, 0.0, 1.0);
    const double cd = 1.0 - nr * (1.0 - 0.4);
    const double rx = -a.noz_l12 + (-length / 2);        // _get_jet_moment_arm, y = z = 0 (robot.py:567-575)
    const double nrm = fabs(rx);                         // np.linalg.norm((rx, 0, 0)) = sqrt(rx^2), exactly |rx|
    const double In = a.noz_mass * (nrm * nrm);
    double I[3];
    I[0] = 0.2 * mass * (hw * hw + hw * hw) + In * 0.0;  // get_inertia_matrix (robot.py:534-551), diagonal
    I[1] = 0.2 * mass * (hl * hl + hw * hw) + In;
    I[2] = 0.2 * mass * (hw * hw + hl * hl) + In;
    c.mass = (T)mass;
    c.inv_m = (T)(1.0 / mass);                           // np.linalg.inv of the diagonal mass matrix
    c.kd = (T)(-0.5 * a.density * area * cd);
    c.kt0 = (T)(-a.density * cd * hw * pow(hl, 4.0));   // (self.length / 2) ** 4
    c.rx = (T)rx;
    c.cm = 0; c.jet_speed = 0;
    if (jet_phase) {                                     // _get_jet_velocity / _get_jet_force (robot.py:586-614)
        const double volume_rate = -(volume - prev_vol) / 0.01;
        c.jet_speed = (T)(volume_rate / a.noz_area);
        const double mass_rate = (water_mass - prev_vol * a.density) / 0.01;
        c.cm = (T)(0.1 * mass_rate);
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        c.dI_rate[i] = (T)((I[i] - prev_I[i]) / 0.01);   // get_inertia_matrix_rate (robot.py:553-565)
        prev_I[i] = I[i];
        c.I[i] = (T)I[i];
        c.inv_I[i] = (T)(1.0 / I[i]);
    }
    prev_vol = volume;
}

// float build: every shape-dependent factor is folded into the coefficients the substep multiplies by (forces
// already divided by the mass, torques by the inertia; I_yy == I_zz, robot.py:546-547, so the gyroscopic
// term needs I_yy - I_xx only).  double is kept only where nearly equal numbers are subtracted (water volume
// and inertia of consecutive substeps); no double divisions, reciprocals on the SFU.
template <> struct CycleCoef<float> {
    float kda;          // drag: acc_i = kda * (|v| + 1) * v_i
    float fja[3];       // jet force / mass
    float kta0, kta1;   // rotational drag: kt0 / I_xx, kt0 / I_yy   (times |omega| * omega_i)
    float dIa0, dIa1;   // deformation: dI/dt / I
    float tja1, tja2;   // jet torque / I_yy
    float g1;           // (I_yy - I_xx) / I_yy
    float ia;           // 0.1 / I_zz  (asymmetry torque, robot.py:513)
};

__device__ __forceinline__ float cyc_rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float cyc_rsqrt_approx(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

// float build.  Shape state is the contraction progress s (length = L0 - s, width = W0 + s).  The rates the
// dynamics need are differences of the volume / inertia of consecutive substeps; instead of subtracting two
// nearly equal numbers they are expanded exactly in the step delta = s - s_prev, which keeps everything in
// float (no double arithmetic, no conversions):
//   g(s) = L W^2                 V = pi/6 g          g(s) - g(s - delta) = delta (2 L1 W1 - W1^2) + delta^2 (L1 - 2 W1) - delta^3
//   q(s) = (L^2 + W^2) / 4                           q(s) - q(s - delta) = delta ((W1 + W2) - (L1 + L2)) / 4
//   rx(s) = -(l1 + l2) - L / 2                       rx^2(s) - rx^2(s - delta) = delta / 2 (rx1 + rx2)
// with index 1 = previous shape, 2 = current shape.
__device__ __forceinline__ void cyc_shape_f(const CycleArgs &a, float s2, float delta, bool jet_phase, const float (&dir)[3],
                                            CycleCoef<float> &c) {
    const float L2 = a.f_L0 - s2, W2 = a.f_W0 + s2, L1 = L2 + delta, W1 = W2 - delta;
    const float kv = 0.52359878f * a.f_density;                       // rho pi / 6
    const float g2 = L2 * W2 * W2;
    const float dg = delta * (fmaf(2.0f * L1, W1, -W1 * W1) + delta * ((L1 - 2.0f * W1) - delta));
    const float m2 = fmaf(kv, g2, a.f_m_fixed), dm = kv * dg, m1 = m2 - dm;
    const float q2 = 0.25f * fmaf(L2, L2, W2 * W2);
    const float dq = 0.25f * delta * ((W1 + W2) - (L1 + L2));
    const float rx2 = -a.f_l12 - 0.5f * L2, rx1 = rx2 - 0.5f * delta;
    const float I1 = fmaf(0.2f * m2, q2, a.f_noz_mass * rx2 * rx2);
    const float dI1 = fmaf(0.2f, fmaf(dm, q2, m1 * dq), a.f_noz_mass * (0.5f * delta) * (rx1 + rx2));
    const float I0 = 0.1f * m2 * (W2 * W2);
    const float dI0 = 0.1f * fmaf(dm, W2 * W2, m1 * delta * (W1 + W2));
    float nr = (L2 * cyc_rcp_approx(W2) - a.f_min_aspect) * a.f_inv_aspect_span;
    nr = fminf(fmaxf(nr, 0.0f), 1.0f);
    const float cd = fmaf(nr, -0.6f, 1.0f);
    const float inv_m = cyc_rcp_approx(m2), inv_I0 = cyc_rcp_approx(I0), inv_I1 = cyc_rcp_approx(I1);
    const float rcd = a.f_density * cd, hl = 0.5f * L2, hl2 = hl * hl;
    c.kda = inv_m * (-0.39269908f * rcd * (L2 * W2));                  // -0.5 rho (pi L W / 4) cd / m
    const float kt0 = -rcd * (0.5f * W2) * (hl2 * hl2);
    c.kta0 = inv_I0 * kt0; c.kta1 = inv_I1 * kt0;
    c.g1 = inv_I1 * (I1 - I0);
    c.ia = 0.1f * inv_I1;
    c.dIa0 = inv_I0 * (dI0 * 100.0f);
    c.dIa1 = inv_I1 * (dI1 * 100.0f);
    float f = 0.0f;
    if (jet_phase) {                                                    // robot.py:586-614
        const float dv = 0.52359878f * dg * 100.0f;                    // dV/dt
        f = (0.1f * (a.f_density * dv)) * (-dv * a.f_inv_noz_area);   // C_discharge * mass_rate * jet_speed
    }
    const float fa = inv_m * f, ta = inv_I1 * (rx2 * f);
    c.fja[0] = fa * dir[0]; c.fja[1] = fa * dir[1]; c.fja[2] = fa * dir[2];
    c.tja1 = -(ta * dir[2]); c.tja2 = ta * dir[1];                     // cross((rx, 0, 0), F_jet) / I
}

struct CycleStatsAcc {
    double ret = 0.0, cycles = 0.0;
    int episodes = 0, reached = 0, truncs = 0, nonfinite = 0, steps = 0;
    long long substeps = 0;
};

// sin/cos, sqrt and reciprocal per build.  double: the libm calls the reference makes.  float: SFU
// approximations (MUFU.SIN/COS after a two-constant reduction to [-pi, pi], MUFU.RSQ, MUFU.RCP; absolute
// error <= 5e-7, below the resolution the float state carries).
template <typename T> __device__ __forceinline__ void cyc_sincos(T x, T &s, T &c);
template <> __device__ __forceinline__ void cyc_sincos<double>(double x, double &s, double &c) { s = sin(x); c = cos(x); }
template <> __device__ __forceinline__ void cyc_sincos<float>(float x, float &s, float &c) {
    // MUFU.SIN/COS of x / (2 pi): the phase error of that float product (6e-8 * |x|) is below the spacing of
    // float angles of that magnitude, so an explicit range reduction would buy nothing
    __sincosf(x, &s, &c);
}
__device__ __forceinline__ double cyc_sqrt(double x) { return sqrt(x); }
__device__ __forceinline__ float cyc_sqrt(float x) { return x * cyc_rsqrt_approx(fmaxf(x, 1e-30f)); }
__device__ __forceinline__ double cyc_rcp(double x) { return 1.0 / x; }
__device__ __forceinline__ float cyc_rcp(float x) { return cyc_rcp_approx(x); }

// One SalpRobotEnv.step for env e (salp_robot_env.py:139-201): plan the cycle, integrate it, reward,
// termination, observation, auto-reset.
template <typename T>
__device__ __forceinline__ void cyc_run_env(const CycleArgs &a, long long e, CycleStatsAcc &acc) {
    CycleEnvRegs<T> s;
    cyc_load<T>(a, e, s);
    const T *act = reinterpret_cast<const T *>(a.actions) + 3 * e;
    CyclePlan<T> p;
    cyc_plan<T>(a, (double)act[0], (double)act[1], (double)act[2], p);
    s.cycle += 1;                                   // set_control (robot.py:349-351)

    double prev_vol = (double)s.volume, prev_I[3] = {(double)s.prev_I[0], (double)s.prev_I[1], (double)s.prev_I[2]};
    double length = (double)s.length, width = (double)s.width;
    double s_prev = a.L0 - length;                   // float build: contraction progress of the previous substep
    int phase = s.phase;
    CycleCoef<T> c;
    bool steady = false;
    const T dt = (T)0.01;
    T v0 = s.vel[0], v1 = s.vel[1], v2 = s.vel[2], w0 = s.w[0], w1 = s.w[1], w2 = s.w[2];
    T e0 = s.eu[0], e1 = s.eu[1], e2 = s.eu[2], p0 = s.pos[0], p1 = s.pos[1], p2 = s.pos[2];
    T vw0 = s.vw[0], vw1 = s.vw[1], vw2 = s.vw[2];
    T sphi, cphi, sth, cth;
    cyc_sincos<T>(e0, sphi, cphi);
    cyc_sincos<T>(e1, sth, cth);

    bool rest_shape = false;                         // the coefficients already describe the rest shape
    for (int k = 1; k <= p.n_total; ++k) {          // step_through_cycle (robot.py:435-436), Robot.step (:387-395)
        if (!steady) {                              // update_state (:360-373), update_properties (:375-385)
            phase = k <= p.kA ? 0 : (k <= p.kB ? 1 : (k <= p.kC ? 2 : 3));
            if (rest_shape) {                       // second substep at rest shape: no deformation torque any more
                steady = true;
                if constexpr (sizeof(T) == 8) { c.dI_rate[0] = 0; c.dI_rate[1] = 0; c.dI_rate[2] = 0; }
                else { c.dIa0 = 0; c.dIa1 = 0; }
            } else {
                if constexpr (sizeof(T) == 8) {
                    if (phase <= 1) {
                        const double ctk = __ldg(a.ct + k);
                        if (phase == 0) {           // get_current_length / width (:702-730)
                            length = a.L0 - ctk * kContractRate;
                            width = a.W0 + ctk * kContractRate;
                        } else {
                            length = a.L0 - p.contraction + (ctk - p.refill) * kReleaseRate;
                            width = a.W0 + p.contraction - (ctk - p.refill) * kReleaseRate;
                        }
                    } else {
                        length = a.L0;
                        width = a.W0;
                        rest_shape = true;
                    }
                    cyc_shape(a, length, width, phase == 1, prev_vol, prev_I, c);
                } else {
                    double s2 = 0.0;
                    if (phase <= 1) {
                        const double ctk = __ldg(a.ct + k);
                        s2 = phase == 0 ? ctk * kContractRate : p.contraction - (ctk - p.refill) * kReleaseRate;
                    } else {
                        rest_shape = true;
                    }
                    cyc_shape_f(a, (float)s2, (float)(s2 - s_prev), phase == 1, p.dir, c);
                    s_prev = s2;
                }
            }
        }
        if constexpr (sizeof(T) == 8) {
            // _newton_equations (:494-505)
            const T cx = w1 * v2 - w2 * v1, cy = w2 * v0 - w0 * v2, cz = w0 * v1 - w1 * v0;
            const T vnorm = cyc_sqrt(v0 * v0 + v1 * v1 + v2 * v2);
            const T kdn = c.kd * vnorm;
            const T fj0 = c.cm * (p.dir[0] * c.jet_speed), fj1 = c.cm * (p.dir[1] * c.jet_speed), fj2 = c.cm * (p.dir[2] * c.jet_speed);
            const T a0 = c.inv_m * (fj0 + (kdn * v0 + c.kd * v0) + c.mass * cx);
            const T a1 = c.inv_m * (fj1 + (kdn * v1 + c.kd * v1) + c.mass * cy);
            const T a2 = c.inv_m * (fj2 + (kdn * v2 + c.kd * v2) + c.mass * cz);
            // _euler_equations (:507-522)
            const T Iw0 = c.I[0] * w0, Iw1 = c.I[1] * w1, Iw2 = c.I[2] * w2;
            const T tc0 = -(w1 * Iw2 - w2 * Iw1), tc1 = -(w2 * Iw0 - w0 * Iw2), tc2 = -(w0 * Iw1 - w1 * Iw0);
            const T wnorm = cyc_sqrt(w0 * w0 + w1 * w1 + w2 * w2);
            const T kt = c.kt0 * wnorm;
            const T tj1 = -(c.rx * fj2), tj2 = c.rx * fj1;   // cross((rx, 0, 0), F_jet)
            const T aa0 = c.inv_I[0] * ((T)0 + kt * w0 + tc0 + (T)0 - c.dI_rate[0] * w0);
            const T aa1 = c.inv_I[1] * (tj1 + kt * w1 + tc1 + (T)0 - c.dI_rate[1] * w1);
            const T aa2 = c.inv_I[2] * (tj2 + kt * w2 + tc2 + (T)0.1 * vnorm - c.dI_rate[2] * w2);
            // _update_motion_states (:524-532)
            v0 += a0 * dt; v1 += a1 * dt; v2 += a2 * dt;
            w0 += aa0 * dt; w1 += aa1 * dt; w2 += aa2 * dt;
            T er0, er1, er2;                             // _to_euler_angle_rate (:447-459) on the pre-update angles
            if constexpr (sizeof(T) == 8) {
                const T tth = tan(e1);
                er0 = 1 * w0 + sphi * tth * w1 + cphi * tth * w2;
                er1 = 0 * w0 + cphi * w1 + -sphi * w2;
                er2 = 0 * w0 + sphi / cth * w1 + cphi / cth * w2;
            } else {
                const T icth = cyc_rcp(cth), tth = sth * icth;
                er0 = w0 + tth * (sphi * w1 + cphi * w2);
                er1 = cphi * w1 - sphi * w2;
                er2 = icth * (sphi * w1 + cphi * w2);
            }
            e0 += er0 * dt; e1 += er1 * dt; e2 += er2 * dt;
            T ss, cs;                                    // _to_world_frame (:461-486) on the updated angles
            cyc_sincos<T>(e0, sphi, cphi);
            cyc_sincos<T>(e1, sth, cth);
            cyc_sincos<T>(e2, ss, cs);
            const T m00 = cs * cth, m01 = -ss, m02 = cs * sth;
            const T m10 = ss * cth, m11 = cs, m12 = ss * sth;
            const T r01 = m01 * cphi + m02 * sphi, r02 = m01 * -sphi + m02 * cphi;
            const T r11 = m11 * cphi + m12 * sphi, r12 = m11 * -sphi + m12 * cphi;
            const T r21 = (T)0 * cphi + cth * sphi, r22 = (T)0 * -sphi + cth * cphi;
            vw0 = m00 * v0 + r01 * v1 + r02 * v2;
            vw1 = m10 * v0 + r11 * v1 + r12 * v2;
            vw2 = -sth * v0 + r21 * v1 + r22 * v2;
            p0 += vw0 * dt; p1 += vw1 * dt; p2 += vw2 * dt;
        } else {
            // the same equations with the shape-only factors pre-folded (CycleCoef<float>)
            const T cx = w1 * v2 - w2 * v1, cy = w2 * v0 - w0 * v2, cz = w0 * v1 - w1 * v0;
            const T vnorm = cyc_sqrt(v0 * v0 + v1 * v1 + v2 * v2);
            const T kdv = fmaf(c.kda, vnorm, c.kda);
            const T a0 = fmaf(kdv, v0, c.fja[0]) + cx, a1 = fmaf(kdv, v1, c.fja[1]) + cy, a2 = fmaf(kdv, v2, c.fja[2]) + cz;
            const T wnorm = cyc_sqrt(w0 * w0 + w1 * w1 + w2 * w2);
            const T kw0 = fmaf(c.kta0, wnorm, -c.dIa0), kw1 = fmaf(c.kta1, wnorm, -c.dIa1);
            const T aa0 = kw0 * w0;                                       // I_yy == I_zz: no gyroscopic roll torque
            const T aa1 = fmaf(c.g1, w0 * w2, fmaf(kw1, w1, c.tja1));
            const T aa2 = fmaf(c.ia, vnorm, fmaf(-c.g1, w0 * w1, fmaf(kw1, w2, c.tja2)));
            v0 = fmaf(a0, dt, v0); v1 = fmaf(a1, dt, v1); v2 = fmaf(a2, dt, v2);
            w0 = fmaf(aa0, dt, w0); w1 = fmaf(aa1, dt, w1); w2 = fmaf(aa2, dt, w2);
            const T icth = cyc_rcp(cth), tth = sth * icth;
            const T t12 = fmaf(sphi, w1, cphi * w2);
            e0 = fmaf(fmaf(tth, t12, w0), dt, e0);
            e1 = fmaf(fmaf(cphi, w1, -sphi * w2), dt, e1);
            e2 = fmaf(icth * t12, dt, e2);
            T ss, cs;
            cyc_sincos<T>(e0, sphi, cphi);
            cyc_sincos<T>(e1, sth, cth);
            cyc_sincos<T>(e2, ss, cs);
            const T m02 = cs * sth, m12 = ss * sth;
            vw0 = fmaf(cs * cth, v0, fmaf(fmaf(m02, sphi, -ss * cphi), v1, fmaf(m02, cphi, ss * sphi) * v2));
            vw1 = fmaf(ss * cth, v0, fmaf(fmaf(m12, sphi, cs * cphi), v1, fmaf(m12, cphi, -cs * sphi) * v2));
            vw2 = fmaf(-sth, v0, cth * fmaf(sphi, v1, cphi * v2));
            p0 = fmaf(vw0, dt, p0); p1 = fmaf(vw1, dt, p1); p2 = fmaf(vw2, dt, p2);
        }
    }
    if (p.n_total > 0) phase = p.n_total <= p.kA ? 0 : (p.n_total <= p.kB ? 1 : (p.n_total <= p.kC ? 2 : 3));
    s.vel[0] = v0; s.vel[1] = v1; s.vel[2] = v2; s.w[0] = w0; s.w[1] = w1; s.w[2] = w2;
    s.eu[0] = e0; s.eu[1] = e1; s.eu[2] = e2; s.pos[0] = p0; s.pos[1] = p1; s.pos[2] = p2;
    s.vw[0] = vw0; s.vw[1] = vw1; s.vw[2] = vw2;
    s.phase = phase;
    if constexpr (sizeof(T) == 4) {                  // the stored shape state, from the last substep's progress
        if (p.n_total > 0) {
            length = a.L0 - s_prev; width = a.W0 + s_prev;
            const double hw = width * 0.5, hl = length * 0.5, rx = -a.noz_l12 - hl;
            prev_vol = (4.0 / 3.0 * kCycPi) * hl * (hw * hw);
            const double m5 = 0.2 * (a.dry_mass + a.density * prev_vol + a.noz_mass), In = a.noz_mass * (rx * rx);
            prev_I[0] = m5 * (hw * hw + hw * hw); prev_I[1] = m5 * (hl * hl + hw * hw) + In; prev_I[2] = prev_I[1];
        }
    }
    s.volume = (T)prev_vol; s.prev_I[0] = (T)prev_I[0]; s.prev_I[1] = (T)prev_I[1]; s.prev_I[2] = (T)prev_I[2];
    s.length = (T)length; s.width = (T)width;

    // _calculate_reward (:203-242); self.action is never assigned, so the smoothness term is -0.1 * 0^2
    const T dx = p0 - (T)s.tx, dy = p1 - (T)s.ty;
    const T dist = sqrt(dx * dx + dy * dy);
    const T r_track = (-dist + s.prev_dist) * (T)100;
    s.prev_dist = dist;
    const T en = dist + (T)1e-6;
    const T ex = -(dx / en), ey = -(dy / en);
    const T hn = sqrt(vw0 * vw0 + vw1 * vw1) + (T)1e-6;
    const T r_heading = (vw0 / hn) * ex + (vw1 / hn) * ey;
    T rew = ((T)1.0 * r_track) + ((T)0.5 * r_heading);
    bool done = false, trunc = false;                // termination (:174-188)
    if (dist < (T)0.01) { done = true; rew += (T)10.0; }
    else if (dist > (T)5.0) { trunc = true; rew -= (T)5.0; }
    if (s.cycle >= a.max_cycles) trunc = true;
    s.ep_return += rew;

    reinterpret_cast<T *>(a.reward)[e] = rew;
    a.terminated[e] = done ? 1 : 0;
    a.truncated[e] = trunc ? 1 : 0;
    if (a.n_sub) a.n_sub[e] = p.n_total;
    acc.steps += 1;
    acc.substeps += p.n_total;
    const bool finite = (p0 == p0) && (p1 == p1) && (fabs((double)p0) <= 1e300) && (fabs((double)p1) <= 1e300) && (rew == rew);
    acc.nonfinite += finite ? 0 : 1;
    if (done || trunc) {
        acc.episodes += 1; acc.ret += (double)s.ep_return; acc.cycles += (double)s.cycle;
        acc.reached += done ? 1 : 0; acc.truncs += trunc ? 1 : 0;
        if (a.final_obs) cyc_write_obs<T>(s, a.final_obs + 6 * e);
        if (a.autoreset) cyc_reset_env<T>(a, e, s);
    }
    cyc_write_obs<T>(s, a.obs + 6 * e);
    cyc_store<T>(a, e, s);
}

__device__ __forceinline__ double cyc_warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Episode statistics: warp shuffles, then one atomic per warp and counter.
__device__ __forceinline__ void cyc_flush_stats(const CycleArgs &a, const CycleStatsAcc &acc) {
    if (!a.stats) return;
    const int ep = __reduce_add_sync(0xffffffffu, acc.episodes);
    const int steps = __reduce_add_sync(0xffffffffu, acc.steps);
    const int reached = __reduce_add_sync(0xffffffffu, acc.reached);
    const int truncs = __reduce_add_sync(0xffffffffu, acc.truncs);
    const int nonf = __reduce_add_sync(0xffffffffu, acc.nonfinite);
    const double sub = cyc_warp_sum((double)acc.substeps);
    double ret = 0.0, cyc = 0.0;
    if (ep) { ret = cyc_warp_sum(acc.ret); cyc = cyc_warp_sum(acc.cycles); }  // ep is warp-uniform
    if ((threadIdx.x & 31) == 0) {
        if (ep) { atomicAdd(a.stats + 0, (double)ep); atomicAdd(a.stats + 1, ret); atomicAdd(a.stats + 2, cyc); }
        if (reached) atomicAdd(a.stats + 3, (double)reached);
        if (truncs) atomicAdd(a.stats + 4, (double)truncs);
        if (nonf) atomicAdd(a.stats + 5, (double)nonf);
        atomicAdd(a.stats + 6, (double)steps);
        atomicAdd(a.stats + 7, sub);
    }
}

constexpr int kCycleSlice = 2048;  // envs sorted at a time by one CTA

// Number of integrator substeps of env e's next cycle and the length of its breathing part, from its action
// alone (same arithmetic as cyc_plan).
template <typename T>
__device__ __forceinline__ void cyc_lengths(const CycleArgs &a, long long e, int &n_total, int &kB) {
    const T *act = reinterpret_cast<const T *>(a.actions) + 3 * e;
    const double contraction = (double)act[0] * 0.06, coast = (double)act[1] * 10.0;
    const double rj = contraction / kContractRate + contraction / kReleaseRate;
    const double total = rj + coast;
    const int K = a.max_substeps;
    const int c = cyc_count_le(a.ct, K, total);
    const int n = (__ldg(a.ct + c) < total) ? c + 1 : c;
    n_total = n > K ? K : n;
    kB = cyc_count_le(a.ct, K, rj);
}

// Work distribution.  Cycles differ in length by the action (0 .. 1450 substeps), so lanes of a warp that took
// envs in index order would idle half of the time.  Each CTA owns a contiguous range of envs (one CTA per SM, all
// resident: one wave, equal expected work), sorts it in slices of kCycleSlice by substep count
// (bitonic sort of (count | local index) keys in shared memory), and its warps pull 32-env chunks of the sorted
// slice, longest first, from a shared counter: the lanes of a warp run cycles of nearly equal length, and a warp
// that drew short cycles simply draws more of them.
template <typename T, int BLOCK, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) salp_cycle_step_kernel(const __grid_constant__ CycleArgs a) {
    __shared__ unsigned s_key[kCycleSlice];
    __shared__ int s_next, s_valid;
    const int tid = threadIdx.x, lane = tid & 31;
    const long long per = (a.n + gridDim.x - 1) / gridDim.x;
    const long long begin = per * blockIdx.x;
    const long long end = begin + per < a.n ? begin + per : a.n;
    CycleStatsAcc acc;
    for (long long base = begin; base < end; base += kCycleSlice) {
        const int cnt = (int)(end - base < kCycleSlice ? end - base : kCycleSlice);
        if (tid == 0) { s_next = 0; s_valid = cnt; }
        // lanes that still breathe make the whole warp run the shape block: with enough chunks in the slice, group
        // envs of similar breathing length first (measured at 55 chunks: 1 group 1.122 ms, 2 1.015, 3 1.006, 4 1.018, 8 1.059)
        const int n_buckets = a.sort_buckets > 0 ? a.sort_buckets : max(1, min(3, ((cnt + 31) >> 5) / 16));
        for (int i = tid; i < kCycleSlice; i += BLOCK) {
            unsigned key = 0xFFFFFFFFu;
            if (i < cnt) {
                int n_total, kB;
                cyc_lengths<T>(a, base + i, n_total, kB);
                const int bucket = n_buckets > 1 ? min(n_buckets - 1, kB * n_buckets / 451) : 0;
                key = ((unsigned)bucket << 24) | ((unsigned)n_total << 11) | (unsigned)i;
            }
            s_key[i] = key;
        }
        __syncthreads();
        for (int size = 2; size <= kCycleSlice; size <<= 1) {          // bitonic sort, ascending
            for (int stride = size >> 1; stride > 0; stride >>= 1) {
                for (int t = tid; t < kCycleSlice / 2; t += BLOCK) {
                    const int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                    const unsigned x = s_key[lo], y = s_key[hi];
                    const bool up = (lo & size) == 0;
                    if ((x > y) == up) { s_key[lo] = y; s_key[hi] = x; }
                }
                __syncthreads();
            }
        }
        const int n_chunks = (cnt + 31) >> 5;
        for (;;) {
            int c = 0;
            if (lane == 0) c = atomicAdd(&s_next, 1);
            c = __shfl_sync(0xffffffffu, c, 0);
            if (c >= n_chunks) break;
            const int rank = cnt - 1 - (c * 32 + lane);                // longest cycles first
            if (rank >= 0) cyc_run_env<T>(a, base + (long long)(s_key[rank] & 2047u), acc);
            __syncwarp();
        }
        __syncthreads();                                                 // s_key is rewritten by the next slice
    }
    cyc_flush_stats(a, acc);
}

template <typename T>
__global__ void __launch_bounds__(256) salp_cycle_reset_kernel(const __grid_constant__ CycleArgs a) {
    const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
    if (e >= a.n) return;
    if (a.mask && !a.mask[e]) return;
    CycleEnvRegs<T> s;
    s.cur = a.rewind_rng ? 0u : (unsigned)a.ctrl[e].y;
    cyc_reset_env<T>(a, e, s);
    cyc_write_obs<T>(s, a.obs + 6 * e);
    cyc_store<T>(a, e, s);
}

cudaError_t launch_cycle_step_f32(const CycleArgs &a, cudaStream_t s);
cudaError_t launch_cycle_step_f64(const CycleArgs &a, cudaStream_t s);
cudaError_t launch_cycle_reset_f32(const CycleArgs &a, cudaStream_t s);
cudaError_t launch_cycle_reset_f64(const CycleArgs &a, cudaStream_t s);

}  // namespace salp
