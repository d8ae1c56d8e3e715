/* salp_cycle_oracle.c -- CPU restatement of the reference's model (B) path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under underwater_swimmer_rl_b200/ may import, link or execute this
 * file; it exists to check the CUDA path (tests/, __graft_entry__.smoke()) and to be timed as the CPU
 * baseline of the robot_cycle workload (bench.py cpu_baseline / --impl reference).
 *
 * What it restates (reference tree, Python/numpy):
 *   src/salp/environments/robot.py           Nozzle :5-178, Robot :181-777
 *   src/salp/environments/salp_robot_env.py  SalpRobotEnv.reset :100-127, _rescale_action :129-137,
 *                                            step :139-201, _calculate_reward :203-242,
 *                                            generate_target_point("random") :266-328,
 *                                            _get_observation :400-420
 * One env step = one breathing cycle: the action (contraction, coast time, nozzle yaw) is rescaled, the
 * 3-link nozzle inverse kinematics is solved, and Robot.step_through_cycle integrates 3-D Newton-Euler
 * dynamics with dt = 0.01 for round-up((refill + jet + coast) / dt) substeps (0 .. 1450).
 *
 * PARITY STATUS: pinned against the unmodified Python reference run in the build container
 * (oracle/validate_cycle_against_reference.py; committed fixtures tests/golden/cycle_*.npz).  The reference
 * computes its 3x3 products and norms through BLAS/LAPACK (np.linalg.inv, @, np.linalg.norm), whose
 * summation order and FMA use are not part of any contract, so the pin is: DISCRETE state (substeps per
 * cycle, phase, cycle count, terminated / truncated) bit-exact, continuous state within 1e-6 abs/rel.
 * The floor is the reference's own: its nozzle IK takes arcsin of a quantity that is analytically
 * -sign(yaw) for every yaw (a^2 + b^2 == sin^2 yaw), so angle1 carries +-sqrt(2 ulp) ~ 2e-8 of
 * libm-dependent rounding noise (numpy's SIMD sin/cos differ from glibc's in the last bit); whenever
 * the IK roundings coincide, whole trajectories agree to 1e-14.
 *
 * Arithmetic notes (each checked against the reference):
 *   - np.linalg.inv of the diagonal mass / inertia matrices is diag(1/d): the product with the force
 *     vector is (1/d)*f (a reciprocal and a multiply, not a division).
 *   - R_br^T @ v, R_mb @ [0,0,l2] and the zero entries of every rotation are exact, so they are folded.
 *   - `self.action` is never assigned by step() (salp_robot_env.py:48,200,223), so nozzle_yaw in the
 *     reward is always 0 and r_smooth is -0.1 * |0 - 0|^2.
 *   - the target point is float32 (astype, :328); distances are computed in double from it.
 *   - np.random.uniform(lo, hi) = lo + (hi - lo) * random_sample(), random_sample = 53-bit MT19937 double.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

typedef struct {
    double dry_mass, init_length, init_width, max_contraction, density; /* Robot.__init__ robot.py:196-222 */
    double noz_l1, noz_l2, noz_l3, noz_area, noz_mass;                  /* Nozzle.__init__ robot.py:19-25  */
    double width, height;                                               /* SalpRobotEnv.__init__ :32-39    */
    int32_t max_cycles;    /* 500, salp_robot_env.py:187 */
    int32_t max_substeps;  /* safety cap on the cycle loop (actions outside the Box); 0 = none */
} CycleParams;

typedef struct {
    double pos[3], vel[3], w[3], euler[3], prev_I[3], vel_world[3];
    double volume, prev_dist, ep_return, time;
    double length, width_b;     /* body length / width after the last substep (info only) */
    float target[2];
    int32_t cycle, n_sub, phase, pad_;
    uint64_t cur;               /* cursor into this env's np.random word row */
} CycleEnv;

typedef struct {
    const uint32_t *words; /* [rows][len] MT19937 outputs of np.random.seed(seed0 + row) */
    int64_t rows, len;
} CycleTables;

int64_t salp_cycle_oracle_env_size(void) { return (int64_t)sizeof(CycleEnv); }

#define DT 0.01
static const double kPi = 3.141592653589793;

static inline double clipd(double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); }

static inline double next_double(const CycleTables *t, int64_t row, uint64_t *cur) {
    const uint32_t *r = t->words + row * t->len;
    uint32_t a = r[(*cur) % (uint64_t)t->len] >> 5;
    uint32_t b = r[(*cur + 1) % (uint64_t)t->len] >> 6;
    *cur += 2;
    return (a * 67108864.0 + b) / 9007199254740992.0;
}

/* ---- geometry helpers (robot.py:732-759, :616-649) ---- */
static inline double vol_of(double length, double width) {
    double hw = width / 2;
    return 4.0 / 3.0 * kPi * (length / 2) * (hw * hw);
}
static inline double area_of(double length, double width) { return kPi * (length / 2) * (width / 2); }

static double drag_coefficient(const CycleParams *p, double length, double width) {
    double aspect = length / width;
    double init_aspect = p->init_length / p->init_width;
    double cl = p->init_length - p->max_contraction;
    double cw = p->init_length - cl + p->init_width;
    double min_aspect = cl / cw;
    double nr = (aspect - min_aspect) / (init_aspect - min_aspect);
    nr = clipd(nr, 0.0, 1.0);
    return 1.0 - nr * (1.0 - 0.4);
}

/* diag of get_inertia_matrix (robot.py:534-551) for total mass m */
static void inertia_diag(const CycleParams *p, double m, double length, double width, double I[3]) {
    double rx = -(p->noz_l1 + p->noz_l2) + (-length / 2); /* _get_jet_moment_arm :567-575, y = z = 0 */
    double nr = sqrt(rx * rx + 0.0 + 0.0);
    double In = p->noz_mass * (nr * nr);
    double hw = width / 2, hl = length / 2;
    I[0] = 0.2 * m * (hw * hw + hw * hw) + In * 0.0;
    I[1] = 0.2 * m * (hl * hl + hw * hw) + In;
    I[2] = 0.2 * m * (hw * hw + hl * hl) + In;
}

static void robot_reset(const CycleParams *p, CycleEnv *e) { /* Robot.reset robot.py:287-312 */
    memset(e->pos, 0, sizeof e->pos);
    memset(e->vel, 0, sizeof e->vel);
    memset(e->w, 0, sizeof e->w);
    memset(e->euler, 0, sizeof e->euler);
    memset(e->vel_world, 0, sizeof e->vel_world);
    e->time = 0.0;
    e->cycle = 0;
    e->phase = 3;
    e->n_sub = 0;
    e->length = p->init_length;
    e->width_b = p->init_width;
    e->volume = vol_of(p->init_length, p->init_width);
    double m = p->dry_mass + p->density * e->volume + p->noz_mass;
    inertia_diag(p, m, p->init_length, p->init_width, e->prev_I);
}

static void write_obs(const CycleEnv *e, float *obs) { /* _get_observation :400-420 */
    obs[0] = (float)(e->pos[0] - (double)e->target[0]);
    obs[1] = (float)(e->pos[1] - (double)e->target[1]);
    obs[2] = (float)e->vel[0];
    obs[3] = (float)e->vel[1];
    obs[4] = (float)e->euler[2];
    obs[5] = (float)e->w[2];
}

static double dist_to_target(const CycleEnv *e) {
    double dx = e->pos[0] - (double)e->target[0], dy = e->pos[1] - (double)e->target[1];
    return sqrt(dx * dx + dy * dy);
}

static void env_reset(const CycleParams *p, CycleEnv *e, const CycleTables *t, int64_t row, float *obs) {
    /* generate_target_point("random") :266-282, clamp :319-328 */
    const double scale = 200.0, margin = 50.0;
    double x_min = (-p->width / 2 + margin) / scale, x_max = (p->width / 2 - margin) / scale;
    double y_min = (-p->height / 2 + margin) / scale, y_max = (p->height / 2 - margin) / scale;
    double tx = x_min + (x_max - x_min) * next_double(t, row, &e->cur);
    double ty = y_min + (y_max - y_min) * next_double(t, row, &e->cur);
    e->target[0] = (float)clipd(tx, x_min, x_max);
    e->target[1] = (float)clipd(ty, y_min, y_max);
    robot_reset(p, e);
    e->prev_dist = dist_to_target(e);
    e->ep_return = 0.0;
    if (obs) write_obs(e, obs);
}

/* ---- Nozzle.solve_angles robot.py:55-83 + _get_rotation_matrices :153-178 + get_nozzle_direction :115-130 ---- */
static void nozzle_direction(double yaw, double *angle1_out, double *angle2_out, double d[3]) {
    /* target = R_br^T @ -(cos yaw, sin yaw, 0) = (-0, -sin yaw, cos yaw) */
    double t1 = -sin(yaw), t2 = cos(yaw);
    double angle2 = acos(clipd(2 * t2 - 1, -1.0, 1.0));
    if (angle2 <= -kPi) angle2 += 2 * kPi;
    else if (angle2 > kPi) angle2 -= 2 * kPi;
    double angle1;
    if (angle2 == 0) {
        angle1 = 0.0;
    } else {
        double a = 0.5 * (cos(angle2) - 1);
        double b = sqrt(2.0) * sin(angle2) / 2;
        double val1 = clipd(t1 / sqrt(a * a + b * b), -1.0, 1.0);
        angle1 = asin(val1) - atan2(b, a);
    }
    if (angle1 <= -kPi) angle1 += 2 * kPi;
    else if (angle1 > kPi) angle1 -= 2 * kPi;
    *angle1_out = angle1;
    *angle2_out = angle2;
    const double gamma = kPi / 4;
    double cg = cos(gamma), sg = sin(gamma);
    double c1 = cos(angle1), s1 = sin(angle1), c2 = cos(angle2), s2 = sin(angle2);
    /* R_nm = R_theta_fixed @ R_nozzle */
    double n00 = cg * c2, n01 = cg * -s2, n02 = -sg;
    double n10 = s2, n11 = c2, n12 = 0.0;
    double n20 = sg * c2, n21 = sg * -s2, n22 = cg;
    /* B = (R_br @ R_mb) @ R_nm, R_br @ R_mb = [[0,0,-1],[s1,c1,0],[c1,-s1,0]] */
    double b00 = -n20, b02 = -n22;
    double b10 = s1 * n00 + c1 * n10, b12 = s1 * n02 + c1 * n12;
    double b20 = c1 * n00 + -s1 * n10, b22 = c1 * n02 + -s1 * n12;
    (void)n01; (void)n11; (void)n21;
    d[0] = b00 * cg + b02 * sg;
    d[1] = b10 * cg + b12 * sg;
    d[2] = b20 * cg + b22 * sg;
}

/* One SalpRobotEnv.step (salp_robot_env.py:139-201) = one breathing cycle of Robot (robot.py:335-445). */
static void env_step(const CycleParams *p, CycleEnv *e, const double *action, float *obs, double *reward,
                     uint8_t *terminated, uint8_t *truncated) {
    /* _rescale_action :129-137 (float64 actions) */
    double contraction = action[0] * 0.06;
    double coast_time = action[1] * 10.0;
    double yaw = action[2] * (kPi / 2);
    double angle1, angle2, dir[3];
    nozzle_direction(yaw, &angle1, &angle2, dir);
    /* set_control :335-358 */
    e->cycle += 1;
    double cycle_time = 0.0;
    const double contract_rate = 0.06 / 3, release_rate = 0.06 / 1.5;
    double refill_time = contraction / contract_rate;
    double jet_time = contraction / release_rate;
    /* step_through_cycle :422-445 */
    double total = refill_time + jet_time + coast_time;
    double rj = refill_time + jet_time;
    double length = e->length, width = e->width_b;
    int n_sub = 0, phase = e->phase;
    double *v = e->vel, *w = e->w, *eu = e->euler;
    while (cycle_time < total) {
        if (p->max_substeps > 0 && n_sub >= p->max_substeps) break;
        ++n_sub;
        /* Robot.step :387-395 */
        cycle_time += DT;
        e->time += DT;
        /* update_state :360-373 */
        if (cycle_time <= refill_time) phase = 0;
        else if (cycle_time <= rj) phase = 1;
        else if (cycle_time <= rj + coast_time) phase = 2;
        else phase = 3;
        /* update_properties :375-385 */
        double prev_vol = e->volume;
        double prev_wm = prev_vol * p->density;
        if (phase == 0) {
            length = p->init_length - cycle_time * contract_rate;
            width = p->init_width + cycle_time * contract_rate;
        } else if (phase == 1) {
            length = p->init_length - contraction + (cycle_time - refill_time) * release_rate;
            width = p->init_width + contraction - (cycle_time - refill_time) * release_rate;
        } else {
            length = p->init_length;
            width = p->init_width;
        }
        double area = area_of(length, width);
        double volume = vol_of(length, width);
        e->volume = volume;
        double water_mass = p->density * volume;
        double mass = p->dry_mass + water_mass + p->noz_mass;
        double cd = drag_coefficient(p, length, width);
        /* _newton_equations :494-505 */
        double cx = w[1] * v[2] - w[2] * v[1], cy = w[2] * v[0] - w[0] * v[2], cz = w[0] * v[1] - w[1] * v[0];
        double fcor[3] = {mass * cx, mass * cy, mass * cz};
        double vnorm = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        double kd = -0.5 * p->density * area * cd;
        double fdrag[3];
        for (int i = 0; i < 3; ++i) fdrag[i] = kd * vnorm * v[i] + kd * v[i];
        double fjet[3] = {0.0, 0.0, 0.0};
        if (phase == 1) {
            double volume_rate = -(volume - prev_vol) / DT;
            double jet_speed = volume_rate / p->noz_area;
            double mass_rate = (water_mass - prev_wm) / DT;
            double cm = 0.1 * mass_rate;
            for (int i = 0; i < 3; ++i) fjet[i] = cm * (dir[i] * jet_speed);
        }
        double inv_m = 1.0 / mass;
        double acc[3];
        for (int i = 0; i < 3; ++i) acc[i] = inv_m * (fjet[i] + fdrag[i] + fcor[i]);
        /* _euler_equations :507-522 */
        double I[3];
        inertia_diag(p, mass, length, width, I);
        double Iw[3] = {I[0] * w[0], I[1] * w[1], I[2] * w[2]};
        double tcor[3] = {-(w[1] * Iw[2] - w[2] * Iw[1]), -(w[2] * Iw[0] - w[0] * Iw[2]), -(w[0] * Iw[1] - w[1] * Iw[0])};
        double wnorm = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
        double hl = length / 2;
        double kt = -p->density * cd * (width / 2) * pow(hl, 4.0) * wnorm;
        double tdrag[3] = {kt * w[0], kt * w[1], kt * w[2]};
        double rx = -(p->noz_l1 + p->noz_l2) + (-length / 2);
        double tjet[3] = {0.0 * fjet[2] - 0.0 * fjet[1], 0.0 * fjet[0] - rx * fjet[2], rx * fjet[1] - 0.0 * fjet[0]};
        double tasym[3] = {0.0, 0.0, 0.1 * vnorm};
        double tdef[3];
        for (int i = 0; i < 3; ++i) {
            tdef[i] = ((I[i] - e->prev_I[i]) / DT) * w[i];
            e->prev_I[i] = I[i];
        }
        double aacc[3];
        for (int i = 0; i < 3; ++i) aacc[i] = (1.0 / I[i]) * (tjet[i] + tdrag[i] + tcor[i] + tasym[i] - tdef[i]);
        /* _update_motion_states :524-532 */
        for (int i = 0; i < 3; ++i) { v[i] += acc[i] * DT; w[i] += aacc[i] * DT; }
        double sphi = sin(eu[0]), cphi = cos(eu[0]), tth = tan(eu[1]), cth = cos(eu[1]);
        double er0 = 1 * w[0] + sphi * tth * w[1] + cphi * tth * w[2];
        double er1 = 0 * w[0] + cphi * w[1] + -sphi * w[2];
        double er2 = 0 * w[0] + sphi / cth * w[1] + cphi / cth * w[2];
        eu[0] += er0 * DT; eu[1] += er1 * DT; eu[2] += er2 * DT;
        /* _to_world_frame :461-486 with the updated angles: R = (Rz @ Ry) @ Rx */
        double sp = sin(eu[0]), cp = cos(eu[0]), st = sin(eu[1]), ct = cos(eu[1]), ss = sin(eu[2]), cs = cos(eu[2]);
        /* Rz@Ry = [[cs*ct, -ss, cs*st],[ss*ct, cs, ss*st],[-st, 0, ct]] */
        double m00 = cs * ct, m01 = -ss, m02 = cs * st;
        double m10 = ss * ct, m11 = cs, m12 = ss * st;
        double m20 = -st, m22 = ct;
        /* (Rz@Ry)@Rx, Rx = [[1,0,0],[0,cp,-sp],[0,sp,cp]] */
        double r00 = m00, r01 = m01 * cp + m02 * sp, r02 = m01 * -sp + m02 * cp;
        double r10 = m10, r11 = m11 * cp + m12 * sp, r12 = m11 * -sp + m12 * cp;
        double r20 = m20, r21 = 0.0 * cp + m22 * sp, r22 = 0.0 * -sp + m22 * cp;
        e->vel_world[0] = r00 * v[0] + r01 * v[1] + r02 * v[2];
        e->vel_world[1] = r10 * v[0] + r11 * v[1] + r12 * v[2];
        e->vel_world[2] = r20 * v[0] + r21 * v[1] + r22 * v[2];
        for (int i = 0; i < 3; ++i) e->pos[i] += e->vel_world[i] * DT;
    }
    e->n_sub = n_sub;
    e->phase = phase;
    e->length = length;
    e->width_b = width;
    /* _calculate_reward :203-242 */
    double dx = e->pos[0] - (double)e->target[0], dy = e->pos[1] - (double)e->target[1];
    double dist = sqrt(dx * dx + dy * dy);
    double r_track = (-dist + e->prev_dist) * 100;
    e->prev_dist = dist;
    double en = dist + 1e-6;
    double ex = -(dx / en), ey = -(dy / en);
    double hn = sqrt(e->vel_world[0] * e->vel_world[0] + e->vel_world[1] * e->vel_world[1]) + 1e-6;
    double hx = e->vel_world[0] / hn, hy = e->vel_world[1] / hn;
    double r_heading = hx * ex + hy * ey;
    double r_smooth = -0.1 * (0.0 * 0.0);
    double rew = (1.0 * r_track) + (0.5 * r_heading) + 0.0 + r_smooth;
    /* termination :174-188 */
    int done = 0, trunc = 0;
    if (dist < 0.01) { done = 1; rew += 10.0; }
    else if (dist > 5.0) { trunc = 1; rew -= 5.0; }
    if (e->cycle >= p->max_cycles) trunc = 1;
    *reward = rew;
    *terminated = (uint8_t)done;
    *truncated = (uint8_t)trunc;
    e->ep_return += rew;
    write_obs(e, obs);
}

/* ---- batch entry points (ctypes: oracle/salp_cycle_oracle.py) ---- */
void salp_cycle_oracle_reset(const CycleParams *p, CycleEnv *envs, int64_t n, const CycleTables *t, int64_t row0,
                             const uint8_t *mask, int rewind_rng, float *obs) {
    for (int64_t i = 0; i < n; ++i) {
        if (mask && !mask[i]) continue;
        if (rewind_rng) envs[i].cur = 0;
        env_reset(p, &envs[i], t, row0 + i, obs ? obs + 6 * i : 0);
    }
}

/* One env step for n envs.  With autoreset, finished envs are reset after their terminal observation was
 * copied to final_obs (the vector env's convention; the reference itself never auto-resets).
 * Returns the number of integrator substeps executed.  OpenMP over envs when built with -fopenmp. */
int64_t salp_cycle_oracle_step(const CycleParams *p, CycleEnv *envs, int64_t n, const CycleTables *t, int64_t row0,
                               const double *actions, float *obs, double *reward, uint8_t *terminated,
                               uint8_t *truncated, int32_t *n_sub, float *final_obs, int autoreset) {
    int64_t total = 0;
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : total)
    for (int64_t i = 0; i < n; ++i) {
        CycleEnv *e = &envs[i];
        env_step(p, e, actions + 3 * i, obs + 6 * i, &reward[i], &terminated[i], &truncated[i]);
        if (n_sub) n_sub[i] = e->n_sub;
        total += e->n_sub;
        if (terminated[i] | truncated[i]) {
            if (final_obs) memcpy(final_obs + 6 * i, obs + 6 * i, 6 * sizeof(float));
            if (autoreset) env_reset(p, e, t, row0 + i, obs + 6 * i);
        }
    }
    return total;
}
