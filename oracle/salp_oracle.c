/*
 * salp_oracle.c -- CPU restatement (fp64, scalar C) of the reference hot path
 *                  SalpSnakeEnv.step/reset over the legacy SalpRobotEnv integrator.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may build, load or call this file.  The product
 * (underwater_swimmer_rl_b200/) never links, imports or falls back to it.
 *
 * Parity status: PINNED.  oracle/validate_against_reference.py runs the unmodified
 * Python reference (oracle/ref_harness.py) in the build container and checks that this
 * file reproduces it BIT-EXACTLY (all continuous state, observations, rewards, flags,
 * RNG consumption) for the four reference YAML configs in both breathing modes; the
 * same trajectories are committed under tests/golden/ and re-checked by the CPU tests.
 *
 * Reference files followed (paths relative to the reference tree):
 *   R = scripts/utilities/salp_robot.py            (legacy model "A")
 *   S = src/salp/environments/salp_snake_env.py
 * Every function below cites the R:/S: lines it restates.  Floating-point operation
 * ORDER follows the Python expressions exactly (Python floats / numpy float64 scalars
 * are IEEE-754 binary64, CPython's math.* calls this same glibc libm), so compile with
 *   gcc -O2 -ffp-contract=off -fno-fast-math
 * Note x**2 in the reference is libm pow(x, 2.0), which glibc 2.39 does NOT round
 * identically to x*x (about 1 in 1100 inputs differs by 1 ulp); pow() is kept here.
 *
 * Randomness: the reference draws food positions from the process-global `random`
 * (S:101-104,127-130,146,239-242,269-272) and thrust jitter from the global `np.random`
 * (R:311).  Both are MT19937; here each env reads its two streams from host-seeded
 * tables of raw 32-bit generator outputs ("words") through a per-env cursor:
 *   random.random() / np.random.random()  == ((w0>>5)*2^26 + (w1>>6)) / 2^53
 *   random.uniform(a,b)                   == a + (b-a)*random()
 *   random.randint(1,n)                   == 1 + _randbelow(n), getrandbits(k)=w>>(32-k)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define SALP_ORACLE_MAX_FOOD 32
#define SALP_ORACLE_MAX_OBS (10 + 4 * 8 + 2)

typedef struct {
    double width, height;            /* S:29 (ints in Python; exact in double) */
    int32_t num_food_items;          /* base_num_food_items, S:36 */
    int32_t max_observed_food;       /* S:44 */
    int32_t forced_breathing;        /* S:43,66 */
    int32_t random_food_count;       /* S:37 */
    int32_t respawn_food;            /* S:38 */
    int32_t max_steps_without_food;  /* S:46 */
    double food_reward, collision_penalty, time_penalty, efficiency_bonus,
        proximity_reward_weight;     /* S:39-45 */
} OracleParams;

typedef struct {
    double px, py, vx, vy, angle, omega;       /* R:61-64 */
    double nozzle, target_nozzle;              /* R:67-68 */
    double wv, a, b;                           /* water_volume, ellipse_a/b R:74-77 */
    double score;                              /* S:57 */
    double food[SALP_ORACLE_MAX_FOOD][2];      /* S:52 food_positions */
    int32_t alive[SALP_ORACLE_MAX_FOOD];       /* 0 == None */
    int32_t n_food;                            /* len(food_positions) == num_food_items */
    int32_t phase, timer, exhale_dur;          /* R:71-72,223  phase: 0 rest 1 inhaling 2 exhaling */
    int32_t is_inhaling;                       /* R:76 */
    int32_t food_collected, steps_since_food;  /* S:58-59 */
    /* bookkeeping for the tests only (never read by the model): which expression last wrote
     * ellipse_a/b -- 0 reset R:112-113, 1 rest R:190-191, 2 inhale R:212-213, 3 exhale R:245-246 --
     * and the breathing_timer value it used.  Mirrors the kernel's packed "shape" fields. */
    int32_t shape_kind, shape_timer, pad_;
    uint64_t cur_food, cur_jit;                /* words consumed from the two streams */
} OracleEnv;

typedef struct {
    const uint32_t *food_words; /* [rows][food_len] */
    const uint32_t *jit_words;  /* [rows][jit_len]  */
    int64_t rows, food_len, jit_len; /* lens are powers of two; cursors wrap */
} OracleTables;

/* R:34-53 constants */
static const double TANK_MARGIN = 50.0, BASE_RADIUS = 30.0, MAX_THRUST = 100.0, DRAG = 0.98;
static const double NOZZLE_RATE = 0.05;
static const int INHALE_DUR = 120, EXHALE_DUR = 150, REST_DUR = 60;
static const double FOOD_RADIUS = 15.0, MIN_FOOD_DIST = 80.0; /* S:53-54 */
#define PI 3.141592653589793 /* math.pi */

static inline double pymax(double x, double y) { return (y > x) ? y : x; } /* max(x,y) */
static inline double pymin(double x, double y) { return (y < x) ? y : x; } /* min(x,y) */
static inline double sq(double x) { return pow(x, 2.0); }                  /* x**2 */

static inline uint32_t next_word(const uint32_t *tab, int64_t row, int64_t len, uint64_t *cur) {
    uint32_t w = tab[row * len + (int64_t)(*cur & (uint64_t)(len - 1))];
    *cur += 1;
    return w;
}
static inline double next_double(const uint32_t *tab, int64_t row, int64_t len, uint64_t *cur) {
    uint32_t a = next_word(tab, row, len, cur) >> 5, b = next_word(tab, row, len, cur) >> 6;
    return ((double)a * 67108864.0 + (double)b) / 9007199254740992.0; /* genrand_res53 */
}
static inline double uniform_food(const OracleTables *t, int64_t row, OracleEnv *e, double lo, double hi) {
    return lo + (hi - lo) * next_double(t->food_words, row, t->food_len, &e->cur_food);
}

/* S:92-131 _generate_food_positions */
static void generate_food(const OracleParams *p, OracleEnv *e, const OracleTables *t, int64_t row) {
    const double xlo = TANK_MARGIN + FOOD_RADIUS, xhi = p->width - TANK_MARGIN - FOOD_RADIUS;
    const double ylo = TANK_MARGIN + FOOD_RADIUS, yhi = p->height - TANK_MARGIN - FOOD_RADIUS;
    int placed = 0;
    for (int i = 0; i < SALP_ORACLE_MAX_FOOD; ++i) e->alive[i] = 0;
    for (int i = 0; i < e->n_food; ++i) {
        int attempts = 0;
        while (attempts < 100) {
            double x = uniform_food(t, row, e, xlo, xhi);
            double y = uniform_food(t, row, e, ylo, yhi);
            int valid = 1;
            for (int j = 0; j < placed; ++j) {
                double d = sqrt(sq(x - e->food[j][0]) + sq(y - e->food[j][1]));
                if (d < MIN_FOOD_DIST) { valid = 0; break; }
            }
            double rd = sqrt(sq(x - p->width / 2) + sq(y - p->height / 2));
            if (rd < MIN_FOOD_DIST) valid = 0;
            if (valid) { e->food[placed][0] = x; e->food[placed][1] = y; e->alive[placed] = 1; ++placed; break; }
            ++attempts;
        }
        if (attempts >= 100) {
            double x = uniform_food(t, row, e, xlo, xhi);
            double y = uniform_food(t, row, e, ylo, yhi);
            e->food[placed][0] = x; e->food[placed][1] = y; e->alive[placed] = 1; ++placed;
        }
    }
}

/* R:371-388 _get_observation + S:366-428 _get_extended_observation */
static void extended_obs(const OracleParams *p, const OracleEnv *e, float *obs) {
    const double diag = sqrt(p->width * p->width + p->height * p->height);
    obs[0] = (float)(e->px / p->width);
    obs[1] = (float)(e->py / p->height);
    obs[2] = (float)(e->vx / 5.0);
    obs[3] = (float)(e->vy / 5.0);
    obs[4] = (float)(e->angle / PI);
    obs[5] = (float)(e->omega / 0.1);
    obs[6] = (float)(pymax(e->a, e->b) / BASE_RADIUS);
    obs[7] = (float)((double)e->phase / 2.0);
    obs[8] = (float)e->wv;
    obs[9] = (float)(e->nozzle / (PI / 3));
    /* live foods with distances, stable-sorted by distance (list.sort is stable) */
    int idx[SALP_ORACLE_MAX_FOOD], n = 0;
    double dist[SALP_ORACLE_MAX_FOOD];
    for (int i = 0; i < e->n_food; ++i)
        if (e->alive[i]) {
            dist[n] = sqrt(sq(e->food[i][0] - e->px) + sq(e->food[i][1] - e->py));
            idx[n] = i; ++n;
        }
    for (int i = 1; i < n; ++i) { /* insertion sort == stable */
        double d = dist[i]; int k = idx[i], j = i - 1;
        while (j >= 0 && dist[j] > d) { dist[j + 1] = dist[j]; idx[j + 1] = idx[j]; --j; }
        dist[j + 1] = d; idx[j + 1] = k;
    }
    float *fo = obs + 10;
    for (int i = 0; i < p->max_observed_food; ++i) {
        if (i < n) {
            const double fx = e->food[idx[i]][0], fy = e->food[idx[i]][1];
            double rel = atan2(fy - e->py, fx - e->px) - e->angle;
            while (rel > PI) rel -= 2 * PI;
            while (rel < -PI) rel += 2 * PI;
            fo[0] = (float)((fx - e->px) / p->width);
            fo[1] = (float)((fy - e->py) / p->height);
            fo[2] = (float)(dist[i] / diag);
            fo[3] = (float)(rel / PI);
        } else { fo[0] = 0.f; fo[1] = 0.f; fo[2] = 1.f; fo[3] = 0.f; }
        fo += 4;
    }
    fo[0] = (float)pymin((double)n / 10.0, 1.0);
    if (n > 0) {
        double s = 0.0; /* sum() starts from int 0 */
        for (int i = 0; i < n; ++i) s = s + dist[i];
        fo[1] = (float)((s / (double)n) / diag);
    } else fo[1] = 1.f;
}

/* R:95-117 SalpRobotEnv.reset + S:133-155 SalpSnakeEnv.reset */
static void env_reset(const OracleParams *p, OracleEnv *e, const OracleTables *t, int64_t row, float *obs) {
    e->px = p->width / 2; e->py = p->height / 2;
    e->vx = 0.0; e->vy = 0.0; e->angle = 0.0; e->omega = 0.0;
    e->nozzle = 0.0; e->target_nozzle = 0.0;
    e->phase = 0; e->timer = 0;
    e->a = BASE_RADIUS; e->b = BASE_RADIUS; e->shape_kind = 0; e->shape_timer = 0;
    e->is_inhaling = 0; e->wv = 0.0;
    e->score = 0.0; e->food_collected = 0; e->steps_since_food = 0;
    if (p->random_food_count) { /* S:144-146 random.randint(1, max(1, base)) */
        uint32_t n = (uint32_t)(p->num_food_items > 1 ? p->num_food_items : 1);
        int k = 0; for (uint32_t v = n; v; v >>= 1) ++k; /* n.bit_length() */
        uint32_t r;
        do { r = next_word(t->food_words, row, t->food_len, &e->cur_food) >> (32 - k); } while (r >= n);
        e->n_food = 1 + (int32_t)r;
    } else e->n_food = p->num_food_items;
    generate_food(p, e, t, row);
    if (obs) extended_obs(p, e, obs);
}

/* R:261-314 _apply_jet_thrust */
static void jet_thrust(OracleEnv *e, const OracleTables *t, int64_t row) {
    const double T = MAX_THRUST * e->wv * 0.4;
    const double th = e->angle - e->nozzle;
    const double tx = cos(th) * T, ty = sin(th) * T;
    e->vx += tx * 0.012; e->vy += ty * 0.012;
    const double primary = -e->nozzle * T * 0.0002;
    const double arm = pymax(e->a, e->b) * 0.7;
    const double perp = T * sin(-e->nozzle);
    const double moment = perp * arm * 0.00005;
    const double shape = -e->nozzle * T * e->wv * 0.00003;
    const double total = primary + moment + shape;
    e->omega += total;
    const double sa = th + PI / 2;
    const double sm = T * fabs(e->nozzle) * 0.3;
    const double sx = cos(sa) * sm, sy = sin(sa) * sm;
    e->vx += sx * 0.008; e->vy += sy * 0.008;
    const double u = next_double(t->jit_words, row, t->jit_len, &e->cur_jit);
    const double na = th + (u - 0.5) * 0.05;
    const double nf = T * 0.04;
    e->vx += cos(na) * nf * 0.002;
    e->vy += sin(na) * nf * 0.002;
}

/* R:119-156 SalpRobotEnv.step (reward/obs/info of the base class are dead values, S:160-169) */
static void robot_step(const OracleParams *p, OracleEnv *e, const double *action, const OracleTables *t, int64_t row) {
    double dir;
    if (p->forced_breathing) { /* R:123-127,158-167 */
        dir = action[0];
        e->is_inhaling = (e->timer % (INHALE_DUR + EXHALE_DUR + REST_DUR)) < INHALE_DUR;
    } else { /* R:128-132 */
        e->is_inhaling = action[0] > 0.5;
        dir = action[1];
    }
    const double maxn = PI / 3;
    e->target_nozzle = dir * maxn; /* R:135 */
    /* R:169-182 _update_nozzle */
    const double diff = e->target_nozzle - e->nozzle;
    if (fabs(diff) > NOZZLE_RATE) { if (diff > 0) e->nozzle += NOZZLE_RATE; else e->nozzle -= NOZZLE_RATE; }
    else e->nozzle = e->target_nozzle;
    e->nozzle = pymax(-maxn, pymin(maxn, e->nozzle));
    /* R:184-259 _update_breathing_cycle */
    if (e->phase == 0) {
        e->a = BASE_RADIUS * 1.3; e->b = BASE_RADIUS * 0.8; e->shape_kind = 1;
        if (e->is_inhaling) { e->phase = 1; e->timer = 0; }
    } else if (e->phase == 1) {
        if (e->is_inhaling && e->timer < INHALE_DUR) {
            e->timer += 1;
            const double pr = (double)e->timer / (double)INHALE_DUR;
            const double sa = BASE_RADIUS * 1.3, sb = BASE_RADIUS * 0.8, ea = BASE_RADIUS * 1.1, eb = BASE_RADIUS * 1.1;
            e->a = sa + (ea - sa) * pr; e->b = sb + (eb - sb) * pr;
            e->shape_kind = 2; e->shape_timer = e->timer;
            e->wv = pr;
        } else if (e->wv > 0.05) {
            e->phase = 2; e->timer = 0;
            e->exhale_dur = (int32_t)((double)EXHALE_DUR * pymax(e->wv, 0.3)); /* int() truncates */
        } else { e->phase = 0; e->timer = 0; e->wv = 0.0; }
    } else {
        e->timer += 1;
        const double pr = (double)e->timer / (double)e->exhale_dur;
        if (pr <= 1.0) {
            const double sa = BASE_RADIUS * 1.1, sb = BASE_RADIUS * 1.1, ea = BASE_RADIUS * 1.3, eb = BASE_RADIUS * 0.8;
            e->a = sa + (ea - sa) * pr; e->b = sb + (eb - sb) * pr;
            e->shape_kind = 3; e->shape_timer = e->timer;
            if (0.1 <= pr && pr <= 0.5) jet_thrust(e, t, row);
            const double w = e->wv * (1.0 - pr);
            e->wv = (w > 0) ? w : 0.0; /* max(0, w) */
        } else { e->phase = 0; e->timer = 0; e->wv = 0.0; }
    }
    /* R:316-352 _update_physics */
    e->vx *= DRAG; e->vy *= DRAG; e->omega *= 0.95;
    e->px += e->vx; e->py += e->vy;
    e->angle += e->omega;
    while (e->angle > PI) e->angle -= 2 * PI;
    while (e->angle < -PI) e->angle += 2 * PI;
    const double margin = TANK_MARGIN + pymax(e->a, e->b);
    if (e->px < margin) { e->px = margin; e->vx = fabs(e->vx) * 0.4; e->omega *= 0.7; }
    else if (e->px > p->width - margin) { e->px = p->width - margin; e->vx = -fabs(e->vx) * 0.4; e->omega *= 0.7; }
    if (e->py < margin) { e->py = margin; e->vy = fabs(e->vy) * 0.4; e->omega *= 0.7; }
    else if (e->py > p->height - margin) { e->py = p->height - margin; e->vy = -fabs(e->vy) * 0.4; e->omega *= 0.7; }
}

/* S:232-276 _respawn_food */
static void respawn_food(const OracleParams *p, OracleEnv *e, const OracleTables *t, int64_t row) {
    const double xlo = TANK_MARGIN + FOOD_RADIUS, xhi = p->width - TANK_MARGIN - FOOD_RADIUS;
    const double ylo = TANK_MARGIN + FOOD_RADIUS, yhi = p->height - TANK_MARGIN - FOOD_RADIUS;
    int attempts = 0;
    double x, y;
    while (attempts < 50) {
        x = uniform_food(t, row, e, xlo, xhi);
        y = uniform_food(t, row, e, ylo, yhi);
        const double rd = sqrt(sq(x - e->px) + sq(y - e->py));
        if (rd < MIN_FOOD_DIST) { ++attempts; continue; }
        int valid = 1;
        for (int j = 0; j < e->n_food; ++j)
            if (e->alive[j]) {
                const double d = sqrt(sq(x - e->food[j][0]) + sq(y - e->food[j][1]));
                if (d < MIN_FOOD_DIST) { valid = 0; break; }
            }
        if (valid) {
            for (int j = 0; j < e->n_food; ++j)
                if (!e->alive[j]) { e->food[j][0] = x; e->food[j][1] = y; e->alive[j] = 1; return; }
        }
        ++attempts;
    }
    x = uniform_food(t, row, e, xlo, xhi);
    y = uniform_food(t, row, e, ylo, yhi);
    for (int j = 0; j < e->n_food; ++j)
        if (!e->alive[j]) { e->food[j][0] = x; e->food[j][1] = y; e->alive[j] = 1; return; }
}

/* S:157-202 SalpSnakeEnv.step.  flags: bit0 terminated, bit1 truncated, bit2 collision, bit3 food hit */
static int env_step(const OracleParams *p, OracleEnv *e, const double *action, const OracleTables *t,
                    int64_t row, float *obs, double *reward) {
    robot_step(p, e, action, t, row);
    const double r = pymax(e->a, e->b);
    /* S:204-217 _check_food_collection */
    int hit = 0;
    for (int i = 0; i < e->n_food; ++i)
        if (e->alive[i]) {
            const double d = sqrt(sq(e->px - e->food[i][0]) + sq(e->py - e->food[i][1]));
            if (d < r + FOOD_RADIUS) { e->alive[i] = 0; hit = 1; break; }
        }
    /* S:219-230 _check_wall_collision */
    const int collision = (e->px - r <= TANK_MARGIN) || (e->px + r >= p->width - TANK_MARGIN) ||
                          (e->py - r <= TANK_MARGIN) || (e->py + r >= p->height - TANK_MARGIN);
    /* S:278-327 _calculate_snake_reward */
    double rew = 0.0;
    if (hit) {
        rew += p->food_reward;
        if (p->efficiency_bonus > 0) {
            const int remaining = p->max_steps_without_food - e->steps_since_food;
            rew += p->efficiency_bonus * (double)remaining;
        }
    }
    if (collision) rew += p->collision_penalty;
    if (p->proximity_reward_weight > 0) {
        int best = -1; double bd = 0.0; /* S:350-364 nearest live food, strict < */
        for (int i = 0; i < e->n_food; ++i)
            if (e->alive[i]) {
                const double d = sqrt(sq(e->px - e->food[i][0]) + sq(e->py - e->food[i][1]));
                if (best < 0 || d < bd) { bd = d; best = i; }
            }
        if (best >= 0) {
            double al = atan2(e->food[best][1] - e->py, e->food[best][0] - e->px) - e->angle;
            while (al > PI) al -= 2 * PI;
            while (al < -PI) al += 2 * PI;
            rew += p->proximity_reward_weight * cos(al);
        }
    }
    rew += p->time_penalty;
    /* S:171-180 */
    e->steps_since_food += 1;
    if (hit) {
        e->food_collected += 1;
        e->score += p->food_reward;
        e->steps_since_food = 0;
        if (p->respawn_food) respawn_food(p, e, t, row);
    }
    /* S:182-189 */
    int flags = (collision ? 4 : 0) | (hit ? 8 : 0);
    if (collision) flags |= 1;
    else if (e->steps_since_food > p->max_steps_without_food) flags |= 2;
    else if (!p->respawn_food) {
        int any = 0;
        for (int i = 0; i < e->n_food; ++i) any |= e->alive[i];
        if (!any) flags |= 1;
    }
    if (obs) extended_obs(p, e, obs);
    *reward = rew;
    return flags;
}

/* ------------------------------------------------------------------------------------ */
/* exported entry points (ctypes)                                                         */
/* ------------------------------------------------------------------------------------ */

int salp_oracle_env_size(void) { return (int)sizeof(OracleEnv); }
int salp_oracle_obs_dim(const OracleParams *p) { return 10 + 4 * p->max_observed_food + 2; }

void salp_oracle_reset(const OracleParams *p, OracleEnv *envs, int64_t n, int64_t env_id0,
                       const OracleTables *t, float *obs /* [n][obs_dim] or NULL */) {
    const int od = salp_oracle_obs_dim(p);
    for (int64_t i = 0; i < n; ++i)
        env_reset(p, &envs[i], t, (env_id0 + i) % t->rows, obs ? obs + i * od : NULL);
}

/*
 * The two word streams of the reference's generators, for callers that must not touch the product library (bench.py's
 * reference arm): row r of dst[rows][len] = the first `len` raw MT19937 outputs after
 *   kind 0: random.seed(seed0 + first_row + r)     CPython random_seed(): init_by_array([seed]) for seed < 2^32
 *   kind 1: np.random.seed(seed0 + first_row + r)  numpy legacy: init_genrand(seed)
 * (Matsumoto & Nishimura's MT19937 reference code, restated; tests/test_rng_tables.py checks both against the
 * Python stdlib and numpy, and against the product's salp_fill_rng_words.)
 */
typedef struct { uint32_t mt[624]; int idx; } OracleMT;
static void mt_init_genrand(OracleMT *g, uint32_t s) {
    g->mt[0] = s;
    for (int i = 1; i < 624; ++i) g->mt[i] = 1812433253u * (g->mt[i - 1] ^ (g->mt[i - 1] >> 30)) + (uint32_t)i;
    g->idx = 624;
}
static void mt_init_by_array1(OracleMT *g, uint32_t key) {
    mt_init_genrand(g, 19650218u);
    int i = 1;
    for (int k = 624; k; --k) {
        g->mt[i] = (g->mt[i] ^ ((g->mt[i - 1] ^ (g->mt[i - 1] >> 30)) * 1664525u)) + key + 0u;
        if (++i >= 624) { g->mt[0] = g->mt[623]; i = 1; }
    }
    for (int k = 623; k; --k) {
        g->mt[i] = (g->mt[i] ^ ((g->mt[i - 1] ^ (g->mt[i - 1] >> 30)) * 1566083941u)) - (uint32_t)i;
        if (++i >= 624) { g->mt[0] = g->mt[623]; i = 1; }
    }
    g->mt[0] = 0x80000000u;
    g->idx = 624;
}
static uint32_t mt_next(OracleMT *g) {
    if (g->idx >= 624) {
        for (int k = 0; k < 624; ++k) {
            const uint32_t y = (g->mt[k] & 0x80000000u) | (g->mt[(k + 1) % 624] & 0x7fffffffu);
            g->mt[k] = g->mt[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        g->idx = 0;
    }
    uint32_t y = g->mt[g->idx++];
    y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
    return y;
}
void salp_oracle_fill_rng_words(uint32_t *dst, int64_t rows, int64_t len, uint64_t seed0, int64_t first_row, int kind) {
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < rows; ++r) {
        OracleMT g;
        const uint32_t s = (uint32_t)(seed0 + (uint64_t)first_row + (uint64_t)r);
        if (kind == 0) mt_init_by_array1(&g, s); else mt_init_genrand(&g, s);
        for (int64_t i = 0; i < len; ++i) dst[r * len + i] = mt_next(&g);
    }
}

/* S:366-428 _get_extended_observation on the CURRENT state of n envs (no step): what the reference's callers
 * get after poking robot_pos / robot_angle / food_positions (eval/collect_navigation_data.py:76-89). */
void salp_oracle_observe(const OracleParams *p, const OracleEnv *envs, int64_t n, float *obs /* [n][obs_dim] */) {
    const int od = salp_oracle_obs_dim(p);
    for (int64_t i = 0; i < n; ++i) extended_obs(p, &envs[i], obs + i * od);
}

/* One step for n envs.  actions [n][act_dim] double.  Outputs may be NULL. */
void salp_oracle_step(const OracleParams *p, OracleEnv *envs, int64_t n, int64_t env_id0,
                      const OracleTables *t, const double *actions, float *obs, double *reward,
                      uint8_t *flags) {
    const int od = salp_oracle_obs_dim(p), ad = p->forced_breathing ? 1 : 2;
    for (int64_t i = 0; i < n; ++i) {
        double r;
        int f = env_step(p, &envs[i], actions + i * ad, t, (env_id0 + i) % t->rows, obs ? obs + i * od : NULL, &r);
        if (reward) reward[i] = r;
        if (flags) flags[i] = (uint8_t)f;
    }
}

/*
 * Rollout of T steps for n envs with caller-style auto-reset ("if done or truncated:
 * env.reset()").  actions [T][n][act_dim] (float32 values widened
 * by the caller, like float(action[0]) in R:125).  Outputs [T][n]... may be NULL:
 *   obs       post-step observation (the TERMINAL one on done steps)
 *   reset_obs observation returned by the auto-reset (rows of non-done steps untouched)
 * Threads: OpenMP over envs when compiled with -fopenmp (envs are independent).
 * Returns total env-steps executed.
 */
int64_t salp_oracle_rollout(const OracleParams *p, OracleEnv *envs, int64_t n, int64_t env_id0,
                            const OracleTables *t, int64_t T, const float *actions, int autoreset,
                            float *obs, float *reset_obs, double *reward, uint8_t *flags,
                            int64_t *stats /* [6]: episodes, food, collisions, truncations, steps, _ */,
                            double *ret_sum) {
    const int od = salp_oracle_obs_dim(p), ad = p->forced_breathing ? 1 : 2;
    int64_t episodes = 0, food = 0, coll = 0, trunc = 0;
    double rsum = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : episodes, food, coll, trunc, rsum)
    for (int64_t i = 0; i < n; ++i) {
        OracleEnv *e = &envs[i];
        const int64_t row = (env_id0 + i) % t->rows;
        float scratch[SALP_ORACLE_MAX_OBS];
        for (int64_t s = 0; s < T; ++s) {
            const float *af = actions + (s * n + i) * ad;
            double a[2] = {(double)af[0], ad > 1 ? (double)af[1] : 0.0};
            double r;
            float *o = obs ? obs + (s * n + i) * od : scratch;
            int f = env_step(p, e, a, t, row, o, &r);
            if (reward) reward[s * n + i] = r;
            if (flags) flags[s * n + i] = (uint8_t)f;
            rsum += r;
            if (f & 8) ++food;
            if (f & 3) {
                ++episodes;
                if (f & 4) ++coll;
                if (f & 2) ++trunc;
                if (autoreset) env_reset(p, e, t, row, reset_obs ? reset_obs + (s * n + i) * od : scratch);
            }
        }
    }
    if (stats) { stats[0] = episodes; stats[1] = food; stats[2] = coll; stats[3] = trunc; stats[4] = n * T; stats[5] = 0; }
    if (ret_sum) *ret_sum = rsum;
    return n * T;
}
