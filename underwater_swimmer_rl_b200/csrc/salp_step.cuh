// salp_step.cuh -- the SALP env-step kernel for sm_100a (one thread per environment).
//
// What it replaces (reference tree, Python):
//   R = scripts/utilities/salp_robot.py             SalpRobotEnv.step            R:119-156
//   S = src/salp/environments/salp_snake_env.py     SalpSnakeEnv.step / reset    S:133-202
// K consecutive env steps per launch: the per-env state is read ONCE with 128-bit (f32) /
// 256-bit (f64) vector loads from structure-of-arrays records, lives in registers across the
// K substeps and is written back ONCE; per substep only action is read and only
// obs / reward / terminated / truncated are written.  The observation tile of a CTA
// (256 envs x obs_dim floats, contiguous in HBM) is staged in shared memory and written with
// one TMA bulk store (cp.async.bulk.global.shared::cta) per substep, double-buffered so the
// store of substep k overlaps the arithmetic of substep k+1.  Rare, divergent work (food
// respawn, episode reset: rejection sampling from the RNG word tables) lives in __noinline__
// slow paths that talk to global memory directly so it costs no registers on the hot path.
// Episode statistics are reduced with warp shuffles and one atomic per warp, only in warps
// where an episode actually ended.
//
// Arithmetic: templated on the real type.  The double instantiation is compiled with
// -fmad=false and keeps the reference's operation order so that discrete events match the
// reference bit-for-bit; the float instantiation is the throughput build.
#pragma once
#include <math_constants.h>

#include "salp_internal.h"
#include "salp_shape_host.h"
#include "salp_sortnets.cuh"

namespace salp {

template <typename T> struct alignas(4 * sizeof(T)) Vec4 { T x, y, z, w; };
template <typename T> struct alignas(2 * sizeof(T)) Vec2 { T x, y; };


template <typename T> struct is_f64 { static constexpr bool value = false; };
template <> struct is_f64<double> { static constexpr bool value = true; };

// Task constants in the build's real type: the doubles of the parameter block (validation build) or their float
// roundings made once on the host (throughput build; a per-thread double -> float conversion is an XU-pipe op).
#define SALP_CONST(name, dfield, ffield)                                                             \
    template <typename T> __device__ __forceinline__ T name(const KernelArgs &a) {                  \
        if constexpr (is_f64<T>::value) return a.dfield; else return a.ffield;                     \
    }
SALP_CONST(c_W, W, fW) SALP_CONST(c_H, H, fH) SALP_CONST(c_diag, diag, f_diag)
SALP_CONST(c_inv_w, inv_w, f_inv_w) SALP_CONST(c_inv_h, inv_h, f_inv_h) SALP_CONST(c_inv_diag, inv_diag, f_inv_diag)
SALP_CONST(c_food_reward, food_reward, f_food_reward) SALP_CONST(c_collision_penalty, collision_penalty, f_collision_penalty)
SALP_CONST(c_time_penalty, time_penalty, f_time_penalty) SALP_CONST(c_efficiency_bonus, efficiency_bonus, f_efficiency_bonus)
SALP_CONST(c_proximity_w, proximity_w, f_proximity_w)
#undef SALP_CONST

#ifdef SALP_LAB_NOMATH  // tools/kernel_lab.cu only: how much of the step is transcendental math?
__device__ __forceinline__ float sqrt_(float x) { return x * 0.5f; }
__device__ __forceinline__ float sin_(float x) { return x * 0.25f; }
__device__ __forceinline__ float cos_(float x) { return 1.0f - x * 0.125f; }
__device__ __forceinline__ void sincos_(float x, float *s, float *c) { *s = x * 0.25f; *c = 1.0f - x * 0.125f; }
__device__ __forceinline__ float atan2_(float y, float x) { return y * 0.001f + x * 0.002f; }
#else
// f32 throughput build: sqrt.approx.f32 (one MUFU op, max relative error 2^-23 per the PTX ISA) instead
// of the IEEE sequence with its special-case branch -- the 12-food configs take 12 of these per step.
__device__ __forceinline__ float sqrt_(float x) {
#ifdef SALP_LAB_IEEE_SQRT
    return sqrtf(x);
#else
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));  // .ftz: no denormal pre-scaling (distances are pixels)
    return r;
#endif
}
__device__ __forceinline__ double sqrt_(double x) { return sqrt(x); }
__device__ __forceinline__ double sin_(double x) { return sin(x); }
__device__ __forceinline__ double cos_(double x) { return cos(x); }
__device__ __forceinline__ void sincos_(double x, double *s, double *c) { sincos(x, s, c); }
#ifdef SALP_LAB_LIBM  // tools/kernel_lab.cu only: CUDA libm instead of the fast paths below
__device__ __forceinline__ float sin_(float x) { return sinf(x); }
__device__ __forceinline__ float cos_(float x) { return cosf(x); }
__device__ __forceinline__ void sincos_(float x, float *s, float *c) { sincosf(x, s, c); }
__device__ __forceinline__ float atan2_(float y, float x) { return atan2f(y, x); }
#else
// f32 throughput build: every angle the model feeds to sin/cos is within [-4pi/3, 4pi/3] (|angle| <= pi,
// |nozzle| <= pi/3), where the SFU approximations (MUFU.SIN/COS after a multiply by 1/2pi) have an absolute
// error of about 2^-21 -- two orders of magnitude below what one f32 rounding of a position in a
// 800-pixel tank already costs (ulp(512) = 6e-5 px), and five below the 1e-3 parity tolerance.
__device__ __forceinline__ float sin_(float x) { return __sinf(x); }
__device__ __forceinline__ float cos_(float x) { return __cosf(x); }
__device__ __forceinline__ void sincos_(float x, float *s, float *c) { __sincosf(x, s, c); }
// atan2 in ~20 instructions: a = min/max in [0,1] (one MUFU.RCP), odd minimax polynomial
// a + a^3 P(a^2) with deg P = 6 (max abs error 1.1e-7 evaluated in f32, fitted in tools/fit_atan.py),
// then octant/quadrant/sign fix-ups.  atan2(0, 0) = 0 like math.atan2.
__device__ __forceinline__ float atan2_(float y, float x) {
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    float rmx;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rmx) : "f"(fmaxf(mx, 1e-30f)));
    const float a = mn * rmx;
    const float s = a * a;
    float r = -0.0043554045259952545f;
    r = fmaf(r, s, 0.023040130734443665f);
    r = fmaf(r, s, -0.05777358263731003f);
    r = fmaf(r, s, 0.09794233739376068f);
    r = fmaf(r, s, -0.13976581394672394f);
    r = fmaf(r, s, 0.19962704181671143f);
    r = fmaf(r, s, -0.3333165943622589f);
    r = fmaf(r * s, a, a);
    r = (ay > ax) ? 1.5707963267948966f - r : r;
    r = (x < 0.f) ? 3.141592653589793f - r : r;
    return copysignf(r, y);
}
#endif
#endif
__device__ __forceinline__ double atan2_(double y, double x) { return atan2(y, x); }
__device__ __forceinline__ float abs_(float x) { return fabsf(x); }
__device__ __forceinline__ double abs_(double x) { return fabs(x); }

// Python's max(x, y) / min(x, y): keep the first argument unless the second is strictly better.
template <typename T> __device__ __forceinline__ T pymax(T x, T y) { return (y > x) ? y : x; }
template <typename T> __device__ __forceinline__ T pymin(T x, T y) { return (y < x) ? y : x; }

// x / c for an OBSERVATION column (c a compile-time constant): multiply by the folded reciprocal in both builds.  The
// result is rounded to float32 right away; in the validation build the product is within 1 ulp(double) of the reference's
// quotient, so the float32 value differs from the reference's in about one value in 2^29 (bar: 2e-6).  An IEEE double
// division is ~30 instructions and the observation has 24 of them: a third of the fp64 step's instructions.
// Nothing that feeds the STATE is computed this way.
template <typename T> __device__ __forceinline__ T divc(T x, double c) { return x * (T)(1.0 / c); }

static __device__ __noinline__ unsigned wall_touch_bits(int skind, int stimer, int exd, double W, double H) {
    double ea, eb;
    shape_axes_f64(skind, stimer, exd, ea, eb);
    return wall_touch_bits_r(eb > ea ? eb : ea, W, H);
}

// R:329-332, S:311-314, S:403-406:  while x > pi: x -= 2pi;  while x < -pi: x += 2pi.
// One predicated correction per direction reproduces the loops' first iterations (same operations,
// same order, bit-identical) and is all that ever runs for values the model itself produces
// (|angle| <= pi, |omega| << pi, |atan2| <= pi); the loops remain as the tail for poked-in states.
template <typename T> __device__ __forceinline__ T wrap_pi(T x) {
    x = (x > (T)kc::PI) ? x - (T)kc::TWO_PI : x;
    x = (x < (T)(-kc::PI)) ? x + (T)kc::TWO_PI : x;
    if (abs_(x) > (T)kc::PI) {
        while (x > (T)kc::PI) x -= (T)kc::TWO_PI;
        while (x < (T)(-kc::PI)) x += (T)kc::TWO_PI;
    }
    return x;
}

// ---------------------------------------------------------------------------------------------
// RNG word tables (see include/salp_b200.h): genrand_res53 from two consecutive MT19937 outputs
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double res53(uint32_t w0, uint32_t w1) {
    return ((double)(w0 >> 5) * 67108864.0 + (double)(w1 >> 6)) * (1.0 / 9007199254740992.0);
}
// Philox4x32-10 (Salmon et al., SC'11).  rng_mode 1: word w of stream s (0 = the `random` stream of food
// positions, 1 = the `np.random` stream of thrust jitter) of the env with GLOBAL id g is
//     philox4x32_10(counter = (w >> 2, s, g_lo, g_hi), key = (seed_lo, seed_hi))[w & 3]
// -- unbounded, aperiodic streams that need no memory and do not depend on how envs are sharded over GPUs.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}
__device__ __forceinline__ uint4 philox_block(const KernelArgs &a, unsigned e, unsigned stream, unsigned w) {
    const unsigned long long g = (unsigned long long)a.env_id0 + e;
    return philox4x32_10(make_uint4(w >> 2, stream, (uint32_t)g, (uint32_t)(g >> 32)),
                         make_uint2((uint32_t)a.rng_seed, (uint32_t)(a.rng_seed >> 32)));
}
__device__ __forceinline__ uint32_t pick_word(uint4 r, unsigned w) {
    const unsigned i = w & 3u;
    return i == 0 ? r.x : (i == 1 ? r.y : (i == 2 ? r.z : r.w));
}
// the pair of words (w, w + 1) of the jitter stream, w even (np.random.random() consumes two words, R:311)
__device__ __forceinline__ uint2 philox_jitter_pair(const KernelArgs &a, unsigned e, unsigned w) {
    const uint4 r = philox_block(a, e, 1u, w);
    return (w & 2u) ? make_uint2(r.z, r.w) : make_uint2(r.x, r.y);
}
// `e` is the env's index within the handle: table row e % rows, or the Philox stream of global env env_id0 + e
__device__ __forceinline__ uint32_t food_word(const KernelArgs &a, unsigned e, unsigned &cur) {
    uint32_t w;
    if (a.rng_mode) w = pick_word(philox_block(a, e, 0u, cur), cur);
    else w = a.food_words[(size_t)(e % a.rows) * a.food_len + (cur & a.food_len_mask)];
    ++cur;
    return w;
}
// the four words of one (x, y) candidate: in Philox mode they are ONE block whenever the cursor sits on a multiple of 4
// (always, unless random_food_count's randint has moved it), instead of one block per word
__device__ __forceinline__ uint4 food_words4(const KernelArgs &a, unsigned e, unsigned &cur) {
    uint4 w;
    if (a.rng_mode && (cur & 3u) == 0u) { w = philox_block(a, e, 0u, cur); cur += 4u; }
    else { w.x = food_word(a, e, cur); w.y = food_word(a, e, cur); w.z = food_word(a, e, cur); w.w = food_word(a, e, cur); }
    return w;
}
// random.uniform(xlo, xhi), random.uniform(ylo, yhi): a + (b - a) * random(), two words per draw
template <typename T> __device__ __forceinline__ void food_xy(const KernelArgs &a, unsigned e, unsigned &cur, T xlo, T xhi, T ylo, T yhi, T &x, T &y) {
    const uint4 w = food_words4(a, e, cur);
    x = xlo + (xhi - xlo) * (T)res53(w.x, w.y);
    y = ylo + (yhi - ylo) * (T)res53(w.z, w.w);
}
template <typename T> __device__ __forceinline__ T food_uniform(const KernelArgs &a, unsigned row, unsigned &cur, T lo, T hi) {
    const uint32_t w0 = food_word(a, row, cur), w1 = food_word(a, row, cur);  // `row` = env index, see food_word
    return lo + (hi - lo) * (T)res53(w0, w1);  // random.uniform: a + (b-a)*random()
}

// ---------------------------------------------------------------------------------------------
// Slow paths: __noinline__, arguments and results BY VALUE (so no hot-path variable ever has its
// address taken), operate on global memory, executed by a handful of lanes per step
// ---------------------------------------------------------------------------------------------
template <typename T> struct RespawnResult { T x, y; unsigned cur, mask; int slot; };
struct ResetResult { unsigned cur, mask; int n_food; };

// S:232-276 _respawn_food.  Live foods are read from HBM (food records only ever change here and in the
// reset paths, so HBM is always current) -- all of them up front with independent loads, so that a respawn
// costs about one memory round trip per attempt instead of one per food per attempt: the slowest lane of a
// launch sets the tail of a one-wave grid.  Returns the refilled slot and its position.
template <typename T, int F>
__device__ __noinline__ RespawnResult<T> respawn_slow(const KernelArgs &a, long long e, T px, T py, unsigned mask,
                                                      int n_food, unsigned cur) {
    const unsigned row = (unsigned)e;  // env index: food_word picks the table row or the Philox stream
    Vec2<T> *food = reinterpret_cast<Vec2<T> *>(a.food);
    const T xlo = (T)(kc::TANK_MARGIN + kc::FOOD_R), xhi = (T)a.W - (T)kc::TANK_MARGIN - (T)kc::FOOD_R;
    const T ylo = xlo, yhi = (T)a.H - (T)kc::TANK_MARGIN - (T)kc::FOOD_R;
    Vec2<T> live[F];
#pragma unroll
    for (int j = 0; j < F; ++j)
        live[j] = (j < n_food && ((mask >> j) & 1u)) ? food[(size_t)j * a.N + e] : Vec2<T>{0, 0};
    int attempts = 0;
    T x = 0, y = 0;
    bool found = false;
    while (attempts < 50) {
        food_xy<T>(a, row, cur, xlo, xhi, ylo, yhi, x, y);
        const T rx = x - px, ry = y - py;
        bool valid = !(sqrt_(rx * rx + ry * ry) < (T)kc::MIN_FOOD_DIST);
#pragma unroll
        for (int j = 0; j < F; ++j)
            if (j < n_food && ((mask >> j) & 1u)) {
                const T dx = x - live[j].x, dy = y - live[j].y;
                if (sqrt_(dx * dx + dy * dy) < (T)kc::MIN_FOOD_DIST) valid = false;
            }
        if (valid) { found = true; break; }
        ++attempts;
    }
    if (!found) food_xy<T>(a, row, cur, xlo, xhi, ylo, yhi, x, y);
    int slot = -1;
    for (int j = 0; j < n_food; ++j)
        if (!((mask >> j) & 1u)) { slot = j; break; }
    if (slot >= 0) {
        food[(size_t)slot * a.N + e] = Vec2<T>{x, y};
        mask |= 1u << slot;
    }
    return RespawnResult<T>{x, y, cur, mask, slot};
}

// Warp-cooperative S:232-276 _respawn_food for ONE env, run by all 32 lanes of a converged warp (every argument
// warp-uniform).  The reference draws up to 50 candidates one at a time and keeps the first that is >= 80 px from the robot
// and from every live food, then draws a 51st unconditionally; here lane L materialises candidate 32 w + L of window w
// (w = 0, 1) and a ballot picks the first valid one -- the same candidate, the same words consumed, in one pass instead
// of a serial rejection loop run by a single lane while 31 wait (the slowest lane of a launch sets a small slab's time).
// Lane 0 writes the new food record; the result is returned in every lane.
template <typename T, int F>
__device__ __noinline__ RespawnResult<T> respawn_warp(const KernelArgs &a, unsigned e, T px, T py, unsigned mask, int n_food,
                                                     unsigned cur, int lane) {
    constexpr unsigned kFull = 0xffffffffu;
    Vec2<T> *food = reinterpret_cast<Vec2<T> *>(a.food);
    const T xlo = (T)(kc::TANK_MARGIN + kc::FOOD_R), xhi = (T)a.W - (T)kc::TANK_MARGIN - (T)kc::FOOD_R;
    const T ylo = xlo, yhi = (T)a.H - (T)kc::TANK_MARGIN - (T)kc::FOOD_R;
    Vec2<T> live[F];  // the same address in every lane: one transaction per food, all in flight together
#pragma unroll
    for (int j = 0; j < F; ++j)
        live[j] = (j < n_food && ((mask >> j) & 1u)) ? food[(size_t)j * a.N + e] : Vec2<T>{0, 0};
    T x = 0, y = 0;
    int chosen = 50;  // candidate index == rejected attempts before it; candidate 50 is the unconditional draw
#pragma unroll 1
    for (int base = 0; base < 64; base += 32) {
        const int idx = base + lane;
        unsigned c2 = cur + 4u * (unsigned)idx;
        T cx, cy;
        food_xy<T>(a, e, c2, xlo, xhi, ylo, yhi, cx, cy);
        const T rx = cx - px, ry = cy - py;
        bool ok = !(sqrt_(rx * rx + ry * ry) < (T)kc::MIN_FOOD_DIST);
#pragma unroll
        for (int j = 0; j < F; ++j)
            if (j < n_food && ((mask >> j) & 1u)) {
                const T dx = cx - live[j].x, dy = cy - live[j].y;
                if (sqrt_(dx * dx + dy * dy) < (T)kc::MIN_FOOD_DIST) ok = false;
            }
        const unsigned m = __ballot_sync(kFull, ok && idx < 50);
        const int c = m ? base + __ffs((int)m) - 1 : (base == 32 ? 50 : -1);
        if (c >= 0) {
            x = __shfl_sync(kFull, cx, c - base);
            y = __shfl_sync(kFull, cy, c - base);
            chosen = c;
            break;
        }
    }
    cur += 4u * (unsigned)(chosen + 1);
    int slot = -1;
    for (int j = 0; j < n_food; ++j)
        if (!((mask >> j) & 1u)) { slot = j; break; }
    if (slot >= 0) {
        if (lane == 0) food[(size_t)slot * a.N + e] = Vec2<T>{x, y};
        mask |= 1u << slot;
    }
    return RespawnResult<T>{x, y, cur, mask, slot};
}

// S:133-155 reset (food part: S:92-131 _generate_food_positions, S:144-148 food count) and the
// reset observation (S:366-428 with the robot at the tank centre, R:95-117) written to obs_row.
template <typename T>
__device__ __noinline__ ResetResult reset_slow(const KernelArgs &a, long long e, unsigned cur, float *obs_row) {
    const unsigned row = (unsigned)e;  // env index: food_word picks the table row or the Philox stream
    Vec2<T> *food = reinterpret_cast<Vec2<T> *>(a.food);
    int n_food = a.num_food;
    if (a.random_food_count) {  // random.randint(1, max(1, base)) == 1 + _randbelow(n)
        const uint32_t n = (uint32_t)(a.num_food > 1 ? a.num_food : 1);
        const int kbits = 32 - __clz(n);
        uint32_t r;
        do { r = food_word(a, row, cur) >> (32 - kbits); } while (r >= n);
        n_food = 1 + (int)r;
    }
    const T xlo = (T)(kc::TANK_MARGIN + kc::FOOD_R), xhi = (T)a.W - (T)kc::TANK_MARGIN - (T)kc::FOOD_R;
    const T ylo = xlo, yhi = (T)a.H - (T)kc::TANK_MARGIN - (T)kc::FOOD_R;
    const T cx = (T)a.W / 2, cy = (T)a.H / 2;
    T dist[kMaxFood];
    int idx[kMaxFood];
    for (int i = 0; i < n_food; ++i) {
        int attempts = 0;
        T x = 0, y = 0;
        while (attempts < 100) {
            food_xy<T>(a, row, cur, xlo, xhi, ylo, yhi, x, y);
            bool valid = true;
            for (int j = 0; j < i; ++j) {
                const Vec2<T> f = food[(size_t)j * a.N + e];
                const T dx = x - f.x, dy = y - f.y;
                if (sqrt_(dx * dx + dy * dy) < (T)kc::MIN_FOOD_DIST) { valid = false; break; }
            }
            const T rx = x - cx, ry = y - cy;
            if (sqrt_(rx * rx + ry * ry) < (T)kc::MIN_FOOD_DIST) valid = false;
            if (valid) break;
            ++attempts;
        }
        if (attempts >= 100) food_xy<T>(a, row, cur, xlo, xhi, ylo, yhi, x, y);
        food[(size_t)i * a.N + e] = Vec2<T>{x, y};
        const T rx = x - cx, ry = y - cy;
        dist[i] = sqrt_(rx * rx + ry * ry);
        idx[i] = i;
    }
    const unsigned mask = n_food >= 32 ? 0xffffffffu : ((1u << n_food) - 1u);
    if (obs_row != nullptr) {
        // observation of the freshly reset env: robot at the centre, all zeros, body radius 30
        obs_row[0] = (float)(cx / (T)a.W); obs_row[1] = (float)(cy / (T)a.H);
        obs_row[2] = 0.f; obs_row[3] = 0.f; obs_row[4] = 0.f; obs_row[5] = 0.f;
        obs_row[6] = (float)((T)kc::BASE_R / (T)kc::BASE_R);
        obs_row[7] = 0.f; obs_row[8] = 0.f; obs_row[9] = 0.f;
        for (int i = 1; i < n_food; ++i) {  // stable insertion sort by distance
            const T d = dist[i];
            const int k = idx[i];
            int j = i - 1;
            while (j >= 0 && dist[j] > d) { dist[j + 1] = dist[j]; idx[j + 1] = idx[j]; --j; }
            dist[j + 1] = d; idx[j + 1] = k;
        }
        const T diag = sqrt_((T)a.W * (T)a.W + (T)a.H * (T)a.H);
        float *fo = obs_row + 10;
        for (int i = 0; i < a.observed; ++i) {
            if (i < n_food) {
                const Vec2<T> f = food[(size_t)idx[i] * a.N + e];
                const T dx = f.x - cx, dy = f.y - cy;
                const T rel = wrap_pi<T>(atan2_(dy, dx) - (T)0);
                fo[0] = (float)(dx / (T)a.W); fo[1] = (float)(dy / (T)a.H);
                fo[2] = (float)(dist[i] / diag); fo[3] = (float)(rel / (T)kc::PI);
            } else { fo[0] = 0.f; fo[1] = 0.f; fo[2] = 1.f; fo[3] = 0.f; }
            fo += 4;
        }
        fo[0] = (float)pymin<T>((T)n_food / (T)10.0, (T)1.0);
        if (n_food > 0) {
            T s = 0;
            for (int i = 0; i < n_food; ++i) s = s + dist[i];
            fo[1] = (float)((s / (T)n_food) / diag);
        } else fo[1] = 1.f;
    }
    return ResetResult{cur, mask, n_food};
}

// "sqrt(d2) < 80" (S:113, S:252: distance below MIN_FOOD_DIST) without the square root, for the many pair tests of
// layout_warp.  In IEEE doubles with a correctly rounded sqrt (CUDA's sqrt.rn.f64, glibc's sqrt in the oracle, CPython's
// math.sqrt) sqrt(x) rounds to 80 exactly for x = 6400 - 2^-40, the double just below 6400, and below 80 for every smaller x,
// so the test is x < 6400 - 2^-40 -- the same verdict for every double (tests/test_forced_cycle.py checks the neighbourhood
// with numpy; the fp64 parity tests check the layouts bit for bit).  The f32 build keeps its approximate square root.
template <typename T> __device__ __forceinline__ bool closer_than_min_food_dist(T d2) {
    static_assert(kc::MIN_FOOD_DIST == 80.0, "the threshold below is sqrt_rn's preimage of [0, 80)");
    if constexpr (is_f64<T>::value) return d2 < 0x1.8ffffffffffffp+12;
    else return sqrt_(d2) < (T)kc::MIN_FOOD_DIST;
}

// reset_slow for the f32 build with at most ONE food (SalpLayout.food_slots == 1): no arrays, no sort, no divisions --
// the reset observation in closed form with the host's reciprocals, as the step's own observation code computes it.
// A finished episode is rare per env but present in some tile of every launch, and a lone lane runs dependent
// instructions at ~5 cycles each: the length of this routine is the tail of a small slab's step.
static __device__ __noinline__ ResetResult reset_single_f32(const KernelArgs &a, unsigned e, unsigned cur, float *obs_row) {
    float2 *food = reinterpret_cast<float2 *>(a.food);
    int n_food = a.num_food;
    if (a.random_food_count) {  // random.randint(1, 1): _randbelow(1) still consumes words (k = 1 bit, until the bit is 0)
        uint32_t r;
        do { r = food_word(a, e, cur) >> 31; } while (r >= 1u);
        n_food = 1;
    }
    const float xlo = (float)(kc::TANK_MARGIN + kc::FOOD_R), xhi = a.fW - (float)kc::TANK_MARGIN - (float)kc::FOOD_R;
    const float ylo = xlo, yhi = a.fH - (float)kc::TANK_MARGIN - (float)kc::FOOD_R;
    const float cx = a.fW * 0.5f, cy = a.fH * 0.5f;
    float rx = 0.f, ry = 0.f, d = 0.f;
    if (n_food > 0) {
        float x = 0.f, y = 0.f;
        int attempts = 0;
        for (;;) {  // S:96-117: up to 100 candidates at least 80 px from the robot, then one unconditional draw
            food_xy<float>(a, e, cur, xlo, xhi, ylo, yhi, x, y);
            rx = x - cx; ry = y - cy;
            d = sqrt_(rx * rx + ry * ry);
            if (attempts >= 100 || !(d < (float)kc::MIN_FOOD_DIST)) break;
            ++attempts;
        }
        food[e] = make_float2(x, y);
    }
    if (obs_row != nullptr) {
        float4 ent = make_float4(0.f, 0.f, 1.f, 0.f);  // S:412 padding
        float avg = 1.f;
        if (n_food > 0) {
            ent = make_float4(rx * a.f_inv_w, ry * a.f_inv_h, d * a.f_inv_diag, wrap_pi<float>(atan2_(ry, rx) - 0.f) * (float)(1.0 / kc::PI));
            avg = d * a.f_inv_diag;
        }
        float4 *r4 = reinterpret_cast<float4 *>(obs_row);
        r4[0] = make_float4(cx * a.f_inv_w, cy * a.f_inv_h, 0.f, 0.f);
        r4[1] = make_float4(0.f, 0.f, 1.0f, 0.f);
        r4[2] = make_float4(0.f, 0.f, ent.x, ent.y);
        r4[3] = make_float4(ent.z, ent.w, 0.f, 0.f);
        r4[4] = make_float4(1.f, 0.f, 0.f, 0.f);
        r4[5] = make_float4(1.f, 0.f, fminf((float)n_food * 0.1f, 1.0f), avg);
    }
    return ResetResult{cur, n_food > 0 ? 1u : 0u, n_food};
}

// Warp-cooperative S:92-131 _generate_food_positions (+ S:144-148 food count) for ONE env, run by all
// 32 lanes of a converged warp (e, cur warp-uniform).
//
// The reference draws candidates one at a time and keeps a candidate iff it is >= 80 px from the tank centre and
// from every food kept so far: a greedy scan in stream order.  Here lane L materialises candidate L of a
// 32-candidate window of the env's `random` stream (4 words each), tests it against the centre and the foods kept
// from earlier windows, and computes once which other candidates of the window lie within 80 px of it (31
// independent distance tests).  Which candidates the greedy scan keeps is then decided for the whole window in a
// few rounds of two ballots: an undecided candidate with no undecided conflicting predecessor is kept, and every
// undecided successor in conflict with a kept one is dropped -- the same set as the serial scan, because its
// decision for a candidate depends only on earlier candidates.  Lane i then fetches the (i+1)-th kept candidate.
// (A lone warp runs dependent instructions at ~5-10 cycles each; the serial scan was ~45 dependent instructions
// per food, 3 300 cycles for twelve, and the slowest reset sets the time of a small slab.)
// The 100-attempt cap with its unconditional 101st draw (S:100-117) can only bind for the first food sought in a
// window (the rejections carried over from earlier windows); that food alone takes the serial path.  Candidate
// order, distance arithmetic and the words consumed are exactly the reference's.
// Returns food j's position in lane j (j < n_food), plus the new cursor and n_food in every lane.
template <typename T> struct LayoutResult { T x, y; unsigned cur; int n_food; };

#ifdef SALP_LAB_CLOCK
static __device__ unsigned long long g_lab_lay[8];  // [0] calls; cycles: [1] window words+uniform [2] kept-food checks [3] conflict mask [4] total; [5] windows
#define SALP_LAY_T(v) const long long v = clock64()
#define SALP_LAY_ADD(i, d) do { if (lane == 0) atomicAdd(&g_lab_lay[i], (unsigned long long)(d)); } while (0)
#else
#define SALP_LAY_T(v) do { } while (0)
#define SALP_LAY_ADD(i, d) do { } while (0)
#endif

template <typename T>
__device__ __noinline__ LayoutResult<T> layout_warp(const KernelArgs &a, unsigned e, unsigned cur, int lane) {
    constexpr unsigned kFull = 0xffffffffu;
    const unsigned row = (a.rows == (unsigned)a.N) ? e : e % a.rows;
    const uint32_t *words = a.rng_mode ? nullptr : a.food_words + (size_t)row * a.food_len;
    int n_food = a.num_food;
    if (a.random_food_count) {  // random.randint(1, max(1, base)) == 1 + _randbelow(n); same words in every lane
        const uint32_t n = (uint32_t)(a.num_food > 1 ? a.num_food : 1);
        const int kbits = 32 - __clz(n);
        uint32_t r;
        do { unsigned c2 = cur; r = food_word(a, e, c2) >> (32 - kbits); ++cur; } while (r >= n);
        n_food = 1 + (int)r;
    }
    const T xlo = (T)(kc::TANK_MARGIN + kc::FOOD_R), xhi = (T)a.W - (T)kc::TANK_MARGIN - (T)kc::FOOD_R;
    const T ylo = xlo, yhi = (T)a.H - (T)kc::TANK_MARGIN - (T)kc::FOOD_R;
    const T cx = (T)a.W / 2, cy = (T)a.H / 2;
    const unsigned below = (1u << lane) - 1u;  // the candidates before `lane`
    T myx = 0, myy = 0;    // position of food `lane`
    int placed = 0;        // foods kept so far
    int attempts = 0;      // candidates rejected so far for food `placed` (S:100 caps them at 100)
    int ptr = 0;           // next untested candidate of the current window
    unsigned base = cur;   // stream position of candidate 0 of the next window
    SALP_LAY_T(t_begin);
    SALP_LAY_ADD(0, 1);
    while (placed < n_food) {
        // ---- a window of 32 candidates
        SALP_LAY_T(t0);
        const unsigned w = base + 4u * (unsigned)lane;
        uint32_t w0, w1, w2, w3;
        if (a.rng_mode) {
            if ((w & 3u) == 0u) {  // a candidate's four words are one Philox block unless randint moved the cursor off 4
                const uint4 r = philox_block(a, e, 0u, w);
                w0 = r.x; w1 = r.y; w2 = r.z; w3 = r.w;
            } else {
                unsigned c2 = w;
                w0 = food_word(a, e, c2); w1 = food_word(a, e, c2); w2 = food_word(a, e, c2); w3 = food_word(a, e, c2);
            }
        } else {
            w0 = words[w & a.food_len_mask]; w1 = words[(w + 1) & a.food_len_mask];
            w2 = words[(w + 2) & a.food_len_mask]; w3 = words[(w + 3) & a.food_len_mask];
        }
        const T candx = xlo + (xhi - xlo) * (T)res53(w0, w1);  // random.uniform: a + (b-a)*random()
        const T candy = ylo + (yhi - ylo) * (T)res53(w2, w3);
        const T rx = candx - cx, ry = candy - cy;
        bool base_ok = !(sqrt_(rx * rx + ry * ry) < (T)kc::MIN_FOOD_DIST);
        SALP_LAY_T(t1);
        for (int j = 0; j < placed; ++j) {  // foods kept from earlier windows
            const T dx = candx - __shfl_sync(kFull, myx, j), dy = candy - __shfl_sync(kFull, myy, j);
            if (sqrt_(dx * dx + dy * dy) < (T)kc::MIN_FOOD_DIST) base_ok = false;
        }
        SALP_LAY_T(t2);
        unsigned conf = 0u;  // bit j: candidate j of this window is within 80 px of candidate `lane`
#pragma unroll 8
        for (int j = 0; j < 32; ++j) {
            const T dx = candx - __shfl_sync(kFull, candx, j), dy = candy - __shfl_sync(kFull, candy, j);
            if (closer_than_min_food_dist<T>(dx * dx + dy * dy)) conf |= 1u << j;
        }
        conf &= ~(1u << lane);
        const unsigned earlier = conf & below;
        unsigned ok = __ballot_sync(kFull, base_ok);  // candidates that are still valid (warp-uniform)
        SALP_LAY_T(t3);
        SALP_LAY_ADD(1, t1 - t0); SALP_LAY_ADD(2, t2 - t1); SALP_LAY_ADD(3, t3 - t2); SALP_LAY_ADD(5, 1);
        base += 128u;
        ptr = 0;
        // ---- serial path, one food at a time, while the 100-attempt cap could bind inside this window
        while (placed < n_food && ptr < 32 && attempts > 68 + ptr) {
            const unsigned m = ok & (kFull << ptr);
            const int first = m ? __ffs((int)m) - 1 : 32;
            const int limit = ptr + (100 - attempts);  // after 100 rejects the next draw is taken unconditionally
            const int c = first < limit ? first : limit;
            if (c >= 32) { attempts += 32 - ptr; ptr = 32; break; }
            const T x = __shfl_sync(kFull, candx, c), y = __shfl_sync(kFull, candy, c);
            if (lane == placed) { myx = x; myy = y; }
            ok &= ~__shfl_sync(kFull, conf, c);  // candidates within 80 px of the new food are out
            ptr = c + 1;
            attempts = 0;
            ++placed;
        }
        // ---- the rest of the window at once: the cap cannot bind (attempts <= 68 + ptr, and 0 after a food is kept)
        if (placed < n_food && ptr < 32) {
            unsigned und = ok & (kFull << ptr);  // undecided candidates
            unsigned kept = 0u;
            while (und != 0u) {
                const bool u = (und >> lane) & 1u;
                const unsigned s = __ballot_sync(kFull, u && (earlier & und) == 0u);   // no undecided conflicting predecessor: kept
                const unsigned d = __ballot_sync(kFull, u && (earlier & s) != 0u);     // in conflict with a kept predecessor: dropped
                kept |= s;
                und &= ~(s | d);
            }
            const int n_kept = __popc(kept);
            const int take = n_kept < n_food - placed ? n_kept : n_food - placed;
            const int r = lane - placed;  // food `lane` is the (r + 1)-th kept candidate
            const bool get = r >= 0 && r < take;
            const unsigned src = get ? __fns(kept, 0u, r + 1) : 0u;
            const T sx = __shfl_sync(kFull, candx, (int)src), sy = __shfl_sync(kFull, candy, (int)src);
            if (get) { myx = sx; myy = sy; }
            if (take > 0) { ptr = (int)__fns(kept, 0u, take) + 1; attempts = 0; }
            placed += take;
            if (placed < n_food) { attempts += 32 - ptr; ptr = 32; }  // window exhausted, the remaining candidates are rejects
        }
    }
    SALP_LAY_T(t_end);
    SALP_LAY_ADD(4, t_end - t_begin);
    if (n_food > 0) cur = base - 128u + 4u * (unsigned)ptr;
    return LayoutResult<T>{myx, myy, cur, n_food};
}

// ---------------------------------------------------------------------------------------------
// TMA bulk store helpers (1-D, no tensor map): shared::cta -> global
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void bulk_store_commit(void *gdst, const void *ssrc, uint32_t bytes) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(ssrc);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(s), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int kPending> __device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kPending) : "memory");
}

// 64-bit load that asks L2 to keep the line (evict_last): an env draws its jitter words sequentially, so
// the 32-byte sector serves four consecutive jet steps if it survives the ~230 MB that stream through L2
// per launch; 1 M envs x 32 B = 32 MB of such sectors fit the 126 MB L2.
__device__ __forceinline__ uint2 ld_evict_last_u2(const uint32_t *p) {
    uint2 v;
#ifdef SALP_LAB_NO_EVICT_LAST
    v = *reinterpret_cast<const uint2 *>(p);
#else
    uint64_t pol;
    asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("ld.global.L2::cache_hint.v2.u32 {%0, %1}, [%2], %3;" : "=r"(v.x), "=r"(v.y) : "l"(p), "l"(pol));
#endif
    return v;
}

__device__ __forceinline__ uint32_t ld_evict_last_u1(const uint32_t *p) {
    uint32_t v;
#ifdef SALP_LAB_NO_EVICT_LAST
    v = *p;
#else
    uint64_t pol;
    asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("ld.global.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
#endif
    return v;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// cp.async (LDGSTS): global -> shared without a register round trip; 4/8/16-byte forms, larger
// records as several 16-byte copies.
template <int BYTES> __device__ __forceinline__ void cp_async(void *sdst, const void *gsrc) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(sdst);
    if constexpr (BYTES == 4 || BYTES == 8 || BYTES == 16) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(s), "l"(gsrc), "n"(BYTES) : "memory");
    } else {
        static_assert(BYTES % 16 == 0, "record size");
#pragma unroll
        for (int o = 0; o < BYTES; o += 16)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(s + o), "l"((const char *)gsrc + o) : "memory");
    }
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// How the observation tile leaves the SM (A/B-measured in tools/kernel_lab.cu):
//   kStoreWarp     each warp stages its 32 rows (32*obs_dim floats, contiguous in HBM) in shared
//                  memory and writes them with fully coalesced 128-bit stores; no CTA barrier
//   kStoreTmaWarp  same staging, one TMA bulk store (cp.async.bulk) per warp per substep
//   kStoreTmaBlock whole-CTA tile, one TMA bulk store per CTA per substep (needs __syncthreads)
//   kStoreDirect   no staging: each thread writes its own row (6 x 128-bit, 96-byte stride)
enum : int { kStoreWarp = 0, kStoreTmaWarp = 1, kStoreTmaBlock = 2, kStoreDirect = 3 };

// shared-memory bytes the observation tiles take (the PF stage starts after them, 128-byte aligned)
template <int STORE, int BLOCK> __host__ __device__ constexpr unsigned obs_tile_buffers() {
    return STORE == kStoreDirect ? 0u : (STORE == kStoreWarp ? 1u : 2u);
}
template <int STORE, int BLOCK> __host__ __device__ inline unsigned stage_offset_bytes(int obs_dim) {
    return (obs_tile_buffers<STORE, BLOCK>() * BLOCK * (unsigned)obs_dim * 4u + 127u) & ~127u;
}
template <typename T, int F, int BLOCK> __host__ __device__ constexpr unsigned stage_bytes() {
    return 2u * (32u * BLOCK + 2u * (unsigned)sizeof(T) * 4u * BLOCK + (unsigned)sizeof(T) * 2u * F * BLOCK + 3u * (unsigned)sizeof(T) * BLOCK);
}

// ---------------------------------------------------------------------------------------------
// The step kernel
//   T     : float | double          F : food records per env (1,4,8,12,16,32)
//   MT    : 3 -> max_observed_food == 3 (reference default; vectorised row assembly)
//           8 -> any max_observed_food in 0..8 (generic scalar row assembly)
//   STORE : see above               BLOCK : threads per CTA (multiple of 32)
//   MINB  : CTAs per SM the register allocation must allow (__launch_bounds__)
//   PF    : 0 one tile per CTA, state loaded straight into registers
//           1 persistent CTAs, next tile's state prefetched into shared memory with cp.async
//   RNG   : 0 thrust jitter from the word table, 1 from the in-kernel Philox stream (== a.rng_mode; a template parameter so
//           that the table-mode kernel carries none of the generator's registers)
// ---------------------------------------------------------------------------------------------
template <typename T, int F, int MT, int STORE, int BLOCK, int MINB, int PF, int RNG = 0>
__global__ void __launch_bounds__(BLOCK, MINB) salp_step_kernel(const __grid_constant__ KernelArgs a) {
    extern __shared__ __align__(128) float s_tile[];  // obs tiles [nbuf][BLOCK][obs_dim], then the PF stage
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int OBS = a.obs_dim;
    const unsigned N = (unsigned)a.N;
    // env indices fit 32 bits (salp_create caps num_envs at 2^30): cheap IMAD.WIDE addressing
    const unsigned end = (unsigned)a.first + (unsigned)a.count;
    const unsigned ntiles = ((unsigned)a.count + BLOCK - 1) / BLOCK;
    constexpr bool kTma = (STORE == kStoreTmaWarp || STORE == kStoreTmaBlock);
    // many foods: episode resets are laid out by the whole warp (layout_warp) and the reset observation
    // comes from a second trip through the observation code; few foods: in-thread reset_slow
    constexpr bool kCoopReset = (F >= 8);
    // throughput build with several foods: the MU nearest live foods are selected on packed integer keys
    // (distance bits | slot) with min/max, and their offsets fetched from a shared-memory scratch by slot
    constexpr bool kFastSel = (!is_f64<T>::value) && (F > 1);
    constexpr int MU = (MT < F) ? MT : F;  // observation entries that can ever be real
    constexpr unsigned kSlotMask = F <= 2 ? 1u : (F <= 4 ? 3u : (F <= 8 ? 7u : (F <= 16 ? 15u : 31u)));  // slot bits of a selection key
    __shared__ float2 s_sel[kFastSel ? F * BLOCK : 1];  // [slot][thread] (dx, dy) of this substep

    Vec4<T> *g_body = reinterpret_cast<Vec4<T> *>(a.body);
    T *g_epret = reinterpret_cast<T *>(a.ep_return);
    const Vec2<T> *g_food = reinterpret_cast<const Vec2<T> *>(a.food);
    T *g_reward = reinterpret_cast<T *>(a.reward);
    const T *g_act = reinterpret_cast<const T *>(a.actions);
    const T W = c_W<T>(a), H = c_H<T>(a);
    const T diag = c_diag<T>(a);  // sqrt(W^2 + H^2), S:396, computed on the host in double
    const T invW = c_inv_w<T>(a), invH = c_inv_h<T>(a), invDiag = c_inv_diag<T>(a);  // observation scales (host reciprocals)

    // programmatic dependent launch (salp_launch.cuh): let the next kernel of the stream be scheduled as soon as every
    // CTA of this grid is resident, and do not touch global memory before the previous kernel has completed
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    // ---- PF: persistent CTAs; the state records of the CTA's next tile are copied global->shared
    //      with cp.async (each thread fetches exactly the records it will read itself, so no CTA
    //      barrier is needed) while the current tile is being integrated.
    constexpr unsigned kOffC0 = 0, kOffC1 = kOffC0 + 16 * BLOCK, kOffB0 = kOffC1 + 16 * BLOCK;
    constexpr unsigned kOffB1 = kOffB0 + sizeof(Vec4<T>) * BLOCK, kOffFood = kOffB1 + sizeof(Vec4<T>) * BLOCK;
    constexpr unsigned kOffEp = kOffFood + sizeof(Vec2<T>) * F * BLOCK, kOffAct = kOffEp + sizeof(T) * BLOCK;
    constexpr unsigned kStageBytes = kOffAct + 2 * sizeof(T) * BLOCK;
    unsigned char *s_stage = reinterpret_cast<unsigned char *>(s_tile) + stage_offset_bytes<STORE, BLOCK>(OBS);
    auto stage_issue = [&](unsigned t, int buf) {
        const unsigned pe = (unsigned)a.first + t * BLOCK + tid;
        if (pe >= end) return;
        unsigned char *sb = s_stage + (size_t)buf * kStageBytes;
        cp_async<16>(sb + kOffC0 + 16 * tid, a.ctrl + pe);
        cp_async<16>(sb + kOffC1 + 16 * tid, a.ctrl + (size_t)N + pe);
        cp_async<sizeof(Vec4<T>)>(sb + kOffB0 + sizeof(Vec4<T>) * tid, g_body + pe);
        cp_async<sizeof(Vec4<T>)>(sb + kOffB1 + sizeof(Vec4<T>) * tid, g_body + (size_t)N + pe);
#pragma unroll
        for (int f = 0; f < F; ++f)
            cp_async<sizeof(Vec2<T>)>(sb + kOffFood + sizeof(Vec2<T>) * (f * BLOCK + tid), g_food + (size_t)f * N + pe);
        cp_async<sizeof(T)>(sb + kOffEp + sizeof(T) * tid, g_epret + pe);
        if (a.act_dim == 1) cp_async<sizeof(T)>(sb + kOffAct + sizeof(T) * tid, g_act + pe);
        else cp_async<2 * sizeof(T)>(sb + kOffAct + 2 * sizeof(T) * tid, g_act + 2 * (size_t)pe);
    };
    if constexpr (PF) {
        if (blockIdx.x < ntiles) { stage_issue(blockIdx.x, 0); cp_async_commit(); }
    }

    unsigned gk = 0;  // substeps this CTA has published so far (obs tile double-buffer parity)
    unsigned it = 0;
    for (unsigned tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const unsigned blk0 = (unsigned)a.first + tile * BLOCK;
        const unsigned e = blk0 + tid;
        const bool valid = e < end;

        // ---- load state: 4 x 128-bit records + ep_return + F food records, all independent loads
        Vec4<T> b0{0, 0, 0, 0}, b1{0, 0, 0, 0};
        int4 c0 = make_int4(0, 0, 0, 0), c1 = make_int4(0, 0, 0, 0);
        T ep_ret = 0;
        T fx[F], fy[F];
    #pragma unroll
        for (int f = 0; f < F; ++f) { fx[f] = 0; fy[f] = 0; }
        if constexpr (PF) {
            // prefetch the NEXT tile of this CTA while this one is computed, then wait for this one
            const unsigned nxt = tile + gridDim.x;
            if (nxt < ntiles) { stage_issue(nxt, (it + 1) & 1); cp_async_commit(); cp_async_wait<1>(); }
            else cp_async_wait<0>();
            const unsigned char *sb = s_stage + (size_t)(it & 1) * kStageBytes;
            if (valid) {
                c0 = reinterpret_cast<const int4 *>(sb + kOffC0)[tid];
                c1 = reinterpret_cast<const int4 *>(sb + kOffC1)[tid];
                b0 = reinterpret_cast<const Vec4<T> *>(sb + kOffB0)[tid];
                b1 = reinterpret_cast<const Vec4<T> *>(sb + kOffB1)[tid];
                ep_ret = reinterpret_cast<const T *>(sb + kOffEp)[tid];
                if constexpr (!kFastSel) {
    #pragma unroll
                    for (int f = 0; f < F; ++f) {
                        const Vec2<T> v = reinterpret_cast<const Vec2<T> *>(sb + kOffFood)[f * BLOCK + tid];
                        fx[f] = v.x; fy[f] = v.y;
                    }
                }
            }
        } else if (valid) {
            c0 = a.ctrl[e];
            c1 = a.ctrl[(size_t)N + e];
            b0 = g_body[e];
            b1 = g_body[(size_t)N + e];
            ep_ret = g_epret[e];
            if constexpr (!kFastSel) {
    #pragma unroll
                for (int f = 0; f < F; ++f) {
                    const Vec2<T> v = g_food[(size_t)f * N + e];
                    fx[f] = v.x; fy[f] = v.y;
                }
            }
        }
        T px = b0.x, py = b0.y, vx = b0.z, vy = b0.w;
        T angle = b1.x, omega = b1.y, nozzle = b1.z, wv = b1.w;
        int phase = c0.x & 3, timer = (c0.x >> 2) & 0xff, exd = (c0.x >> 10) & 0xff;
        int skind = (c0.x >> 18) & 3, stimer = (c0.x >> 20) & 0xff;
        int steps_since_food = c0.y, food_collected = c0.z;
        unsigned mask = (unsigned)c0.w;
        unsigned cur_food = (unsigned)c1.x, cur_jit = (unsigned)c1.y;
        int ep_len = c1.z & kEpLenMask, n_food = (int)((unsigned)c1.z >> 26);
        unsigned jcache = (unsigned)c1.w;  // first word of the jitter pair at cur_jit (see salp_b200.h)

        const unsigned trow = (a.rows == N) ? e : e % a.rows;
        const uint32_t *jit_row = RNG ? nullptr : a.jit_words + (size_t)trow * a.jit_len;

        for (int k = 0; k < a.K; ++k, ++gk) {
            // ---- where this thread's observation row goes
            float *s_buf = nullptr;   // tile base (block or warp) in shared memory
            float *row;
            if constexpr (STORE == kStoreDirect) {
                row = a.obs + ((size_t)k * N + (valid ? e : (unsigned)a.first)) * OBS;
            } else if constexpr (STORE == kStoreTmaBlock) {
                s_buf = s_tile + (size_t)(gk & 1) * BLOCK * OBS;
                row = s_buf + (size_t)tid * OBS;
                if (gk >= 2) {  // the bulk store issued two substeps ago must have finished reading s_buf
                    if (tid == 0) bulk_wait_read<1>();
                    __syncthreads();
                }
            } else {
                const int nbuf = (STORE == kStoreTmaWarp) ? 2 : 1;
                s_buf = s_tile + ((size_t)(tid >> 5) * nbuf + (STORE == kStoreTmaWarp ? (gk & 1) : 0)) * 32 * OBS;
                row = s_buf + (size_t)lane * OBS;
                if constexpr (STORE == kStoreTmaWarp) {
                    if (gk >= 2) { if (lane == 0) bulk_wait_read<1>(); __syncwarp(); }
                } else if (gk >= 1) __syncwarp();  // previous substep's cooperative copy has read the tile
            }
            // ---- does this step fire the jet?  (exhaling, and 0.1 <= (timer+1)/D <= 0.5, R:231-250.)
            //      Known before anything else is computed, so the jitter word load is issued first.
            const int t1 = timer + 1;
            const bool thrust = (phase == kPhaseExhale) && (t1 <= exd) && (10 * t1 >= exd) && (2 * t1 <= exd);
            // f32 build: the np.random word this step draws is already here (ctrl rec1.w, loaded with the state)
            // and the word for the FOLLOWING draw is fetched now, needed only at write-back -- the
            // cursor -> table DRAM round trip is off the critical path.  (23 bits of the draw are plenty for a
            // +-0.025 rad jitter in f32.)  The validation build needs both words of the pair bit-exactly and
            // keeps the dependent load; it is not a throughput configuration.
            uint2 jw = make_uint2(jcache, 0u);
#ifndef SALP_LAB_NOJIT
            if (thrust) {
                if constexpr (RNG != 0) jw = philox_jitter_pair(a, e, cur_jit);  // RNG: compile-time copy of a.rng_mode
                else if constexpr (is_f64<T>::value) jw = ld_evict_last_u2(jit_row + (cur_jit & a.jit_len_mask));
                else jcache = ld_evict_last_u1(jit_row + ((cur_jit + 2u) & a.jit_len_mask));
            }
#endif
            // ---- kFastSel: the food records are read here, every substep (HBM at k = 0, L1/L2 afterwards;
            //      they only ever change through the slow paths, which write HBM), so that no food array
            //      has to stay in registers across the observation assembly.  Issued before the integrator
            //      so the loads are in flight while it runs.
            Vec2<T> fv[kFastSel ? F : 1];
            if constexpr (kFastSel) {
    #pragma unroll
                for (int f = 0; f < F; ++f) fv[f] = valid ? g_food[(size_t)f * N + e] : Vec2<T>{0, 0};
            }
            // ---- action (R:121-135)
            T act0 = 0, act1 = 0;
            if (PF && k == 0) {
                const T *sa = reinterpret_cast<const T *>(s_stage + (size_t)(it & 1) * kStageBytes + kOffAct);
                if (valid) {
                    if (a.act_dim == 1) act0 = sa[tid];
                    else { act0 = sa[2 * tid]; act1 = sa[2 * tid + 1]; }
                }
            } else if (valid) {
                if (a.act_dim == 1) act0 = g_act[(size_t)k * N + e];
                else {
                    const Vec2<T> v = reinterpret_cast<const Vec2<T> *>(g_act)[(size_t)k * N + e];
                    act0 = v.x; act1 = v.y;
                }
            }
#ifdef SALP_LAB_SKELETON  // tools/kernel_lab.cu only: memory skeleton of the step (no model arithmetic)
            T radius = (T)30; bool hit = false; const bool collision = false; const T near_rel = 0;
            const bool terminated = false, truncated = false;
            px += act0 * (T)0.001 + (T)jw.x * (T)1e-12; T reward = act0 + act1;
            T dx[F], dy[F], dd[F];
#pragma unroll
            for (int f = 0; f < F; ++f) { dx[f] = fx[f] - px; dy[f] = fy[f] - py; dd[f] = dx[f] + dy[f]; }
            unsigned key[MU]; T dist_sum = 0;
    #pragma unroll
            for (int i = 0; i < MU; ++i) key[i] = (unsigned)i;
            if constexpr (kFastSel) {
    #pragma unroll
                for (int f = 0; f < F; ++f) s_sel[f * BLOCK + tid] = make_float2((float)dx[f], (float)dy[f]);
            }
            auto reselect = [&]() {};
#else
            bool inhaling;
            T dir;
            if (a.forced) {  // R:158-167: (timer % 330) < 120; timer <= 151 so the modulo is the identity
                dir = act0;
                inhaling = timer < kc::INHALE_DUR;
            } else {
                inhaling = act0 > (T)0.5;
                dir = act1;
            }
            // ---- nozzle slew (R:169-182)
            {
                const T target = dir * (T)kc::MAX_NOZZLE;
                const T diff = target - nozzle;
                if (abs_(diff) > (T)kc::NOZZLE_RATE) nozzle = (diff > 0) ? nozzle + (T)kc::NOZZLE_RATE : nozzle - (T)kc::NOZZLE_RATE;
                else nozzle = target;
                nozzle = pymax<T>((T)(-kc::MAX_NOZZLE), pymin<T>((T)kc::MAX_NOZZLE, nozzle));
            }
            // ---- breathing FSM on integers (R:184-259); thresholds 0.1<=t/D<=0.5 and t/D<=1 are
            //      exact integer compares (margins >= 1/(10*150) >> ulp), see tests/test_fsm_invariants.py
            int wv_mode = 0;  // 0 keep, 1 wv=p, 2 wv=max(0,wv*(1-p)), 3 wv=0
            if (phase == kPhaseRest) {
                skind = kShapeRest;
                if (inhaling) { phase = kPhaseInhale; timer = 0; }
            } else if (phase == kPhaseInhale) {
                if (inhaling && timer < kc::INHALE_DUR) {
                    ++timer; skind = kShapeInhale; stimer = timer; wv_mode = 1;
                } else if (wv > (T)0.05) {
                    exd = a.exhale_lut[timer];
                    phase = kPhaseExhale; timer = 0;
                } else { phase = kPhaseRest; timer = 0; wv_mode = 3; }
            } else {
                ++timer;
                if (timer <= exd) { skind = kShapeExhale; stimer = timer; wv_mode = 2; }
                else { phase = kPhaseRest; timer = 0; wv_mode = 3; }
            }
            // ---- ellipse axes from the shape fields (R:190-191, 207-213, 240-246)
            T ea, eb, prog = 0;
            if (skind == kShapeReset) { ea = (T)kc::BASE_R; eb = (T)kc::BASE_R; }
            else if (skind == kShapeRest) { ea = (T)kc::A_REST; eb = (T)kc::B_REST; }
            else {
                const bool inh = (skind == kShapeInhale);
                prog = (T)stimer / (T)(inh ? kc::INHALE_DUR : exd);
                const T sa = inh ? (T)kc::A_REST : (T)kc::A_FULL, sb = inh ? (T)kc::B_REST : (T)kc::B_FULL;
                const T da = inh ? (T)(kc::A_FULL - kc::A_REST) : (T)(kc::A_REST - kc::A_FULL);
                const T db = inh ? (T)(kc::B_FULL - kc::B_REST) : (T)(kc::B_REST - kc::B_FULL);
                ea = sa + da * prog; eb = sb + db * prog;
            }
            T radius = pymax<T>(ea, eb);
            if (wv_mode == 1) wv = prog;
            // ---- jet thrust through the steerable nozzle (R:261-314)
            if (thrust) {
                const T Tm = (T)100.0 * wv * (T)0.4;
                const T th = angle - nozzle;
                T s0, c0_;
                sincos_(th, &s0, &c0_);
                vx += (c0_ * Tm) * (T)0.012;
                vy += (s0 * Tm) * (T)0.012;
                const T primary = -nozzle * Tm * (T)0.0002;
                const T arm = radius * (T)0.7;
                const T perp = Tm * sin_(-nozzle);
                const T moment = perp * arm * (T)0.00005;
                const T shape = -nozzle * Tm * wv * (T)0.00003;
                omega += primary + moment + shape;
                const T sm = Tm * abs_(nozzle) * (T)0.3;
                T s1, c1_;
                if constexpr (is_f64<T>::value) sincos_(th + (T)kc::HALF_PI, &s1, &c1_);
                else { s1 = c0_; c1_ = -s0; }  // cos(x+pi/2) = -sin x, sin(x+pi/2) = cos x
                vx += (c1_ * sm) * (T)0.008;
                vy += (s1 * sm) * (T)0.008;
                cur_jit += 2;
                T u;  // np.random.random(), R:311
                if constexpr (is_f64<T>::value) u = res53(jw.x, jw.y);
                else u = (float)(jw.x >> 5) * (1.0f / 134217728.0f);
                const T dn = (u - (T)0.5) * (T)0.05;
                const T nf = Tm * (T)0.04;
                T s2, c2;
                if constexpr (is_f64<T>::value) sincos_(th + dn, &s2, &c2);
                else {  // sin/cos(th + dn) by angle addition; |dn| <= 0.025 so the series error is < 1e-13
                    const T d2 = dn * dn;
                    const T sd = dn * ((T)1.0 - d2 * ((T)(1.0 / 6.0) - d2 * (T)(1.0 / 120.0)));
                    const T cd = (T)1.0 - d2 * ((T)0.5 - d2 * (T)(1.0 / 24.0));
                    s2 = s0 * cd + c0_ * sd;
                    c2 = c0_ * cd - s0 * sd;
                }
                vx += c2 * nf * (T)0.002;
                vy += s2 * nf * (T)0.002;
            }
            if (wv_mode == 2) { const T w2 = wv * ((T)1.0 - prog); wv = (w2 > 0) ? w2 : (T)0; }
            else if (wv_mode == 3) wv = 0;
            // ---- drag, integrate, wrap, wall bounce (R:316-352)
            vx *= (T)0.98; vy *= (T)0.98; omega *= (T)0.95;
            px += vx; py += vy;
            angle = wrap_pi<T>(angle + omega);
            unsigned clamped = 0u;  // f32 build: walls the robot was clamped onto this step (1 low x / y, 2 right, 4 bottom)
            {
                const T margin = (T)kc::TANK_MARGIN + radius;
                if (px < margin) { px = margin; vx = abs_(vx) * (T)0.4; omega *= (T)0.7; clamped |= 1u; }
                else if (px > W - margin) { px = W - margin; vx = -abs_(vx) * (T)0.4; omega *= (T)0.7; clamped |= 2u; }
                if (py < margin) { py = margin; vy = abs_(vy) * (T)0.4; omega *= (T)0.7; clamped |= 1u; }
                else if (py > H - margin) { py = H - margin; vy = -abs_(vy) * (T)0.4; omega *= (T)0.7; clamped |= 4u; }
            }
            // ---- food distances (shared by hit test S:204-217, reward S:350-364 and obs S:375-379)
            T dx[kFastSel ? 1 : F], dy[kFastSel ? 1 : F], dd[kFastSel ? 1 : F];
            bool hit = false;
            bool any_food = false;
            T near_x = 0, near_y = 0, near_rel = 0;
            unsigned key[MU];   // kFastSel: sorted keys of the MU nearest live foods
            T dist_sum = 0;     // kFastSel: sum of live-food distances (slot order)
            // key = distance in 32-bit fixed point (diag -> 2^32, saturating) with the slot in the low bits:
            // unsigned order == (distance, slot) order at a resolution of diag / 2^(32 - slot bits)
            // (3.7e-6 px for 12 foods in the 800 x 600 tank, finer than one f32 ulp of the distance itself).
            // Eaten foods sort last.
            const float key_scale = a.f_inv_diag * (4294967296.0f * 0.9999f);
            auto make_key = [&](T d, int f) -> unsigned { return (__float2uint_rz((float)d * key_scale) & ~kSlotMask) | (unsigned)f; };
            auto insert_key = [&](unsigned c) {
    #pragma unroll
                for (int i = 0; i < MU; ++i) { const unsigned lo = min(key[i], c); c = max(key[i], c); key[i] = lo; }
            };
            // rare paths (a food was eaten / the env was reset): redo the selection from the (dx, dy) scratch
            auto reselect = [&]() {
                dist_sum = 0;
    #pragma unroll
                for (int i = 0; i < MU; ++i) key[i] = 0xffffffffu;
    #pragma unroll 1
                for (int f = 0; f < F; ++f) {
                    const float2 v = s_sel[f * BLOCK + tid];
                    const T d = sqrt_((T)v.x * (T)v.x + (T)v.y * (T)v.y);
                    const bool alive = (mask >> f) & 1u;
                    if (alive) dist_sum += d;
                    insert_key(alive ? make_key(d, f) : (~kSlotMask | (unsigned)f));
                }
            };
            if constexpr (kFastSel) {
                // one pass per food: offset, distance, hit bit, key insertion, scratch store -- nothing but the
                // MU keys, the hit bits and the distance sum survives it
                const T reach = radius + (T)kc::FOOD_R;
                unsigned hitbits = 0u;
    #pragma unroll
                for (int i = 0; i < MU; ++i) key[i] = 0xffffffffu;
    #pragma unroll
                for (int f = 0; f < F; ++f) {
                    const T ox = fv[f].x - px, oy = fv[f].y - py;
                    const T d = sqrt_(ox * ox + oy * oy);
                    s_sel[f * BLOCK + tid] = make_float2((float)ox, (float)oy);
                    const bool alive = (mask >> f) & 1u;
                    if (d < reach) hitbits |= 1u << f;
                    if (alive) dist_sum += d;
                    insert_key(alive ? make_key(d, f) : (~kSlotMask | (unsigned)f));
                }
                hitbits &= mask;  // S:204-217: the first live food within reach is eaten
                if (hitbits != 0u) {
                    hit = true;
                    mask &= ~(hitbits & (0u - hitbits));
                    reselect();
                }
            } else {
    #pragma unroll
                for (int f = 0; f < F; ++f) {
                    dx[f] = fx[f] - px; dy[f] = fy[f] - py;
                    dd[f] = sqrt_(dx[f] * dx[f] + dy[f] * dy[f]);
                }
                const T reach = radius + (T)kc::FOOD_R;
    #pragma unroll
                for (int f = 0; f < F; ++f)
                    if (!hit && ((mask >> f) & 1u) && dd[f] < reach) { hit = true; mask &= ~(1u << f); }
            }
            // ---- wall contact (S:219-230), rounding-critical: same operands/order as the reference
            bool collision;
            if constexpr (is_f64<T>::value) {
                collision = (px - radius <= (T)kc::TANK_MARGIN) || (px + radius >= W - (T)kc::TANK_MARGIN) ||
                            (py - radius <= (T)kc::TANK_MARGIN) || (py + radius >= H - (T)kc::TANK_MARGIN);
            } else {
                // The bounce above puts the robot EXACTLY on the line 50 + r, so whether that touch registers as a contact
                // is decided by the rounding of (50 + r) - r <= 50 in the reference's doubles (~5 % of the
                // touches bounce and survive).  A float comparison cannot reproduce that; the radius, however, is a pure
                // function of the integer shape fields, so on the (rare) clamp branch the reference's expression is
                // evaluated in double from them.  Away from a clamp the robot is strictly inside and nothing registers.
                collision = false;
                if (clamped != 0u) collision = (wall_touch_bits(skind, stimer, exd, a.W, a.H) & clamped) != 0u;
            }
            // ---- reward (S:278-327)
            T reward = 0;
            if (hit) {
                reward += c_food_reward<T>(a);
                if (c_efficiency_bonus<T>(a) > 0) reward += c_efficiency_bonus<T>(a) * (T)(a.max_steps - steps_since_food);
            }
            if (collision) reward += c_collision_penalty<T>(a);
            // nearest live food (strict <, first index wins: S:350-364).  Its bearing is used by the
            // alignment reward AND, when no food was eaten this step, is the first observation entry.
            if constexpr (kFastSel) {
                any_food = mask != 0u;
                const float2 nv = s_sel[(key[0] & kSlotMask) * BLOCK + tid];
                near_x = nv.x; near_y = nv.y;
            } else {
                T near_d = 0;
    #pragma unroll
                for (int f = 0; f < F; ++f)
                    if (((mask >> f) & 1u) && (!any_food || dd[f] < near_d)) { any_food = true; near_d = dd[f]; near_x = dx[f]; near_y = dy[f]; }
            }
            if (any_food && (c_proximity_w<T>(a) > 0 || a.observed > 0)) near_rel = wrap_pi<T>(atan2_(near_y, near_x) - angle);
            if (c_proximity_w<T>(a) > 0 && any_food) reward += c_proximity_w<T>(a) * cos_(near_rel);
            reward += c_time_penalty<T>(a);
            // ---- counters, respawn (S:171-180)
            ++steps_since_food;
            if (hit) {
                ++food_collected;
                steps_since_food = 0;
                if (a.respawn) {
                    const RespawnResult<T> r = respawn_slow<T, F>(a, e, px, py, mask, n_food, cur_food);
                    mask = r.mask; cur_food = r.cur;
                    if constexpr (kFastSel) {  // the new food joins the selection made before it existed
                        if (r.slot >= 0) {
                            const T ox = r.x - px, oy = r.y - py;
                            const T d = sqrt_(ox * ox + oy * oy);
                            s_sel[r.slot * BLOCK + tid] = make_float2((float)ox, (float)oy);
                            dist_sum += d;
                            insert_key(make_key(d, r.slot));
                        }
                    } else {
    #pragma unroll
                        for (int f = 0; f < F; ++f)
                            if (f == r.slot) {
                                fx[f] = r.x; fy[f] = r.y;
                                dx[f] = r.x - px; dy[f] = r.y - py;
                                dd[f] = sqrt_(dx[f] * dx[f] + dy[f] * dy[f]);
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
#endif
            ep_ret += reward;
            ++ep_len;

            // The observation is assembled once per substep -- and a second time, in warps where an episode
            // ended under auto-reset, from the freshly reset state (the row the caller sees is the new
            // episode's first observation; the terminal one goes to final_obs).  One copy of the code, two trips.
    #pragma unroll 1
            for (int pass = 0;; ++pass) {
            // ---- observation (R:377-388 + S:366-428), assembled in registers, written as 128-bit rows
            float4 v0, v1;
            float o8, o9;
            v0.x = (float)(px * invW); v0.y = (float)(py * invH);  // host reciprocals in both builds, see divc
            v0.z = (float)divc<T>(vx, 5.0); v0.w = (float)divc<T>(vy, 5.0);
            v1.x = (float)divc<T>(angle, kc::PI); v1.y = (float)divc<T>(omega, 0.1);
            v1.z = (float)divc<T>(radius, kc::BASE_R);
            v1.w = (float)((T)phase * (T)0.5);
            o8 = (float)wv;
            o9 = (float)divc<T>(nozzle, kc::MAX_NOZZLE);
            const int n_live = __popc(mask);
            // The MU nearest live foods in order (stable: equal distances keep slot order, S:382) and the
            // mean distance (S:419).
            T top_d[MU], top_x[MU], top_y[MU];
            T obs_sum = 0;
            if constexpr (F == 1) {
                top_d[0] = dd[0]; top_x[0] = dx[0]; top_y[0] = dy[0];
                obs_sum = dd[0];
            } else if constexpr (is_f64<T>::value) {
                // validation build: full sorting network on (distance, slot) so that the mean is summed
                // in sorted order exactly like sum() over the sorted list
                int id[F];
    #pragma unroll
                for (int f = 0; f < F; ++f) {
                    id[f] = f;
                    if (!((mask >> f) & 1u)) dd[f] = (T)CUDART_INF;
                }
    #define SALP_CE(i, j)                                                                          \
        {                                                                                          \
            const bool sw = (dd[j] < dd[i]) || (dd[j] == dd[i] && id[j] < id[i]);                  \
            const T t0 = sw ? dd[j] : dd[i], t1_ = sw ? dd[i] : dd[j];                             \
            const int u0 = sw ? id[j] : id[i], u1 = sw ? id[i] : id[j];                            \
            dd[i] = t0; dd[j] = t1_; id[i] = u0; id[j] = u1;                                       \
        }
                if constexpr (F == 2) { SALP_SORTNET_2(SALP_CE) }
                else if constexpr (F == 4) { SALP_SORTNET_4(SALP_CE) }
                else if constexpr (F == 8) { SALP_SORTNET_8(SALP_CE) }
                else if constexpr (F == 12) { SALP_SORTNET_12(SALP_CE) }
                else if constexpr (F == 16) { SALP_SORTNET_16(SALP_CE) }
                else if constexpr (F == 32) { SALP_SORTNET_32(SALP_CE) }
    #undef SALP_CE
    #pragma unroll
                for (int f = 0; f < F; ++f)
                    if (f < n_live) obs_sum = obs_sum + dd[f];
    #pragma unroll
                for (int i = 0; i < MU; ++i) {
                    T sx = 0, sy = 0;
    #pragma unroll
                    for (int f = 0; f < F; ++f)
                        if (id[i] == f) { sx = dx[f]; sy = dy[f]; }
                    top_d[i] = dd[i]; top_x[i] = sx; top_y[i] = sy;
                }
            } else {
                // throughput build: the MU nearest were selected on packed keys (select_top); fetch their offsets
                // by slot from the shared-memory scratch.  The mean is summed in slot order (differs from the
                // sorted-order sum by rounding only).
                obs_sum = dist_sum;
    #pragma unroll
                for (int i = 0; i < MU; ++i) {
                    const float2 v = s_sel[(key[i] & kSlotMask) * BLOCK + tid];
                    top_x[i] = v.x; top_y[i] = v.y;
                    top_d[i] = sqrt_(v.x * v.x + v.y * v.y);
                }
            }
            // i-th nearest live food -> [dx/W, dy/H, d/diag, wrap(atan2(dy,dx) - angle)/pi]; S:412 padding [0,0,1,0]
            auto food_entry = [&](int i) -> float4 {
                float4 q = make_float4(0.f, 0.f, 1.f, 0.f);
                if (i < MU && i < n_live) {
                    const int j = i < MU ? i : 0;
                    const T sx = top_x[j], sy = top_y[j], sd = top_d[j];
                    // entry 0 is the nearest food: its bearing was already computed for the reward unless
                    // a food was eaten (and possibly respawned) after that
                    T rel;
                    if (i == 0) rel = hit ? wrap_pi<T>(atan2_(sy, sx) - angle) : near_rel;
                    else rel = wrap_pi<T>(atan2_(sy, sx) - angle);
                    q.x = (float)(sx * invW); q.y = (float)(sy * invH); q.z = (float)(sd * invDiag);
                    q.w = (float)divc<T>(rel, kc::PI);
                }
                return q;
            };
            float cnt, avg = 1.f;
            cnt = (float)pymin<T>((T)n_live * (T)0.1, (T)1.0);
            if (n_live > 0) {
                const T mean = obs_sum / (T)n_live;
                avg = (float)(mean * invDiag);
            }
            const bool write_row = (STORE != kStoreDirect) || valid;
            if (write_row) {
                float4 *r4 = reinterpret_cast<float4 *>(row);
                r4[0] = v0; r4[1] = v1;
                if constexpr (MT == 3) {
                    const float4 e0 = food_entry(0), e1 = food_entry(1), e2 = food_entry(2);
                    r4[2] = make_float4(o8, o9, e0.x, e0.y);
                    r4[3] = make_float4(e0.z, e0.w, e1.x, e1.y);
                    r4[4] = make_float4(e1.z, e1.w, e2.x, e2.y);
                    r4[5] = make_float4(e2.z, e2.w, cnt, avg);
                } else {
                    float2 *r2 = reinterpret_cast<float2 *>(row);
                    r2[4] = make_float2(o8, o9);
    #pragma unroll
                    for (int i = 0; i < MT; ++i)
                        if (i < a.observed) {
                            const float4 q = food_entry(i);
                            r2[5 + 2 * i] = make_float2(q.x, q.y);
                            r2[6 + 2 * i] = make_float2(q.z, q.w);
                        }
                    r2[5 + 2 * a.observed] = make_float2(cnt, avg);
                }
            }

            if (kCoopReset && pass) break;
            // ---- per-step outputs
            const bool done = terminated || truncated;
            if (valid) {
                const size_t oi = (size_t)k * N + e;
                g_reward[oi] = reward;
                a.terminated[oi] = terminated ? 1 : 0;
                a.truncated[oi] = truncated ? 1 : 0;
                if (a.events) a.events[oi] = (uint8_t)((terminated ? 1 : 0) | (truncated ? 2 : 0) | (collision ? 4 : 0) | (hit ? 8 : 0));
            }
            // ---- episode statistics: warp-shuffle reduction, only in warps where an episode ended
            const unsigned done_mask = __ballot_sync(0xffffffffu, done);
            if (done_mask == 0u) break;
            if (a.stats) {
                const double r = warp_sum(done ? (double)ep_ret : 0.0);
                const double l = warp_sum(done ? (double)ep_len : 0.0);
                const double fc = warp_sum(done ? (double)food_collected : 0.0);
                const unsigned coll_mask = __ballot_sync(0xffffffffu, done && collision);
                const unsigned trunc_mask = __ballot_sync(0xffffffffu, done && truncated);
                const unsigned bad_mask = __ballot_sync(0xffffffffu, done && !(isfinite((double)px) && isfinite((double)py)));
                if (lane == 0) {
                    atomicAdd(a.stats + 0, (double)__popc(done_mask));
                    atomicAdd(a.stats + 1, r);
                    atomicAdd(a.stats + 2, l);
                    atomicAdd(a.stats + 3, fc);
                    if (coll_mask) atomicAdd(a.stats + 4, (double)__popc(coll_mask));
                    if (trunc_mask) atomicAdd(a.stats + 5, (double)__popc(trunc_mask));
                    if (bad_mask) atomicAdd(a.stats + 6, (double)__popc(bad_mask));
                }
            }
            if (done && a.final_info)
                reinterpret_cast<Vec4<T> *>(a.final_info)[(size_t)k * N + e] =
                    Vec4<T>{ep_ret, (T)ep_len, (T)food_collected, (T)steps_since_food};
            if (!a.autoreset) break;
            if (done && a.final_obs) {  // terminal observation (gymnasium "final_observation")
                float4 *dst = reinterpret_cast<float4 *>(a.final_obs + ((size_t)k * N + e) * OBS);
                for (int j = 0; j < OBS / 4; ++j) dst[j] = *reinterpret_cast<const float4 *>(row + 4 * j);
            }
            if constexpr (!kCoopReset) {
                // few foods: the in-thread slow path is cheap enough and keeps the hot path free of the loop
                if (done) {
                    const ResetResult rr = reset_slow<T>(a, e, cur_food, row);
                    cur_food = rr.cur; mask = rr.mask; n_food = rr.n_food;
                    px = W / 2; py = H / 2; vx = 0; vy = 0; angle = 0; omega = 0; nozzle = 0; wv = 0;
                    phase = kPhaseRest; timer = 0; skind = kShapeReset; stimer = 0;
                    steps_since_food = 0; food_collected = 0; ep_ret = 0; ep_len = 0;
                    if constexpr (!kFastSel) {
    #pragma unroll
                        for (int f = 0; f < F; ++f) {
                            if (f < n_food) {
                                const Vec2<T> v = g_food[(size_t)f * N + e];
                                fx[f] = v.x; fy[f] = v.y;
                            }
                        }
                    }
                }
                break;
            } else {
            // ---- auto-reset (S:133-155): the whole warp lays out the food of each finished env in turn
            for (unsigned m = done_mask; m != 0u; m &= m - 1u) {
                const int src = __ffs((int)m) - 1;
                const unsigned e_src = __shfl_sync(0xffffffffu, e, src);
                const LayoutResult<T> L = layout_warp<T>(a, e_src, __shfl_sync(0xffffffffu, cur_food, src), lane);
                if (lane < L.n_food) reinterpret_cast<Vec2<T> *>(a.food)[(size_t)lane * N + e_src] = Vec2<T>{L.x, L.y};
                if constexpr (!kFastSel) {
    #pragma unroll
                    for (int f = 0; f < F; ++f) {  // food f's position lives in lane f: hand it to the env's own lane
                        const T tx = __shfl_sync(0xffffffffu, L.x, f), ty = __shfl_sync(0xffffffffu, L.y, f);
                        if (lane == src && f < L.n_food) { fx[f] = tx; fy[f] = ty; }
                    }
                }
                if (lane == src) {
                    cur_food = L.cur; n_food = L.n_food;
                    mask = L.n_food >= 32 ? 0xffffffffu : ((1u << L.n_food) - 1u);
                }
            }
            if (done) {
                // R:95-117 deterministic robot reset, S:139-141 counters
                px = W / 2; py = H / 2; vx = 0; vy = 0; angle = 0; omega = 0; nozzle = 0; wv = 0;
                phase = kPhaseRest; timer = 0; skind = kShapeReset; stimer = 0;
                steps_since_food = 0; food_collected = 0; ep_ret = 0; ep_len = 0;
                radius = (T)kc::BASE_R;
                hit = true;  // the bearing computed for the reward belongs to the finished episode
            }
            if constexpr (kFastSel) {
                // second trip: offsets of the (new) foods from the (new) position, re-read from HBM where the
                // other lanes of this warp have just written them, then the selection again
                __syncwarp();
                if (done) {
                    Vec2<T> nf[F];  // all loads in flight together: this lane's latency is the tail of a one-wave grid
    #pragma unroll
                    for (int f = 0; f < F; ++f) nf[f] = g_food[(size_t)f * N + e];
    #pragma unroll
                    for (int f = 0; f < F; ++f) s_sel[f * BLOCK + tid] = make_float2((float)(nf[f].x - px), (float)(nf[f].y - py));
                    reselect();
                }
            } else {
                // distances for the second trip, for every lane (the validation build sorts dd[] in place)
    #pragma unroll
                for (int f = 0; f < F; ++f) {
                    dx[f] = fx[f] - px; dy[f] = fy[f] - py;
                    dd[f] = sqrt_(dx[f] * dx[f] + dy[f] * dy[f]);
                }
            }
            }
            }  // pass
            // ---- publish the observation tile
#ifndef SALP_LAB_NOOBS
            if constexpr (STORE == kStoreTmaBlock) {
                fence_proxy_async_smem();
                __syncthreads();
                const int n_rows = (int)((end - blk0) < (unsigned)BLOCK ? (end - blk0) : (unsigned)BLOCK);
                if (tid == 0) bulk_store_commit(a.obs + ((size_t)k * N + blk0) * OBS, s_buf, (uint32_t)(n_rows * OBS * sizeof(float)));
            } else if constexpr (STORE != kStoreDirect) {
                const unsigned w0 = blk0 + (tid & ~31);            // first env of this warp
                const int w_rows = w0 >= end ? 0 : (end - w0 >= 32u ? 32 : (int)(end - w0));
                float *g_tile = a.obs + ((size_t)k * N + w0) * OBS;
                if constexpr (STORE == kStoreTmaWarp) {
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0 && w_rows > 0) bulk_store_commit(g_tile, s_buf, (uint32_t)(w_rows * OBS * sizeof(float)));
                } else {
                    __syncwarp();
                    const int chunks = w_rows * (OBS / 4);  // 16-byte chunks, contiguous in HBM
                    const float4 *src = reinterpret_cast<const float4 *>(s_buf);
                    float4 *dst = reinterpret_cast<float4 *>(g_tile);
                    for (int c = lane; c < chunks; c += 32) dst[c] = src[c];
                }
            }
#endif
        }

        // ---- write state back once (food records are only written by the slow paths)
        if (valid) {
            g_body[e] = Vec4<T>{px, py, vx, vy};
            g_body[(size_t)N + e] = Vec4<T>{angle, omega, nozzle, wv};
            const int fsm = phase | (timer << 2) | (exd << 10) | (skind << 18) | (stimer << 20);
            a.ctrl[e] = make_int4(fsm, steps_since_food, food_collected, (int)mask);
            a.ctrl[(size_t)N + e] = make_int4((int)cur_food, (int)cur_jit, (ep_len < kEpLenMask ? ep_len : kEpLenMask) | (n_food << 26), (int)jcache);
            g_epret[e] = ep_ret;
        }

    }  // tiles
    if (blockIdx.x == 0 && tid == 0 && a.stats) atomicAdd(a.stats + 7, (double)a.count * (double)a.K);
    if constexpr (kTma) {  // smem must stay allocated until the last bulk store has read it
        if (STORE == kStoreTmaBlock ? (tid == 0) : (lane == 0)) bulk_wait_read<0>();
    }
}

// ---------------------------------------------------------------------------------------------
// Reset kernel: SalpSnakeEnv.reset for masked envs (S:133-155); writes obs[0][e][:]
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) salp_reset_kernel(const __grid_constant__ KernelArgs a) {
    const long long e = a.first + (long long)blockIdx.x * 256 + threadIdx.x;
    if (e >= a.first + a.count) return;
    if (a.reset_mask && a.reset_mask[e] == 0) return;
    const long long N = a.N;
    Vec4<T> *g_body = reinterpret_cast<Vec4<T> *>(a.body);
    const int4 c1 = a.ctrl[N + e];
    const unsigned cur_food = a.rewind_rng ? 0u : (unsigned)c1.x;
    const unsigned cur_jit = a.rewind_rng ? 0u : (unsigned)c1.y;
    const ResetResult rr = reset_slow<T>(a, e, cur_food, a.obs ? a.obs + (size_t)e * a.obs_dim : nullptr);
    g_body[e] = Vec4<T>{(T)a.W / 2, (T)a.H / 2, 0, 0};
    g_body[N + e] = Vec4<T>{0, 0, 0, 0};
    const int fsm = kPhaseRest | (kShapeReset << 18);
    a.ctrl[e] = make_int4(fsm, 0, 0, (int)rr.mask);
    const unsigned trow = (unsigned)e % a.rows;
    const uint32_t jcache = a.rng_mode ? 0u : a.jit_words[(size_t)trow * a.jit_len + (cur_jit & a.jit_len_mask)];
    a.ctrl[N + e] = make_int4((int)rr.cur, (int)cur_jit, rr.n_food << 26, (int)jcache);
    reinterpret_cast<T *>(a.ep_return)[e] = 0;
}

// ---------------------------------------------------------------------------------------------
// Observation of the CURRENT state, no step (salp_observe): S:366-428 _get_extended_observation over
// R:371-388 _get_observation -- what the reference's callers get after poking robot_pos / robot_angle /
// food_positions (eval/collect_navigation_data.py:76-89, eval/watch_navigation.py:63).  Generic in the
// food count and in max_observed_food; off the hot path (one thread per env, live foods read from HBM,
// stable insertion sort exactly like list.sort on (distance, position) tuples with distinct distances).
// ---------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ void shape_axes(int fsm, T &ea, T &eb) {
    const int exd = (fsm >> 10) & 0xff, skind = (fsm >> 18) & 3, stimer = (fsm >> 20) & 0xff;
    if (skind == kShapeReset) { ea = (T)kc::BASE_R; eb = (T)kc::BASE_R; }            // R:111-113
    else if (skind == kShapeRest) { ea = (T)kc::A_REST; eb = (T)kc::B_REST; }       // R:190-191
    else {                                                                          // R:207-213, 240-246
        const bool inh = (skind == kShapeInhale);
        const T prog = (T)stimer / (T)(inh ? kc::INHALE_DUR : (exd > 0 ? exd : 1));
        const T sa = inh ? (T)kc::A_REST : (T)kc::A_FULL, sb = inh ? (T)kc::B_REST : (T)kc::B_FULL;
        const T da = inh ? (T)(kc::A_FULL - kc::A_REST) : (T)(kc::A_REST - kc::A_FULL);
        const T db = inh ? (T)(kc::B_FULL - kc::B_REST) : (T)(kc::B_REST - kc::B_FULL);
        ea = sa + da * prog; eb = sb + db * prog;
    }
}

template <typename T>
__global__ void __launch_bounds__(128) salp_observe_kernel(const __grid_constant__ KernelArgs a) {
    const long long e = a.first + (long long)blockIdx.x * 128 + threadIdx.x;
    if (e >= a.first + a.count) return;
    const long long N = a.N;
    const Vec4<T> b0 = reinterpret_cast<const Vec4<T> *>(a.body)[e], b1 = reinterpret_cast<const Vec4<T> *>(a.body)[N + e];
    const int4 c0 = a.ctrl[e], c1 = a.ctrl[N + e];
    const Vec2<T> *food = reinterpret_cast<const Vec2<T> *>(a.food);
    const T px = b0.x, py = b0.y, angle = b1.x;
    const T W = (T)a.W, H = (T)a.H, diag = (T)a.diag;
    T ea, eb;
    shape_axes<T>(c0.x, ea, eb);
    float *o = a.obs + (size_t)e * a.obs_dim;
    o[0] = (float)(px / W); o[1] = (float)(py / H);
    o[2] = (float)(b0.z / (T)5.0); o[3] = (float)(b0.w / (T)5.0);
    o[4] = (float)(angle / (T)kc::PI); o[5] = (float)(b1.y / (T)0.1);
    o[6] = (float)(pymax<T>(ea, eb) / (T)kc::BASE_R);
    o[7] = (float)((T)(c0.x & 3) / (T)2.0);
    o[8] = (float)b1.w;
    o[9] = (float)(b1.z / (T)kc::MAX_NOZZLE);
    int n_food = (int)((unsigned)c1.z >> 26);
    if (n_food > kMaxFood) n_food = kMaxFood;
    const unsigned mask = (unsigned)c0.w;
    T dist[kMaxFood];
    int idx[kMaxFood];
    int n = 0;
    for (int j = 0; j < n_food; ++j)
        if ((mask >> j) & 1u) {
            const Vec2<T> f = food[(size_t)j * N + e];
            const T dx = f.x - px, dy = f.y - py;
            const T d = sqrt_(dx * dx + dy * dy);
            int k = n - 1;  // stable insertion: equal distances keep slot order
            while (k >= 0 && dist[k] > d) { dist[k + 1] = dist[k]; idx[k + 1] = idx[k]; --k; }
            dist[k + 1] = d; idx[k + 1] = j;
            ++n;
        }
    float *fo = o + 10;
    for (int i = 0; i < a.observed; ++i) {
        if (i < n) {
            const Vec2<T> f = food[(size_t)idx[i] * N + e];
            const T dx = f.x - px, dy = f.y - py;
            const T rel = wrap_pi<T>(atan2_(dy, dx) - angle);
            fo[0] = (float)(dx / W); fo[1] = (float)(dy / H);
            fo[2] = (float)(dist[i] / diag); fo[3] = (float)(rel / (T)kc::PI);
        } else { fo[0] = 0.f; fo[1] = 0.f; fo[2] = 1.f; fo[3] = 0.f; }
        fo += 4;
    }
    fo[0] = (float)pymin<T>((T)n / (T)10.0, (T)1.0);
    if (n > 0) {
        T s = 0;
        for (int i = 0; i < n; ++i) s = s + dist[i];
        fo[1] = (float)((s / (T)n) / diag);
    } else fo[1] = 1.f;
}

// ctrl rec1.w = first jitter word at every env's jitter cursor (salp_refresh_rng)
static __global__ void __launch_bounds__(256) salp_refresh_rng_kernel(const __grid_constant__ KernelArgs a) {
    const long long e = a.first + (long long)blockIdx.x * 256 + threadIdx.x;
    if (e >= a.first + a.count) return;
    int4 c1 = a.ctrl[a.N + e];
    const unsigned trow = (unsigned)e % a.rows;
    c1.w = a.rng_mode ? 0 : (int)a.jit_words[(size_t)trow * a.jit_len + ((unsigned)c1.y & a.jit_len_mask)];
    a.ctrl[a.N + e] = c1;
}

}  // namespace salp
