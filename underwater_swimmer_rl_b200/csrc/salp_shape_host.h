// salp_shape_host.h -- model constants and the two pieces of reference arithmetic that BOTH the kernels and the host code
// of the C ABI evaluate in double: the ellipse axes from the integer shape fields, and the rounding-critical wall-contact
// test after a bounce.  R = scripts/utilities/salp_robot.py, S = src/salp/environments/salp_snake_env.py.
#pragma once
#include "salp_internal.h"

#ifndef __CUDACC__
#ifndef __host__
#define __host__
#define __device__
#endif
#endif

namespace salp {

// Model constants, evaluated in double exactly as the Python expressions are (R:34-53, S:53-54).
namespace kc {
constexpr double PI = 3.141592653589793;   // math.pi
constexpr double TWO_PI = 2 * PI;
constexpr double MAX_NOZZLE = PI / 3;      // R:47
constexpr double HALF_PI = PI / 2;         // R:301
constexpr double TANK_MARGIN = 50.0, BASE_R = 30.0, FOOD_R = 15.0, MIN_FOOD_DIST = 80.0;
constexpr double A_REST = BASE_R * 1.3, B_REST = BASE_R * 0.8;  // R:190-191
constexpr double A_FULL = BASE_R * 1.1, B_FULL = BASE_R * 1.1;  // R:209-210
constexpr double NOZZLE_RATE = 0.05;       // R:48
constexpr int INHALE_DUR = 120;            // R:51
}  // namespace kc

__host__ __device__ inline double dmul_rn(double x, double y) {
#ifdef __CUDA_ARCH__
    return __dmul_rn(x, y);
#else
    volatile double p = x * y;  // keep the host compiler from contracting as well
    return p;
#endif
}
__host__ __device__ inline double dadd_rn(double x, double y) {
#ifdef __CUDA_ARCH__
    return __dadd_rn(x, y);
#else
    return x + y;
#endif
}

// Does a clamp onto the low (x or y) / right / bottom wall register as a wall contact?  R:335-352 clamps the position to
// m = 50 + r resp. W - m, H - m and S:225-228 then tests pos - r <= 50 resp. pos + r >= W - 50 -- in doubles, so the answer
// is a rounding property of r = max(ellipse_a, ellipse_b), itself a function of the integer shape fields of the fsm word.
// bit 0: low wall, bit 1: right wall, bit 2: bottom wall.  Double arithmetic in the reference's operation order.
__host__ __device__ inline unsigned wall_touch_bits_r(double r, double W, double H) {
    const double m = kc::TANK_MARGIN + r;
    unsigned bits = 0u;
    if (m - r <= kc::TANK_MARGIN) bits |= 1u;
    if ((W - m) + r >= W - kc::TANK_MARGIN) bits |= 2u;
    if ((H - m) + r >= H - kc::TANK_MARGIN) bits |= 4u;
    return bits;
}
__host__ __device__ inline void shape_axes_f64(int skind, int stimer, int exd, double &ea, double &eb) {
    if (skind == kShapeReset) { ea = kc::BASE_R; eb = kc::BASE_R; }
    else if (skind == kShapeRest) { ea = kc::A_REST; eb = kc::B_REST; }
    else {
        const bool inh = (skind == kShapeInhale);
        const double prog = (double)stimer / (double)(inh ? kc::INHALE_DUR : exd);
        // a + (b - a) * p with the product rounded BEFORE the sum (the reference's two operations): in device code of the
        // f32 translation unit nvcc would otherwise contract it into one DFMA and change the last bit of the radius
        ea = dadd_rn((inh ? kc::A_REST : kc::A_FULL), dmul_rn((inh ? (kc::A_FULL - kc::A_REST) : (kc::A_REST - kc::A_FULL)), prog));
        eb = dadd_rn((inh ? kc::B_REST : kc::B_FULL), dmul_rn((inh ? (kc::B_FULL - kc::B_REST) : (kc::B_REST - kc::B_FULL)), prog));
    }
}
}  // namespace salp
