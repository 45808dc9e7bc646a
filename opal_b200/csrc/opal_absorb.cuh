// opal_absorb.cuh -- device functions of the photon-electron pair processes: Airy Ai on the positive axis
// (qed/special_functions/airy.rs:19-69), the one-photon absorption and stimulated-emission cross sections
// (qed/photon_absorption.rs:17-35, qed/stimulated_emission.rs:17-37).
//
// Evaluated in the reference's operation order and compiled with --fmad=false: what differs from the CPU statement is
// pow/exp (CUDA libm, <= 2 ulp from glibc), far inside the 1e-12 the reference's own Airy tests ask for.
// The quadrature tables sit in constant memory: every lane of a warp walks the same node index at the same time.
#pragma once
#include "opal_types.h"

namespace opal {
namespace pairs {

#include "airy_tables.h"      // generated: tools/gen_airy_tables.py

__constant__ double c_series_a[14] = {AIRY_SERIES_A_LIST};
__constant__ double c_gl_x[60] = {AIRY_GL40_X_LIST, AIRY_GL16_X_LIST, AIRY_GL4_X_LIST};      // 40 | 16 | 4 nodes
__constant__ double c_gl_w[60] = {AIRY_GL40_W_LIST, AIRY_GL16_W_LIST, AIRY_GL4_W_LIST};

constexpr double AIRY_SCALE = 0.262183997088323;            // airy.rs:40 (the reference's 15-digit literal of 1/(sqrt(pi) 48^(1/6) Gamma(5/6)))
constexpr double CLASSICAL_ELECTRON_RADIUS = 2.817940e-15;  // constants.rs:28
constexpr double ALPHA_FINE = 7.29735257e-3;                // constants.rs:24
constexpr double PI_ = 3.14159265358979323846;

// f64::powi (llvm.powi -> __powidf2): square-and-multiply from the low bit
__device__ __forceinline__ double powi(double a, int b) {
    double r = 1.0;
    for (;;) { if (b & 1) r *= a; b >>= 1; if (b == 0) break; a *= a; }
    return r;
}

__device__ inline double airy_quadrature(double x, int first, int n) {       // airy.rs:39-45
    const double s = 2.0 * pow(x, 1.5) / 3.0;
    const double a = AIRY_SCALE * pow(s, -1.0 / 6.0) * exp(-s);
    double integral = 0.0;
    for (int i = 0; i < n; i++) integral += c_gl_w[first + i] * pow(2.0 + c_gl_x[first + i] / s, -1.0 / 6.0);
    return a * integral;
}

// false where the reference returns None: x < 0, x >= 50 (Ai(50) < 4.5e-104) or NaN
__device__ inline bool airy_ai_for_positive(double x, double* ai) {
    if (x < 0.0) return false;
    if (x < 1.0) {
        double sum = 0.0;
        for (int i = 0; i < 14; i++) sum += c_series_a[i] * powi(x, (i >> 1) * 3 + (i & 1));     // powers 0,1,3,4,6,7,...
        *ai = sum;
    } else if (x < 2.0) *ai = airy_quadrature(x, 0, 40);
    else if (x < 10.0) *ai = airy_quadrature(x, 40, 16);
    else if (x < 50.0) *ai = airy_quadrature(x, 56, 4);
    else return false;
    return true;
}

// sign = +1: absorption, -1: stimulated emission.  k, p: four-momenta / mc of the photon and the electron.
template <int SIGN>
__device__ inline bool pair_cross_section(const double k[4], const double p[4], double chi_gamma, double chi_e, double* sigma) {
    if (SIGN < 0 && (chi_gamma >= chi_e || k[0] >= p[0])) return false;       // an electron cannot emit more than it has
    if (chi_e <= 0.0 || chi_gamma <= 0.0) return false;
    const double den = chi_e * (SIGN > 0 ? chi_e + chi_gamma : chi_e - chi_gamma);
    const double g = 0.5 + 0.25 * (chi_gamma * chi_gamma) / den;
    const double z = pow(chi_gamma / den, 2.0 / 3.0);
    const double k_p = k[0] * p[0] - k[1] * p[1] - k[2] * p[2] - k[3] * p[3];
    const double zbar = 2.0 * z * chi_e * k_p / chi_gamma;
    const double zbar_z = 2.0 * p[0] * k_p / k[0];                              // "guarantees positivity of cross section"
    double ai;
    if (!airy_ai_for_positive(zbar, &ai)) return false;
    const double r = 2.0 * PI_ * CLASSICAL_ELECTRON_RADIUS;
    *sigma = (r * r) * chi_e * z * (4.0 * g * zbar_z - 1.0) * ai / (ALPHA_FINE * chi_gamma * k[0] * p[0]);
    return true;
}

// ---- the same cross sections for 32 pairs at once, the Airy quadratures shared out over the warp ------------------------
// One lane per pair would leave most of the warp idle: the pairs of a round fall into different ranges of the Airy
// evaluation (series, 40-, 16-, 4-node rule) and the four branches would run one after the other.  Here every lane
// prepares its pair up to the Airy argument; then, rule by rule, the warp evaluates the nodes of the pairs that need that
// rule side by side -- 32 / N pairs per pass for the 16- and 4-node rules, one pair (32 nodes) plus a quarter pass (its
// 8 remaining nodes) for the 40-node rule -- and sums each group with a butterfly.  The node sum is thus associated
// differently from the reference's sequential one: a few 1e-16 on Ai, inside the 1e-12 its own tests ask for.
struct PairPre {
    bool ok;            // the branch conditions hold (the cross section can still be None through the Airy range)
    double zbar;        // the Airy argument
    double factor;      // (2 pi r_e)^2 chi_e z (4 g zbar_z - 1)
    double denom;       // alpha chi_gamma k0 p0
};

// sign = +1: absorption, -1: stimulated emission (a run-time argument: ONE copy of the code, see airy_warp)
__device__ __forceinline__ PairPre pair_prepare(int sign, const double k[4], const double p[4], double chi_gamma, double chi_e) {
    PairPre r = {false, 0.0, 0.0, 1.0};
    if (sign < 0 && (chi_gamma >= chi_e || k[0] >= p[0])) return r;
    if (chi_e <= 0.0 || chi_gamma <= 0.0) return r;
    const double den = chi_e * (sign > 0 ? chi_e + chi_gamma : chi_e - chi_gamma);
    const double g = 0.5 + 0.25 * (chi_gamma * chi_gamma) / den;
    const double z = pow(chi_gamma / den, 2.0 / 3.0);
    const double k_p = k[0] * p[0] - k[1] * p[1] - k[2] * p[2] - k[3] * p[3];
    const double zbar_z = 2.0 * p[0] * k_p / k[0];
    const double rr = 2.0 * PI_ * CLASSICAL_ELECTRON_RADIUS;
    r.ok = true;
    r.zbar = 2.0 * z * chi_e * k_p / chi_gamma;
    r.factor = (rr * rr) * chi_e * z * (4.0 * g * zbar_z - 1.0);
    r.denom = ALPHA_FINE * chi_gamma * k[0] * p[0];
    return r;
}

// every lane of `mask` gets sum_{i < gs} w[first + i] (2 + x[first + i] / s)^(-1/6) for ITS s; the warp works on 32 / gs
// of them per pass (gs: 32, 16, 8 or 4).  sx / sw: the node tables in shared memory (lanes read different nodes).
// Warp-uniform call.
__device__ __forceinline__ double quadrature_groups(unsigned mask, double s, int first, int gs, const double* sx, const double* sw, int lane) {
    const int per_pass = 32 / gs;
    const int cnt = __popc(mask);
    const int rank = __popc(mask & ((1u << lane) - 1u));
    const bool member = (mask >> lane) & 1u;
    const int i = lane & (gs - 1), sub = lane / gs;
    double result = 0.0;
#pragma unroll 1
    for (int q0 = 0; q0 < cnt; q0 += per_pass) {
        const int q = q0 + sub;
        const unsigned src = __fns(mask, 0, q + 1);                 // the lane of member number q (none: 0xffffffff)
        const double sq = __shfl_sync(0xffffffffu, s, (int)(src & 31u));
        double term = q < cnt ? sw[first + i] * pow(2.0 + sx[first + i] / sq, -1.0 / 6.0) : 0.0;
        for (int o = gs >> 1; o > 0; o >>= 1) term += __shfl_xor_sync(0xffffffffu, term, o);
        const double got = __shfl_sync(0xffffffffu, term, ((rank - q0) * gs) & 31);
        if (member && rank >= q0 && rank < q0 + per_pass) result = got;
    }
    return result;
}

// Ai(x) for the lanes with `want` set; false where the reference returns None.  Warp-uniform call.  The four node sets
// (40 = 32 + 8, 16, 4) go through ONE copy of the pow() loop: with a copy per rule and per cross section the kernel
// outgrew the instruction cache and stalled on fetches two thirds of the time (ncu: stalled_no_instructions).
__device__ __noinline__ bool airy_warp(bool want, double x, const double* sx, const double* sw, int lane, double* ai) {
    const int rule = !want || !(x >= 0.0) ? 0 : (x < 1.0 ? 1 : (x < 2.0 ? 40 : (x < 10.0 ? 16 : (x < 50.0 ? 4 : 0))));
    double s = 1.0, a = 0.0;
    if (rule > 1) {
        s = 2.0 * pow(x, 1.5) / 3.0;
        a = AIRY_SCALE * pow(s, -1.0 / 6.0) * exp(-s);
    }
    double integral = 0.0;
#pragma unroll 1
    for (int part = 0; part < 4; part++) {
        const int of_rule = part < 2 ? 40 : (part == 2 ? 16 : 4);
        const int first = part == 0 ? 0 : (part == 1 ? 32 : (part == 2 ? 40 : 56));
        const int gs = part == 0 ? 32 : (part == 1 ? 8 : (part == 2 ? 16 : 4));
        const unsigned mask = __ballot_sync(0xffffffffu, rule == of_rule);
        if (!mask) continue;
        const double v = quadrature_groups(mask, s, first, gs, sx, sw, lane);
        if (rule == of_rule) integral += v;
    }
    if (rule == 1) {
        double sum = 0.0;
        for (int i = 0; i < 14; i++) sum += c_series_a[i] * powi(x, (i >> 1) * 3 + (i & 1));
        *ai = sum;
    } else *ai = a * integral;
    return rule != 0;
}

}  // namespace pairs
}  // namespace opal
