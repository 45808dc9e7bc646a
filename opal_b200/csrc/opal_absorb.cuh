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

}  // namespace pairs
}  // namespace opal
