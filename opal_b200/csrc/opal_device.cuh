// opal_device.cuh -- device-side arithmetic of the opal PIC hot path (sm_100a).
//
// Every function cites the reference statement it implements.  Operation ORDER follows the
// reference so that the strict build (--fmad=false) reproduces the CPU oracle's IEEE results
// bit-for-bit for + - * / sqrt; only libm calls (log/exp/cbrt/pow/sincos/acos) differ by <= 2 ulp.
#pragma once
#include <cstdint>
#include <math_constants.h>

namespace opal {

// ---- src/constants.rs:4-28 ---------------------------------------------------------------
#define OPAL_C 2.997925e8
#define OPAL_C_SQD 89875517873681764.0
#define OPAL_EPS0 8.854188e-12
#define OPAL_MU0 1.256637e-6
#define OPAL_QE (-1.602177e-19)
#define OPAL_E (1.602177e-19)
#define OPAL_ME 9.109383e-31
#define OPAL_MP 1.672622e-27
#define OPAL_ME_MEV 0.510999
#define OPAL_ECRIT 1.323285e18
#define OPAL_ALPHA_FINE 7.29735257e-3
#define OPAL_COMPTON_TIME 1.28808867e-21
#define OPAL_PI 3.14159265358979323846264338327950288
#define OPAL_FRAC_PI_3 1.04719755119659774615421446109316763

struct V3 { double x, y, z; };
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 operator*(double s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ V3 operator/(V3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

// ---- Cell::weight  yee.rs:140-149 ------------------------------------------------------
__device__ __forceinline__ double weight(double xi) {
    double xhat = fabs(xi);
    double inner = 0.75 - xhat * xhat;
    double outer = 1.125 - 1.5 * xhat + 0.5 * (xhat * xhat);
    double w = xhat < 0.5 ? inner : outer;
    return xhat > 1.5 ? 0.0 : w;
}

// ---- Cell::flux  yee.rs:185-204 ----------------------------------------------------------
__device__ __forceinline__ double flux(double x_i, double x_f) {
    double ai = fabs(x_i), af = fabs(x_f);
    if (ai < 1.0) {
        if (af >= 1.0) {
            double t = 1.0 - ai;
            return copysign(0.5 * (t * t), -x_i);
        } else if (x_i * x_f >= 0.0) {
            double tf = 1.0 - af, ti = 1.0 - ai;
            return copysign(0.5 * (tf * tf) - 0.5 * (ti * ti), x_i - x_f);
        } else {
            return copysign(ai * (1.0 - 0.5 * ai) + af * (1.0 - 0.5 * af), x_i);
        }
    } else if (af < 1.0) {
        double t = 1.0 - af;
        return copysign(0.5 * (t * t), x_f);
    }
    return 0.0;
}

// The 18 field values a particle in cell j can see (yee.rs:499-529): centre-staggered components
// from cells j-1..j+1, node components from j-1..j+2.  One copy per warp (all lanes share j).
struct CellFields {
    double ex[3], ey[4], ez[4], bx, by[3], bz[3];
};

// ---- YeeGrid::fields_at  yee.rs:499-529 --------------------------------------------------
__device__ __forceinline__ void fields_at(const CellFields& f, double xi, V3& E, V3& B) {
    double wc0 = weight(0.5 + xi), wc1 = weight(0.5 - xi), wc2 = weight(1.5 - xi);
    double wn0 = weight(1.0 + xi), wn1 = weight(xi), wn2 = weight(1.0 - xi), wn3 = weight(2.0 - xi);
    E.x = f.ex[0] * wc0 + f.ex[1] * wc1 + f.ex[2] * wc2;
    E.y = f.ey[0] * wn0 + f.ey[1] * wn1 + f.ey[2] * wn2 + f.ey[3] * wn3;
    E.z = f.ez[0] * wn0 + f.ez[1] * wn1 + f.ez[2] * wn2 + f.ez[3] * wn3;
    B.x = f.bx;
    B.y = f.by[0] * wc0 + f.by[1] * wc1 + f.by[2] * wc2;
    B.z = f.bz[0] * wc0 + f.bz[1] * wc1 + f.bz[2] * wc2;
}

// ---- photon_emission::rate  qed/photon_emission.rs:59-79 --------------------------------
// ln_h: the 41 log h(chi) samples (photon_emission.rs:14-56, second column), staged in shared memory
// by the calling kernel.
#define OPAL_LN_H_X0 (-4.60517019e+0)
#define OPAL_DELTA_LN_CHI 0.230258509299
__device__ __forceinline__ double qed_rate(double chi, double gamma, const double* __restrict__ ln_h) {
    double h;
    if (chi < 0.01) {
        h = 5.0 * OPAL_FRAC_PI_3 * (1.0 - 8.0 * chi / (5.0 * sqrt(3.0)));
    } else if (chi >= 100.0) {
        double cb = cbrt(chi);
        h = -1019.4661473121777 + 1786.716527650374 * cb * cb;
        h = 1750.6263395722715 + cb * cb * h;
        h = -2260.1819695887225 + cb * h;
        h = 0.00296527643253334 * h / (chi * chi);
    } else {
        double index = (log(chi) - OPAL_LN_H_X0) / OPAL_DELTA_LN_CHI;
        double w = index - trunc(index);
        int i = (int)floor(index);
        i = i > 39 ? 39 : i;
        double ln_h_v = (1.0 - w) * ln_h[i] + w * ln_h[i + 1];
        h = exp(ln_h_v);
    }
    return sqrt(3.0) * OPAL_ALPHA_FINE * chi * h / (2.0 * OPAL_PI * gamma * OPAL_COMPTON_TIME);
}
__device__ __forceinline__ double qed_classical_rate(double chi, double gamma) {   // :82-85
    double h = 5.0 * OPAL_FRAC_PI_3;
    return sqrt(3.0) * OPAL_ALPHA_FINE * chi * h / (2.0 * OPAL_PI * gamma * OPAL_COMPTON_TIME);
}

// ---- position / cell update shared by all pushers (electron.rs:311-329) ---------------------
// returns the cell delta in {-1, 0, +1}; sets cfl = true when the reference's assert!(dxi < 1.0) fails
__device__ __forceinline__ int move_x(double& x, double& prev_x, double dxi, bool& cfl) {
    prev_x = x;
    if (!(dxi < 1.0)) cfl = true;
    x = x + dxi;
    double fl = floor(x);
    int d = fl < 0.0 ? -1 : (fl > 0.0 ? 1 : 0);
    prev_x -= fl;
    x -= fl;
    return d;
}

// Full per-particle state of a charged macroparticle (the reference's Electron / Ion fields that
// evolve).  For ions `gamma` holds gamma - 1 (ion.rs:83-85).
struct PState {
    double x, prev_x, w;
    V3 u;
    double gamma, chi, tau, work, y, z;
};

// ---- Electron::vay_push  electron.rs:268-330 ----------------------------------------------
// FULL: maintain work, chi, tau, y, z exactly as the reference does every step (quirk Q4).
// !FULL (lean layout): gamma is recomputed from u; work/chi/tau/y/z are not maintained.
template <bool FULL>
__device__ __forceinline__ int electron_push(PState& p, V3 E, V3 B, double dx, double dt, uint32_t flags,
                                             const double* __restrict__ ln_h, bool& cfl) {
    if (!FULL) p.gamma = sqrt(1.0 + dot(p.u, p.u));
    V3 v = (OPAL_C * p.u) / p.gamma;
    const double alpha = OPAL_QE * dt / (2.0 * OPAL_ME * OPAL_C);
    V3 u_half = p.u + alpha * (E + cross(v, B));
    if (FULL) {
        p.work += OPAL_QE * OPAL_C * dot(u_half, E) * dt / sqrt(1.0 + dot(u_half, u_half));
        double gamma_half = sqrt(1.0 + dot(u_half, u_half));
        double eu = dot(E, u_half);
        V3 f = gamma_half * E + OPAL_C * cross(u_half, B);
        p.chi = sqrt(dot(f, f) - eu * eu) / OPAL_ECRIT;
        double r = (flags & 1u) ? qed_classical_rate(p.chi, gamma_half) : qed_rate(p.chi, gamma_half, ln_h);
        p.tau = p.tau - r * dt;
    }
    V3 u_prime = u_half + alpha * E;
    double gamma_prime_sqd = 1.0 + dot(u_prime, u_prime);
    V3 tau = (alpha * OPAL_C) * B;
    double u_star = dot(u_prime, tau);
    double sigma = gamma_prime_sqd - dot(tau, tau);
    p.gamma = sqrt(0.5 * sigma + sqrt(0.25 * (sigma * sigma) + dot(tau, tau) + u_star * u_star));
    V3 t = tau / p.gamma;
    double s = 1.0 / (1.0 + dot(t, t));
    p.u = s * (u_prime + dot(u_prime, t) * t + cross(u_prime, t));
    double dxi = OPAL_C * p.u.x * dt / (dx * p.gamma);
    int d = move_x(p.x, p.prev_x, dxi, cfl);
    if (FULL) {
        p.y = p.y + v.y * dt;     // quirk Q3: pre-push velocity
        p.z = p.z + v.z * dt;
    }
    return d;
}

// ---- Ion::push (Boris)  ion.rs:168-214 ------------------------------------------------------
template <bool FULL>
__device__ __forceinline__ int ion_push(PState& p, V3 E, V3 B, double dx, double dt, double z_star, double a, bool& cfl) {
    V3 cB = OPAL_C * B;
    double q = z_star * OPAL_E;
    double M = a * OPAL_MP;
    double alpha = q * dt / (2.0 * M * OPAL_C);
    V3 u_minus = p.u + alpha * E;
    double um2 = dot(u_minus, u_minus);
    double gamma = 1.0 + um2 / (1.0 + sqrt(1.0 + um2));
    double t = alpha / gamma;
    V3 u_prime = u_minus + t * cross(u_minus, cB);
    double t_prime = 2.0 * t / (1.0 + (t * t) * dot(cB, cB));
    V3 u_plus = u_minus + t_prime * cross(u_prime, cB);
    p.u = u_plus + alpha * E;
    double u2 = dot(p.u, p.u);
    p.gamma = u2 / (1.0 + sqrt(1.0 + u2));   // gamma - 1
    double dxi = OPAL_C * p.u.x * dt / (dx * (1.0 + p.gamma));
    return move_x(p.x, p.prev_x, dxi, cfl);
}

// velocity used by the deposit: Particle::velocity (electron.rs:148-151, ion.rs:123-126)
__device__ __forceinline__ V3 electron_velocity(const PState& p) { return (OPAL_C * p.u) / p.gamma; }
__device__ __forceinline__ V3 ion_velocity(const PState& p) { return (OPAL_C * p.u) / (1.0 + p.gamma); }

// ---- Photon::push  photon.rs:150-183 --------------------------------------------------------
struct PhState { double x, prev_x, w; V3 k; double chi, y, z; };
__device__ __forceinline__ int photon_push(PhState& p, V3 E, V3 B, double dx, double dt, bool& cfl) {
    double k0 = sqrt(dot(p.k, p.k));
    V3 v = (OPAL_C * p.k) / k0;
    double ek = dot(E, p.k);
    V3 f = k0 * E + OPAL_C * cross(p.k, B);
    p.chi = sqrt(dot(f, f) - ek * ek) / OPAL_ECRIT;
    double dxi = v.x * dt / dx;
    int d = move_x(p.x, p.prev_x, dxi, cfl);
    p.y = p.y + v.y * dt;
    p.z = p.z + v.z * dt;
    return d;
}

// ---- YeeGrid::deposit, per-particle contributions  yee.rs:584-609 -----------------------------
// Stencil relative to the particle's (new) cell index `idx`:
//   jx  at idx-2..idx+2  (out[0..4]),  jy at idx-1..idx+1 (out[5..7]), jz at idx-1..idx+1 (out[8..10]),
//   rho at idx-2..idx+1 (out[11..14]).
// The reference's 4th jy/jz point carries weight(2+x) == 0 for x in [0,1) (quirk Q1) and is
// therefore not materialised; rho's 4th point goes to idx-2 with weight(2-x) (quirk Q2).
__device__ __forceinline__ void deposit_terms(double Q, double prev_x, double x, double vy, double vz, double dx, double dt,
                                              double out[15]) {
    out[0] = Q * flux(-1.5 - prev_x, -1.5 - x) / dt;
    out[1] = Q * flux(-0.5 - prev_x, -0.5 - x) / dt;
    out[2] = Q * flux(0.5 - prev_x, 0.5 - x) / dt;
    out[3] = Q * flux(1.5 - prev_x, 1.5 - x) / dt;
    out[4] = Q * flux(2.5 - prev_x, 2.5 - x) / dt;
    double w0 = weight(1.0 + x), w1 = weight(x), w2 = weight(1.0 - x), w3 = weight(2.0 - x);
    out[5] = Q * vy * w0 / dx;
    out[6] = Q * vy * w1 / dx;
    out[7] = Q * vy * w2 / dx;
    out[8] = Q * vz * w0 / dx;
    out[9] = Q * vz * w1 / dx;
    out[10] = Q * vz * w2 / dx;
    out[11] = Q * w3 / dx;   // -> idx-2
    out[12] = Q * w0 / dx;   // -> idx-1
    out[13] = Q * w1 / dx;   // -> idx
    out[14] = Q * w2 / dx;   // -> idx+1
}

// ---- Philox4x32-10 (counter-based RNG; replaces the cloned+jumped xoshiro streams of
// interactions.rs:62-65, which make the reference's own draws thread-count dependent) ---------
__device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
        uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
// two uniform doubles in [0,1) with 53 random bits each from one Philox block
__device__ __forceinline__ void philox_uniform2(uint64_t seed, uint64_t ctr_lo, uint32_t ctr_hi, uint32_t stream, double& a, double& b) {
    uint32_t c[4] = {(uint32_t)ctr_lo, (uint32_t)(ctr_lo >> 32), ctr_hi, stream};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    uint64_t r0 = ((uint64_t)c[1] << 32) | c[0], r1 = ((uint64_t)c[3] << 32) | c[2];
    a = (double)(r0 >> 11) * (1.0 / 9007199254740992.0);
    b = (double)(r1 >> 11) * (1.0 / 9007199254740992.0);
}

}  // namespace opal
