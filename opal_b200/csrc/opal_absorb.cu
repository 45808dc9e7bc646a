// opal_absorb.cu -- photon absorption and stimulated emission (particle/interactions.rs:145-340, photon.rs:312-362,
// qed/photon_absorption.rs, qed/stimulated_emission.rs, qed/special_functions/airy.rs).
//
// Compiled with --fmad=false (see opal_absorb.cuh).
#include <algorithm>
#include <cstdint>

#include "opal_device.cuh"
#include "opal_absorb.cuh"
#include "opal_b200.h"

namespace opal {
namespace pairs {

__global__ void k_airy(int n, const double* __restrict__ x, double* __restrict__ ai, int* __restrict__ defined) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = 0.0;
    const bool ok = airy_ai_for_positive(x[i], &v);
    ai[i] = ok ? v : 0.0;
    defined[i] = ok ? 1 : 0;
}

__global__ void k_pair_cross_sections(int n, const double* __restrict__ k4, const double* __restrict__ p4, const double* __restrict__ chi_gamma,
                                      const double* __restrict__ chi_e, double* __restrict__ sigma_abs, double* __restrict__ sigma_st,
                                      int* __restrict__ defined) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double k[4] = {k4[4 * i], k4[4 * i + 1], k4[4 * i + 2], k4[4 * i + 3]};
    const double p[4] = {p4[4 * i], p4[4 * i + 1], p4[4 * i + 2], p4[4 * i + 3]};
    double sa = 0.0, ss = 0.0;
    const bool a = pair_cross_section<+1>(k, p, chi_gamma[i], chi_e[i], &sa);
    const bool s = pair_cross_section<-1>(k, p, chi_gamma[i], chi_e[i], &ss);
    sigma_abs[i] = a ? sa : 0.0;
    sigma_st[i] = s ? ss : 0.0;
    defined[i] = (a ? 1 : 0) | (s ? 2 : 0);
}

namespace {
// a few device arrays for the lifetime of one pure-function call
struct Scratch {
    void* p[8] = {nullptr};
    int n = 0;
    bool ok = true;
    template <class T>
    T* in(const T* host, size_t count) {
        T* d = out<T>(count);
        if (d && cudaMemcpy(d, host, count * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) ok = false;
        return d;
    }
    template <class T>
    T* out(size_t count) {
        void* d = nullptr;
        if (cudaMalloc(&d, count * sizeof(T)) != cudaSuccess) { ok = false; return nullptr; }
        p[n++] = d;
        return (T*)d;
    }
    template <class T>
    void back(T* host, const T* dev, size_t count) {
        if (ok && cudaMemcpy(host, dev, count * sizeof(T), cudaMemcpyDeviceToHost) != cudaSuccess) ok = false;
    }
    ~Scratch() { for (int i = 0; i < n; i++) cudaFree(p[i]); }
};
}  // namespace

}  // namespace pairs
}  // namespace opal

using namespace opal::pairs;

extern "C" {

opal_status opal_qed_airy(int64_t n, const double* x, double* ai, int* defined) {
    if (n < 0 || n > INT32_MAX || (n > 0 && (!x || !ai || !defined))) return OPAL_EINVAL;
    if (n == 0) return OPAL_OK;
    Scratch s;
    const double* d_x = s.in(x, (size_t)n);
    double* d_ai = s.out<double>((size_t)n);
    int* d_ok = s.out<int>((size_t)n);
    if (!s.ok) return OPAL_ECUDA;
    k_airy<<<(unsigned)((n + 127) / 128), 128>>>((int)n, d_x, d_ai, d_ok);
    if (cudaGetLastError() != cudaSuccess) return OPAL_ECUDA;
    s.back(ai, d_ai, (size_t)n);
    s.back(defined, d_ok, (size_t)n);
    return s.ok ? OPAL_OK : OPAL_ECUDA;
}

opal_status opal_qed_pair_cross_sections(int64_t n, const double* k4, const double* p4, const double* chi_gamma, const double* chi_e,
                                         double* sigma_abs, double* sigma_st, int* defined) {
    if (n < 0 || n > INT32_MAX / 4 || (n > 0 && (!k4 || !p4 || !chi_gamma || !chi_e || !sigma_abs || !sigma_st || !defined))) return OPAL_EINVAL;
    if (n == 0) return OPAL_OK;
    Scratch s;
    const double* d_k = s.in(k4, (size_t)n * 4);
    const double* d_p = s.in(p4, (size_t)n * 4);
    const double* d_cg = s.in(chi_gamma, (size_t)n);
    const double* d_ce = s.in(chi_e, (size_t)n);
    double* d_sa = s.out<double>((size_t)n);
    double* d_ss = s.out<double>((size_t)n);
    int* d_ok = s.out<int>((size_t)n);
    if (!s.ok) return OPAL_ECUDA;
    k_pair_cross_sections<<<(unsigned)((n + 127) / 128), 128>>>((int)n, d_k, d_p, d_cg, d_ce, d_sa, d_ss, d_ok);
    if (cudaGetLastError() != cudaSuccess) return OPAL_ECUDA;
    s.back(sigma_abs, d_sa, (size_t)n);
    s.back(sigma_st, d_ss, (size_t)n);
    s.back(defined, d_ok, (size_t)n);
    return s.ok ? OPAL_OK : OPAL_ECUDA;
}

}  // extern "C"

// =====================================================================================================================
// absorb (particle/interactions.rs:145-340)
//
// The reference walks, photon by photon, the electrons of the photon's cell in store order, spending the photon's two
// optical depths pair by pair and stopping at the first event.  Here one WARP takes one photon: the 32 lanes evaluate
// the two cross sections of 32 electrons of the cell bucket at once (that is where the time goes: up to 40 pow() per
// Airy function; the quadratures are shared out over the warp, see airy_warp in opal_absorb.cuh), then every lane replays the sequential optical-depth bookkeeping over those 32 partial probabilities
// in store order -- the same subtractions in the same order as the reference, so the event, the electron it happens at
// and the depths left over are those of the sequential loop.  Events are recorded per photon (no atomics); the kicks,
// the copies made by stimulated emission and the deletion of absorbed photons follow in separate passes, as the
// reference applies them after the pair loop (interactions.rs:291-330).
// =====================================================================================================================
namespace opal {
namespace pairs {

constexpr unsigned FULLMASK = 0xffffffffu;
#ifndef OPAL_ABSORB_MINBLK
#define OPAL_ABSORB_MINBLK 8    // measured: 4 -> 8.06 ms, 6 -> 7.16 ms, 8 -> 6.76 ms per 16.8e6 pairs (64 regs, occupancy wins over the spills)
#endif
constexpr double PHOTON_E_ECRIT_CUTOFF = 1.0e-8;     // interactions.rs:146
constexpr unsigned ABSORB_STREAM = 8u;               // Philox streams 8..11: apart from emit_radiation's 0..3

struct AbsorbArgs {
    double t, max_displacement, stop_time;           // NaN == None
    unsigned long long seed;
    unsigned call_index;
    int no_stimulated;
};

__global__ void k_last_occupied(BucketDev e, int* __restrict__ last) {
    int best = -1;
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < e.ncell; c += gridDim.x * blockDim.x)
        if (e.count[c] > 0) best = c;
    if (best >= 0) atomicMax(last, best);
}

// kind[i]: 0 nothing, 1 absorbed, 2 stimulated emission; slot[i]: bucket slot of the electron it happened at
__global__ void __launch_bounds__(128, OPAL_ABSORB_MINBLK) k_absorb_scan(GridDev g, BucketDev e, PhotonDev ph, AbsorbArgs a, const int* __restrict__ last_occupied,
                                                      int* __restrict__ kind, long long* __restrict__ slot, StatusDev* __restrict__ st) {
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    const int n = *ph.n;
    const int last = *last_occupied;
    const double cdt_dx = OPAL_C * g.dt / g.dx;
    bool bad = false;
    __shared__ double s_gl_x[60], s_gl_w[60];               // the quadrature nodes: lanes read different ones (see airy_warp)
    if (threadIdx.x < 60) { s_gl_x[threadIdx.x] = c_gl_x[threadIdx.x]; s_gl_w[threadIdx.x] = c_gl_w[threadIdx.x]; }
    __syncthreads();
    for (int i = warp; i < n; i += nwarps) {
        if (lane == 0) { kind[i] = 0; slot[i] = -1; }
        const double kx = ph.f[P_KX][i], ky = ph.f[P_KY][i], kz = ph.f[P_KZ][i];
        const double k0 = sqrt(kx * kx + ky * ky + kz * kz);
        const double chi_g = ph.f[P_CHI][i];
        if (chi_g * OPAL_ME_MEV / (k0 * OPAL_ME_MEV) < PHOTON_E_ECRIT_CUTOFF) continue;
        if (!isnan(a.stop_time) && a.t - ph.f[P_BIRTH][i] > a.stop_time) continue;
        if (!isnan(a.max_displacement) && hypot(ph.f[P_Y][i], ph.f[P_Z][i]) > a.max_displacement) continue;
        const int c = ph.cell[i];
        if (c < 0 || c >= e.ncell || c == last) continue;      // (the last occupied cell: see the header of opal_absorb)
        const int ne = e.count[c];
        if (ne == 0) continue;
        const long long base = e.start[c];
        const double k4[4] = {k0, kx, ky, kz};
        double tau_abs = ph.f[P_TAU_ABS][i], tau_st = ph.f[P_TAU_ST][i];
        int what = 0;
        long long where = -1;
        for (int j0 = 0; j0 < ne && what == 0; j0 += 32) {
            double prob[2] = {0.0, 0.0};                          // [0] absorption, [1] stimulated emission
            double p4[4] = {1.0, 0.0, 0.0, 0.0}, chi_e = 0.0, w = 0.0;
            const bool have = j0 + lane < ne;
            if (have) {
                const long long s = base + j0 + lane;
                p4[0] = e.f[F_GAMMA][s]; p4[1] = e.f[F_UX][s]; p4[2] = e.f[F_UY][s]; p4[3] = e.f[F_UZ][s];
                chi_e = e.f[F_CHI][s];
                w = e.f[F_W][s];
            }
#pragma unroll 1
            for (int which = 0; which < (a.no_stimulated ? 1 : 2); which++) {
                PairPre pre = {false, 0.0, 0.0, 1.0};
                if (have) pre = pair_prepare(which == 0 ? +1 : -1, k4, p4, chi_g, chi_e);
                if (!__any_sync(FULLMASK, pre.ok)) continue;
                double ai;
                if (airy_warp(pre.ok, pre.zbar, s_gl_x, s_gl_w, lane, &ai)) {
                    const double partial = w * cdt_dx * (pre.factor * ai / pre.denom);
                    if (!(partial >= 0.0)) bad = true;
                    if (which == 0) prob[0] = partial; else prob[1] = partial;
                }
            }
            const double pa = prob[0], ps = prob[1];
            const int m = min(32, ne - j0);
            for (int l = 0; l < m; l++) {                        // Photon::interacts_with (photon.rs:312-362), in store order
                const double prob_abs = __shfl_sync(FULLMASK, pa, l), prob_st = __shfl_sync(FULLMASK, ps, l);
                tau_abs -= prob_abs;                             // a None cross section subtracts nothing: prob == 0
                tau_st -= prob_st;
                if (tau_abs < 0.0 && tau_st < 0.0) {
                    double r, r_abs, r_st, unused;
                    philox_uniform2(a.seed, (uint64_t)i, a.call_index, ABSORB_STREAM, r, r_abs);
                    philox_uniform2(a.seed, (uint64_t)i, a.call_index, ABSORB_STREAM + 1u, r_st, unused);
                    if (r < prob_abs / (prob_abs + prob_st)) what = 1;
                    else { tau_abs = -log1p(-r_abs); tau_st = -log1p(-r_st); what = 2; }
                } else if (tau_abs < 0.0) {
                    what = 1;
                } else if (tau_st < 0.0) {
                    double r_st, unused;
                    philox_uniform2(a.seed, (uint64_t)i, a.call_index, ABSORB_STREAM + 1u, r_st, unused);
                    tau_st = -log1p(-r_st);
                    what = 2;
                }
                if (what) { where = base + j0 + l; break; }
            }
        }
        if (lane == 0) {
            ph.f[P_TAU_ABS][i] = tau_abs; ph.f[P_TAU_ST][i] = tau_st;
            kind[i] = what; slot[i] = where;
        }
    }
    if (bad) atomicOr(&st->err, ERR_PROB);
}

// Electron::kick (electron.rs:256-262) for every event, and the copy a stimulated emission adds (interactions.rs:305-310)
__global__ void k_absorb_apply(BucketDev e, PhotonDev ph, AbsorbArgs a, int n0, const int* __restrict__ kind, const long long* __restrict__ slot,
                               StatusDev* __restrict__ st) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n0; i += gridDim.x * blockDim.x) {
        const int what = kind[i];
        if (what == 0) continue;
        const long long s = slot[i];
        const double k[3] = {ph.f[P_KX][i] * OPAL_ME_MEV, ph.f[P_KY][i] * OPAL_ME_MEV, ph.f[P_KZ][i] * OPAL_ME_MEV};     // Photon::momentum
        const double we = e.f[F_W][s];
        // absorbed: kick(k, photon weight); stimulated: kick(-k, electron weight)
        const double scale = what == 1 ? ph.f[P_W][i] / we : we / we;
        const double sgn = what == 1 ? 1.0 : -1.0;
        atomicAdd(&e.f[F_UX][s], scale * (sgn * k[0]) / OPAL_ME_MEV);
        atomicAdd(&e.f[F_UY][s], scale * (sgn * k[1]) / OPAL_ME_MEV);
        atomicAdd(&e.f[F_UZ][s], scale * (sgn * k[2]) / OPAL_ME_MEV);
        if (what == 2) {
            const int o = atomicAdd(ph.n, 1);
            if (o >= ph.cap) { atomicOr(&st->err, ERR_CAPACITY | ERR_CAP_PHOTONS); continue; }
            double r0, r1, r2, r3;
            philox_uniform2(a.seed, (uint64_t)i, a.call_index, ABSORB_STREAM + 2u, r0, r1);
            philox_uniform2(a.seed, (uint64_t)i, a.call_index, ABSORB_STREAM + 3u, r2, r3);
            ph.cell[o] = ph.cell[i];
            for (int f = 0; f < NF_PHOTON; f++) ph.f[f][o] = ph.f[f][i];
            ph.f[P_W][o] = we;                                   // photon.with_weight(e.weight())
            ph.f[P_TAU][o] = -log1p(-r0);                        // with_optical_depths: tau[0] (tau[1] is never read on this path)
            (void)r1;
            ph.f[P_TAU_ABS][o] = -log1p(-r2); ph.f[P_TAU_ST][o] = -log1p(-r3);
            ph.f[P_BIRTH][o] = a.t;                              // at_time(t)
        }
    }
}

__global__ void k_absorb_gamma(BucketDev e, int n0, const int* __restrict__ kind, const long long* __restrict__ slot) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n0; i += gridDim.x * blockDim.x) {
        if (kind[i] == 0) continue;
        const long long s = slot[i];
        const double ux = e.f[F_UX][s], uy = e.f[F_UY][s], uz = e.f[F_UZ][s];
        e.f[F_GAMMA][s] = sqrt(1.0 + (ux * ux + uy * uy + uz * uz));
    }
}

// the photon store without the absorbed ones (interactions.rs:291-301); copies appended by k_absorb_apply sit at >= n0
__global__ void k_absorb_compact(PhotonDev src, PhotonDev dst, int n0, const int* __restrict__ kind) {
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int n = min(*src.n, src.cap);
    const int n_round = (n + 31) & ~31;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x) {
        const bool keep = i < n && (i >= n0 || kind[i] != 1);
        const unsigned mk = __ballot_sync(FULLMASK, keep);
        if (!mk) continue;
        int base = 0;
        const int leader = __ffs(mk) - 1;
        if (lane == leader) base = atomicAdd(dst.n, __popc(mk));
        base = __shfl_sync(FULLMASK, base, leader);
        if (keep) {
            const int o = base + __popc(mk & lt_mask);
            dst.cell[o] = src.cell[i];
            for (int f = 0; f < NF_PHOTON; f++) dst.f[f][o] = src.f[f][i];
        }
    }
}

}  // namespace pairs

// host side of opal_absorb (opal_host.cu owns the context): 5 launches on `stream`
int absorb_launch(cudaStream_t stream, const GridDev& g, const BucketDev& e, const PhotonDev& src, const PhotonDev& dst, int n0, double t,
                  double max_displacement, double stop_time, unsigned long long seed, unsigned call_index, int no_stimulated, int* d_last,
                  int* d_kind, long long* d_slot, StatusDev* d_status) {
    using namespace pairs;
    const AbsorbArgs a = {t, max_displacement, stop_time, seed, call_index, no_stimulated};
    cudaMemsetAsync(d_last, 0xFF, sizeof(int), stream);
    cudaMemsetAsync(dst.n, 0, sizeof(int), stream);
    k_last_occupied<<<148, 256, 0, stream>>>(e, d_last);
    const int warps_per_block = 4;
    const int blocks = std::min((n0 + warps_per_block - 1) / warps_per_block, 148 * 16);
    k_absorb_scan<<<blocks, warps_per_block * 32, 0, stream>>>(g, e, src, a, d_last, d_kind, d_slot, d_status);
    const int gr = std::min((n0 + 255) / 256, 148 * 8);
    k_absorb_apply<<<gr, 256, 0, stream>>>(e, src, a, n0, d_kind, d_slot, d_status);
    k_absorb_gamma<<<gr, 256, 0, stream>>>(e, n0, d_kind, d_slot);
    k_absorb_compact<<<std::min((2 * n0 + 255) / 256, 148 * 8), 256, 0, stream>>>(src, dst, n0, d_kind);
    return 5;
}

}  // namespace opal
