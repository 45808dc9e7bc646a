// opal_hgram.cu -- device-side distribution functions (particle/hgram.rs, particle/mod.rs:383-568).
//
// At 1e9 particles per GPU the particle data cannot leave HBM, so the two passes of
// Histogram::generate_{1d,2d} run where the particles live: pass 1 reduces min/max of the axis
// quantities, pass 2 bins with one f64 atomicAdd per particle into a grid of at most
// ceil(2 N^(1/3))^2 bins (L2-resident).  Both passes are pure streaming reads of the bucket / photon
// SoA arrays: HBM-bound, (8 B x fields touched) per particle.
//
// This translation unit is compiled with --fmad=false and the axis quantities are evaluated in the
// reference's operation order, so every +,-,*,/,sqrt result -- and therefore every bin INDEX derived
// from them -- is bit-identical to the CPU statement of hgram.rs.  (acos/asin/atan2/hypot/ln are CUDA
// libm: <= 2 ulp from glibc, which can move a value that sits on a bin edge.)
#include <cfloat>

#include "opal_device.cuh"
#include "opal_types.h"

namespace opal {

namespace {

struct HgramArgs {
    int ndim, q[2], log_scaled[2];
    int weight_kind, log_adjust;
    double gmin[2], bin_sz[2];
    unsigned long long nbins[2], len;
    double xmin, dx;
    double ion_mass_ratio;     // A * mp / me  (ion.rs:128, 140)
};

struct Pt {                    // what the accessors of mod.rs:388-449 can ask of one particle
    double pos, r, energy, p[3], work, chi, weight;
};

__device__ __forceinline__ double quantity(const Pt& t, int q) {
    switch (q) {
        case OPAL_Q_X: return t.pos;
        case OPAL_Q_R: return t.r;
        case OPAL_Q_ENERGY: return t.energy;
        case OPAL_Q_PX: return t.p[0];
        case OPAL_Q_PY: return t.p[1];
        case OPAL_Q_PZ: return t.p[2];
        case OPAL_Q_P_PERP: return hypot(t.p[1], t.p[2]);                                  // mod.rs:417-420
        case OPAL_Q_THETA: {                                                               // mod.rs:427-431
            const double mag = sqrt(t.p[0] * t.p[0] + t.p[1] * t.p[1] + t.p[2] * t.p[2]);
            return acos(t.p[0] / mag);
        }
        case OPAL_Q_PHI: return atan2(t.p[2], t.p[1]);                                     // mod.rs:434-437
        case OPAL_Q_LONGITUDE: return atan2(t.p[1], -t.p[0]);                              // mod.rs:440-443
        case OPAL_Q_LATITUDE: {                                                            // mod.rs:445-449
            const double mag = sqrt(t.p[0] * t.p[0] + t.p[1] * t.p[1] + t.p[2] * t.p[2]);
            return asin(t.p[2] / mag);
        }
        case OPAL_Q_WORK: return t.work;
        default: return t.chi;
    }
}

template <int SP, bool FULL>
__device__ __forceinline__ Pt load_charged(const BucketDev& b, int c, long long s, const HgramArgs& a) {
    Pt t;
    const double x = b.f[F_X][s];
    const V3 u = {b.f[F_UX][s], b.f[F_UY][s], b.f[F_UZ][s]};
    t.weight = b.f[F_W][s];
    t.pos = a.xmin + ((double)c + x) * a.dx;                                               // mod.rs:388-391
    if (SP == OPAL_ELECTRON) {
        const double gamma = FULL ? b.f[F_GAMMA][s] : sqrt(1.0 + dot(u, u));
        t.energy = gamma * OPAL_ME_MEV;                                                    // electron.rs:117
        t.p[0] = u.x * OPAL_ME_MEV; t.p[1] = u.y * OPAL_ME_MEV; t.p[2] = u.z * OPAL_ME_MEV; // electron.rs:156
        t.r = FULL ? hypot(b.f[F_Y][s], b.f[F_Z][s]) : 0.0;                                // electron.rs:109
        t.work = FULL ? b.f[F_WORK][s] : 0.0;
        t.chi = FULL ? b.f[F_CHI][s] : 0.0;
    } else {
        const double u2 = dot(u, u);
        const double gm1 = FULL ? b.f[F_GAMMA][s] : u2 / (1.0 + sqrt(1.0 + u2));
        t.energy = gm1 * a.ion_mass_ratio * OPAL_ME_MEV;                                   // ion.rs:128
        t.p[0] = u.x * a.ion_mass_ratio * OPAL_ME_MEV;                                     // ion.rs:140
        t.p[1] = u.y * a.ion_mass_ratio * OPAL_ME_MEV;
        t.p[2] = u.z * a.ion_mass_ratio * OPAL_ME_MEV;
        t.r = 0.0; t.chi = 0.0;                                                            // ion.rs:107, 151
        t.work = FULL ? b.f[F_WORK][s] : 0.0;
    }
    return t;
}

__device__ __forceinline__ Pt load_photon(const PhotonDev& ph, int i, const HgramArgs& a) {
    Pt t;
    const V3 k = {ph.f[P_KX][i], ph.f[P_KY][i], ph.f[P_KZ][i]};
    t.weight = ph.f[P_W][i];
    t.pos = a.xmin + ((double)ph.cell[i] + ph.f[P_X][i]) * a.dx;
    t.energy = sqrt(dot(k, k)) * OPAL_ME_MEV;                                              // photon.rs:216
    t.p[0] = k.x * OPAL_ME_MEV; t.p[1] = k.y * OPAL_ME_MEV; t.p[2] = k.z * OPAL_ME_MEV;     // photon.rs:205
    t.r = hypot(ph.f[P_Y][i], ph.f[P_Z][i]);                                               // photon.rs:189
    t.work = 0.0;
    t.chi = ph.f[P_CHI][i];
    return t;
}

// ---- pass 1 -------------------------------------------------------------------------------
struct MinMax { double mn[2], mx[2]; };

__device__ __forceinline__ void mm_take(MinMax& m, const Pt& t, const HgramArgs& a) {
    for (int d = 0; d < 2; d++) {
        if (d >= a.ndim) break;
        double v = quantity(t, a.q[d]);
        if (a.log_scaled[d]) v = log(v);
        if (isfinite(v)) { m.mn[d] = fmin(m.mn[d], v); m.mx[d] = fmax(m.mx[d], v); }       // hgram.rs:110-121
    }
}

// order-preserving map double <-> unsigned so min/max can use integer atomics
__device__ __forceinline__ unsigned long long ord(double v) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}

__device__ void mm_flush(MinMax m, unsigned long long* out /* min0,max0,min1,max1 */) {
    for (int d = 0; d < 2; d++) {
        double mn = m.mn[d], mx = m.mx[d];
        for (int o = 16; o; o >>= 1) {
            mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
            mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(out + 2 * d, ord(mn));
            atomicMax(out + 2 * d + 1, ord(mx));
        }
    }
}

template <int SP, bool FULL>
__global__ void __launch_bounds__(256) k_hgram_minmax(BucketDev b, HgramArgs a, unsigned long long* out) {
    MinMax m = {{DBL_MAX, DBL_MAX}, {-DBL_MAX, -DBL_MAX}};
    const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
    for (int c = blockIdx.x * wpb + (threadIdx.x >> 5); c < b.ncell; c += gridDim.x * wpb) {
        const int n = b.count[c];
        const long long base = b.start[c];
        for (int i = lane; i < n; i += 32) mm_take(m, load_charged<SP, FULL>(b, c, base + i, a), a);
    }
    mm_flush(m, out);
}

__global__ void __launch_bounds__(256) k_hgram_minmax_photon(PhotonDev ph, HgramArgs a, unsigned long long* out) {
    MinMax m = {{DBL_MAX, DBL_MAX}, {-DBL_MAX, -DBL_MAX}};
    const int n = *ph.n;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) mm_take(m, load_photon(ph, i, a), a);
    mm_flush(m, out);
}

// ---- pass 2 -------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long to_usize(double v) {      // Rust's saturating `as usize`
    if (!(v > 0.0)) return 0ull;                                         // negative and NaN -> 0
    if (v >= 18446744073709551615.0) return ~0ull;
    return (unsigned long long)v;
}

__device__ __forceinline__ double linear_bin_vol(double mn, double bin_sz, unsigned long long bin) {   // hgram.rs:126-128
    return exp(mn + (double)bin * bin_sz) * expm1(bin_sz);
}

__device__ __forceinline__ double bin_take(const Pt& t, const HgramArgs& a, double* counts) {
    double w = a.weight_kind == OPAL_W_ENERGY ? t.energy * t.weight : t.weight;             // mod.rs:496-499
    const double total = w;
    double v[2];
    bool finite = true;
    for (int d = 0; d < a.ndim; d++) {
        v[d] = quantity(t, a.q[d]);
        if (a.log_scaled[d]) v[d] = log(v[d]);
        finite = finite && isfinite(v[d]);
    }
    if (!finite) return total;                                                              // hgram.rs:234, 340
    unsigned long long bin[2] = {0ull, 0ull};
    for (int d = 0; d < a.ndim; d++) {
        bin[d] = a.bin_sz[d] == 0.0 ? 0ull : to_usize(floor((v[d] - a.gmin[d]) / a.bin_sz[d]));   // hgram.rs:228, 343-346
        if (a.log_scaled[d] && a.log_adjust)
            w *= a.bin_sz[d] == 0.0 ? 1.0 : a.bin_sz[d] / linear_bin_vol(a.gmin[d], a.bin_sz[d], bin[d]);   // hgram.rs:349-355
    }
    const unsigned long long fbin = a.ndim == 1 ? bin[0] : bin[1] * a.nbins[1] + bin[0];    // hgram.rs:357
    if (fbin < a.len) atomicAdd(counts + fbin, w);
    return total;
}

__device__ void total_flush(double tot, double* total) {
    for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    if ((threadIdx.x & 31) == 0 && tot != 0.0) atomicAdd(total, tot);
}

template <int SP, bool FULL>
__global__ void __launch_bounds__(256) k_hgram_bin(BucketDev b, HgramArgs a, double* counts, double* total) {
    double tot = 0.0;
    const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
    for (int c = blockIdx.x * wpb + (threadIdx.x >> 5); c < b.ncell; c += gridDim.x * wpb) {
        const int n = b.count[c];
        const long long base = b.start[c];
        for (int i = lane; i < n; i += 32) tot += bin_take(load_charged<SP, FULL>(b, c, base + i, a), a, counts);
    }
    total_flush(tot, total);
}

__global__ void __launch_bounds__(256) k_hgram_bin_photon(PhotonDev ph, HgramArgs a, double* counts, double* total) {
    double tot = 0.0;
    const int n = *ph.n;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) tot += bin_take(load_photon(ph, i, a), a, counts);
    total_flush(tot, total);
}

__global__ void k_hgram_init(unsigned long long* mm) {
    if (threadIdx.x < 4) mm[threadIdx.x] = (threadIdx.x & 1) ? ord(-DBL_MAX) : ord(DBL_MAX);
}
__global__ void k_hgram_decode(const unsigned long long* mm, double* out) {
    if (threadIdx.x < 4) {
        const unsigned long long o = mm[threadIdx.x];
        const unsigned long long bits = (o >> 63) ? (o & 0x7fffffffffffffffull) : ~o;
        out[threadIdx.x] = __longlong_as_double((long long)bits);
    }
}

HgramArgs make_args(const GridDev& g, double mass_number, int ndim, const int* q, const int* lg) {
    HgramArgs a = {};
    a.ndim = ndim;
    for (int d = 0; d < ndim; d++) { a.q[d] = q[d]; a.log_scaled[d] = lg ? lg[d] : 0; }
    a.xmin = g.xmin; a.dx = g.dx;
    a.ion_mass_ratio = mass_number * OPAL_MP / OPAL_ME;
    return a;
}

// grid: 148 SMs x 8 resident 256-thread blocks
constexpr int HGRAM_GRID = 148 * 8;

}  // namespace

// scratch: 4 ull + 4 doubles (min0,max0,min1,max1); out4 is a device pointer to the 4 decoded doubles
void hgram_minmax_charged(cudaStream_t s, const GridDev& g, const BucketDev& b, int species, bool full, int ndim, const int* q, const int* lg,
                          unsigned long long* d_mm, double* d_out4) {
    const HgramArgs a = make_args(g, b.mass_number, ndim, q, lg);
    k_hgram_init<<<1, 32, 0, s>>>(d_mm);
    if (species == OPAL_ELECTRON) {
        if (full) k_hgram_minmax<OPAL_ELECTRON, true><<<HGRAM_GRID, 256, 0, s>>>(b, a, d_mm);
        else k_hgram_minmax<OPAL_ELECTRON, false><<<HGRAM_GRID, 256, 0, s>>>(b, a, d_mm);
    } else {
        if (full) k_hgram_minmax<OPAL_ION, true><<<HGRAM_GRID, 256, 0, s>>>(b, a, d_mm);
        else k_hgram_minmax<OPAL_ION, false><<<HGRAM_GRID, 256, 0, s>>>(b, a, d_mm);
    }
    k_hgram_decode<<<1, 32, 0, s>>>(d_mm, d_out4);
}

void hgram_minmax_photon(cudaStream_t s, const GridDev& g, const PhotonDev& ph, int ndim, const int* q, const int* lg,
                         unsigned long long* d_mm, double* d_out4) {
    const HgramArgs a = make_args(g, 1.0, ndim, q, lg);
    k_hgram_init<<<1, 32, 0, s>>>(d_mm);
    k_hgram_minmax_photon<<<HGRAM_GRID, 256, 0, s>>>(ph, a, d_mm);
    k_hgram_decode<<<1, 32, 0, s>>>(d_mm, d_out4);
}

static HgramArgs bin_args(const GridDev& g, double mass_number, int ndim, const int* q, const int* lg, const double* gmin, const double* bin_sz,
                          const long long* nbins, int weight_kind, int log_adjust) {
    HgramArgs a = make_args(g, mass_number, ndim, q, lg);
    a.weight_kind = weight_kind; a.log_adjust = log_adjust;
    a.len = 1;
    for (int d = 0; d < ndim; d++) { a.gmin[d] = gmin[d]; a.bin_sz[d] = bin_sz[d]; a.nbins[d] = (unsigned long long)nbins[d]; a.len *= a.nbins[d]; }
    return a;
}

void hgram_bin_charged(cudaStream_t s, const GridDev& g, const BucketDev& b, int species, bool full, int ndim, const int* q, const int* lg,
                       const double* gmin, const double* bin_sz, const long long* nbins, int weight_kind, int log_adjust,
                       double* d_counts, double* d_total) {
    const HgramArgs a = bin_args(g, b.mass_number, ndim, q, lg, gmin, bin_sz, nbins, weight_kind, log_adjust);
    if (species == OPAL_ELECTRON) {
        if (full) k_hgram_bin<OPAL_ELECTRON, true><<<HGRAM_GRID, 256, 0, s>>>(b, a, d_counts, d_total);
        else k_hgram_bin<OPAL_ELECTRON, false><<<HGRAM_GRID, 256, 0, s>>>(b, a, d_counts, d_total);
    } else {
        if (full) k_hgram_bin<OPAL_ION, true><<<HGRAM_GRID, 256, 0, s>>>(b, a, d_counts, d_total);
        else k_hgram_bin<OPAL_ION, false><<<HGRAM_GRID, 256, 0, s>>>(b, a, d_counts, d_total);
    }
}

void hgram_bin_photon(cudaStream_t s, const GridDev& g, const PhotonDev& ph, int ndim, const int* q, const int* lg, const double* gmin,
                      const double* bin_sz, const long long* nbins, int weight_kind, int log_adjust, double* d_counts, double* d_total) {
    const HgramArgs a = bin_args(g, 1.0, ndim, q, lg, gmin, bin_sz, nbins, weight_kind, log_adjust);
    k_hgram_bin_photon<<<HGRAM_GRID, 256, 0, s>>>(ph, a, d_counts, d_total);
}

}  // namespace opal
