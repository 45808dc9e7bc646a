// opal_kernels.cu -- sm_100a kernels of the opal PIC hot path.
//
// Compiled twice into one library: -DOPAL_KNS=fast (FMA contraction on) and -DOPAL_KNS=strict
// (--fmad=false: the rustc-like arithmetic used for bit-level comparisons with the CPU oracle).
//
// Layout recap: charged species live in cell-major gapped buckets, SoA f64 streams,
// one warp per cell.  Every lane of a warp therefore shares the 4 grid cells it gathers from
// (held in registers), the cell index is implicit (never loaded or stored), and the deposit of
// the ~97 % of particles that stay in their cell is a per-lane register accumulation followed by
// ONE warp shuffle-reduction and <= 13 global atomics per CELL (not per particle).
#include "opal_types.h"
#include "opal_device.cuh"

#ifndef OPAL_KNS
#define OPAL_KNS fast
#endif

namespace opal {
namespace OPAL_KNS {

constexpr unsigned FULLMASK = 0xffffffffu;
constexpr int WARPS_PER_BLOCK = 8;
constexpr int BLOCK = WARPS_PER_BLOCK * 32;

// photon_emission.rs:14-56, second column (the first is OPAL_LN_H_X0 + i * OPAL_DELTA_LN_CHI)
__constant__ double c_ln_h[41] = {
    1.64660829e+0, 1.64437994e+0, 1.64162210e+0, 1.63822181e+0, 1.63404803e+0, 1.62895132e+0, 1.62276454e+0,
    1.61530492e+0, 1.60637755e+0, 1.59578044e+0, 1.58331053e+0, 1.56877061e+0, 1.55197597e+0, 1.53276058e+0,
    1.51098175e+0, 1.48652339e+0, 1.45929745e+0, 1.42924408e+0, 1.39633080e+0, 1.36055104e+0, 1.32192247e+0,
    1.28048533e+0, 1.23630061e+0, 1.18944834e+0, 1.14002554e+0, 1.08814410e+0, 1.03392829e+0, 9.77512163e-1,
    9.19036694e-1, 8.58647059e-1, 7.96489918e-1, 7.32710951e-1, 6.67452686e-1, 6.00852676e-1, 5.33042051e-1,
    4.64144437e-1, 3.94275229e-1, 3.23541171e-1, 2.52040210e-1, 1.79861566e-1, 1.07085976e-1};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULLMASK, v, o);
    return v;
}

// Bursts (graph-replayed steps): may this kernel run?  No if an earlier kernel raised the halt; no, too, if the advance of
// phase g.prev_phase parked movers or tripped an assert -- then this kernel raises the halt in that phase's name (what a
// k_burst_mark launch of its own used to do).  Every thread gets the same answer: n_overflow is written by the charged
// species' push / scatter only, and an assert is fatal anyway.
__device__ __forceinline__ bool burst_blocked(const GridDev& g) {
    if (!g.halt) return false;
    if (*(volatile const int*)g.halt) return true;
    StatusDev* const st = g.bst;
    if (st && g.prev_phase >= 0) {
        bool filling = false;                                      // a mover segment more than half full (block-wide vote)
        if (g.prev_phase <= 1) {                                   // (only a charged species' advance fills segments)
            const int* seg = reinterpret_cast<const int*>(st + 1);
            for (int i = threadIdx.x; i < OPAL_MOVER_SEGMENTS; i += blockDim.x) filling |= 2 * seg[i] > g.seg_cap;
            filling = __syncthreads_or(filling);
        }
        if (st->n_overflow > 0 || st->err || filling) {
            if (blockIdx.x == 0 && threadIdx.x == 0) { st->halt_species = g.prev_phase; __threadfence(); *(volatile int*)&st->halt = 1; }
            return true;
        }
    }
    return false;
}

__device__ __forceinline__ void load_cell_fields(const GridDev& g, int j, CellFields& f) {
    // j = c + left_bdy_size; cells j-1 .. j+2 always exist (>= 4 ghost cells each side)
#pragma unroll
    for (int k = 0; k < 3; k++) {
        f.ex[k] = __ldg(g.ex + j - 1 + k);
        f.by[k] = __ldg(g.by + j - 1 + k);
        f.bz[k] = __ldg(g.bz + j - 1 + k);
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        f.ey[k] = __ldg(g.ey + j - 1 + k);
        f.ez[k] = __ldg(g.ez + j - 1 + k);
    }
    f.bx = __ldg(g.bx + j);
}

__device__ __forceinline__ double species_charge(int sp, double z_star) {
    return sp == OPAL_ELECTRON ? OPAL_QE : z_star * OPAL_E;    // electron.rs:140, ion.rs:115
}

__device__ __forceinline__ void record_from_state(opal_migrant& m, long long cell, const PState& p, int sp) {
    m.cell = cell; m.x = p.x; m.prev_x = p.prev_x; m.u[0] = p.u.x; m.u[1] = p.u.y; m.u[2] = p.u.z; m.weight = p.w;
    if (sp == OPAL_ION) { m.gamma = 1.0 + p.gamma; m.pad[0] = p.gamma; } else { m.gamma = p.gamma; m.pad[0] = 0.0; }
    m.chi = p.chi; m.tau = p.tau; m.work = p.work; m.y = p.y; m.z = p.z; m.birth_time = 0.0; m.pad[1] = 0.0;
}

// atomics of one particle's deposit into the grid (movers, migrants); idx = cell + left_bdy_size
__device__ __forceinline__ void deposit_atomic(const GridDev& g, int idx, double Q, double prev_x, double x, double vy, double vz) {
    double t[15];
    deposit_terms(Q, prev_x, x, vy, vz, g.dx, g.dt, t);
#pragma unroll
    for (int k = 0; k < 5; k++) if (t[k] != 0.0) atomicAdd(g.jx + idx - 2 + k, t[k]);
#pragma unroll
    for (int k = 0; k < 3; k++) {
        if (t[5 + k] != 0.0) atomicAdd(g.jy + idx - 1 + k, t[5 + k]);
        if (t[8 + k] != 0.0) atomicAdd(g.jz + idx - 1 + k, t[8 + k]);
    }
#pragma unroll
    for (int k = 0; k < 4; k++) if (t[11 + k] != 0.0) atomicAdd(g.rho + idx - 2 + k, t[11 + k]);
}

// ============================================================================================
// K1: gather + push (+ fused deposit of stayers) + in-place bucket compaction + mover emission.
//     Replaces the push loop of Population::advance (mod.rs:263-282) and, when DEP, the stayers'
//     share of YeeGrid::deposit (yee.rs:551-641).
// ============================================================================================
template <int SP, bool FULL, bool DEP>
__global__ void __launch_bounds__(BLOCK) k_push(GridDev g, BucketDev b, StepParams sp, opal_migrant* __restrict__ movers,
                                                 int* __restrict__ seg_count, StatusDev* __restrict__ st) {
    if (g.halt && *(volatile const int*)g.halt) return;   // burst mode: an earlier kernel needs the host first
    __shared__ double s_ln_h[41];
    if (FULL && SP == OPAL_ELECTRON) {
        if (threadIdx.x < 41) s_ln_h[threadIdx.x] = c_ln_h[threadIdx.x];
        __syncthreads();
    }
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int warps_total = gridDim.x * WARPS_PER_BLOCK;
    const double q = species_charge(SP, b.z_star);
    bool cfl = false;
    bool cap_err = false;

    for (int c = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5); c < b.ncell; c += warps_total) {
        const int n = b.count[c];
        if (n == 0) continue;
        const long long base = b.start[c];
        const int j = c + g.left;
        CellFields cf;
        load_cell_fields(g, j, cf);

        // stayers' deposit accumulators: jx at j-1..j+1, jy, jz at j-1..j+1, rho at j-2..j+1
        double ajx[3] = {0.0, 0.0, 0.0}, ajy[3] = {0.0, 0.0, 0.0}, ajz[3] = {0.0, 0.0, 0.0}, arho[4] = {0.0, 0.0, 0.0, 0.0};

        // software pipeline: the next chunk's loads are issued before this chunk's stores
        double nx_[NF_FULL];
        {
            const int i = lane;
            if (i < n) {
#pragma unroll
                for (int k = 0; k < (FULL ? NF_FULL : NF_LEAN); k++) nx_[k] = b.f[k][base + i];
            }
        }
        int written = 0;
        for (int k0 = 0; k0 < n; k0 += 32) {
            const int i = k0 + lane;
            const bool valid = i < n;
            PState p;
            p.x = nx_[F_X]; p.u = {nx_[F_UX], nx_[F_UY], nx_[F_UZ]}; p.w = nx_[F_W];
            if (FULL) { p.prev_x = nx_[F_PREVX]; p.gamma = nx_[F_GAMMA]; p.chi = nx_[F_CHI]; p.tau = nx_[F_TAU]; p.work = nx_[F_WORK]; p.y = nx_[F_Y]; p.z = nx_[F_Z]; }
            else { p.prev_x = 0.0; p.gamma = 1.0; p.chi = 0.0; p.tau = 0.0; p.work = 0.0; p.y = 0.0; p.z = 0.0; }
            {
                const int in = k0 + 32 + lane;
                if (in < n) {
#pragma unroll
                    for (int k = 0; k < (FULL ? NF_FULL : NF_LEAN); k++) nx_[k] = b.f[k][base + in];
                }
            }
            int d = 0;
            if (valid) {
                V3 E, B;
                fields_at(cf, p.x, E, B);
                if (SP == OPAL_ELECTRON) d = electron_push<FULL>(p, E, B, g.dx, g.dt, sp.flags, s_ln_h, cfl);
                else d = ion_push<FULL>(p, E, B, g.dx, g.dt, b.z_star, b.mass_number, cfl);
            }
            const bool stay = valid && d == 0;
            const bool move = valid && d != 0;
            const unsigned m_stay = __ballot_sync(FULLMASK, stay);
            const unsigned m_move = __ballot_sync(FULLMASK, move);
            // reserve mover slots early (one atomic per chunk that has movers, on one of
            // OPAL_MOVER_SEGMENTS counters) so its latency overlaps the stayers' deposit below
            // (segment = hash(cell, chunk): a few cells holding most of the slab -- a compression front -- still spread over all segments)
            int mbase = 0;
            const int seg = (int)(((unsigned)c * 0x9E3779B1u + (unsigned)(k0 >> 5) * 0x85EBCA77u) >> 22);   // top 10 bits: OPAL_MOVER_SEGMENTS == 1024
            if (m_move) {
                const int leader = __ffs(m_move) - 1;
                if (lane == leader) mbase = atomicAdd(seg_count + seg, __popc(m_move));
                mbase = __shfl_sync(FULLMASK, mbase, leader);
            }
            if (stay) {
                const int pos = written + __popc(m_stay & lt_mask);
                const long long s = base + pos;
                b.f[F_X][s] = p.x; b.f[F_UX][s] = p.u.x; b.f[F_UY][s] = p.u.y; b.f[F_UZ][s] = p.u.z;
                if (FULL) {
                    b.f[F_W][s] = p.w; b.f[F_PREVX][s] = p.prev_x; b.f[F_GAMMA][s] = p.gamma; b.f[F_CHI][s] = p.chi;
                    b.f[F_TAU][s] = p.tau; b.f[F_WORK][s] = p.work; b.f[F_Y][s] = p.y; b.f[F_Z][s] = p.z;
                } else if (pos != i) {
                    b.f[F_W][s] = p.w;
                }
                if (DEP) {
                    // a stayer has prev_x and x in [0,1): the faces at -1.5 and +2.5 carry zero flux
                    const double Q = p.w * q;
                    const V3 v = SP == OPAL_ELECTRON ? electron_velocity(p) : ion_velocity(p);
                    ajx[0] += Q * flux(-0.5 - p.prev_x, -0.5 - p.x) / g.dt;
                    ajx[1] += Q * flux(0.5 - p.prev_x, 0.5 - p.x) / g.dt;
                    ajx[2] += Q * flux(1.5 - p.prev_x, 1.5 - p.x) / g.dt;
                    const double w0 = weight(1.0 + p.x), w1 = weight(p.x), w2 = weight(1.0 - p.x), w3 = weight(2.0 - p.x);
                    ajy[0] += Q * v.y * w0 / g.dx; ajy[1] += Q * v.y * w1 / g.dx; ajy[2] += Q * v.y * w2 / g.dx;
                    ajz[0] += Q * v.z * w0 / g.dx; ajz[1] += Q * v.z * w1 / g.dx; ajz[2] += Q * v.z * w2 / g.dx;
                    arho[0] += Q * w3 / g.dx; arho[1] += Q * w0 / g.dx; arho[2] += Q * w1 / g.dx; arho[3] += Q * w2 / g.dx;
                }
            }
            if (move) {
                const int r = mbase + __popc(m_move & lt_mask);
                if (r < sp.mover_cap) {
                    opal_migrant m;
                    record_from_state(m, (long long)c + d, p, SP);
                    movers[(long long)seg * sp.mover_cap + r] = m;
                } else {
                    cap_err = true;
                }
            }
            written += __popc(m_stay);
        }
        if (lane == 0) b.count[c] = written;

        if (DEP) {
            // one shuffle reduction and <= 13 atomics per cell
            double v[13] = {ajx[0], ajx[1], ajx[2], ajy[0], ajy[1], ajy[2], ajz[0], ajz[1], ajz[2], arho[0], arho[1], arho[2], arho[3]};
#pragma unroll
            for (int k = 0; k < 13; k++) v[k] = warp_sum(v[k]);
            double mine = 0.0;
            double* dst = nullptr;
#pragma unroll
            for (int k = 0; k < 13; k++) {
                if (lane == k) {
                    mine = v[k];
                    dst = k < 3 ? g.jx + j - 1 + k : (k < 6 ? g.jy + j - 1 + (k - 3) : (k < 9 ? g.jz + j - 1 + (k - 6) : g.rho + j - 2 + (k - 9)));
                }
            }
            if (dst && mine != 0.0) atomicAdd(dst, mine);
        }
    }
    if (cfl) atomicOr(&st->err, ERR_CFL);
    if (cap_err) atomicOr(&st->err, ERR_CAPACITY | ERR_CAP_MOVERS);
}

#ifdef OPAL_FAST
#include "opal_push_lean.cuh"
#include "opal_push_full.cuh"
#endif  // OPAL_FAST

// ============================================================================================
// K2: scatter of boundary-crossing particles: local movers of K1, migrants received from a
//     neighbour slab, or the retry list after a re-layout.  Routes slab leavers (periodic wrap /
//     outbox / delete: mod.rs:290-373), appends to the destination bucket, and deposits (the
//     movers' share of yee.rs:551-641) when the push was fused.
// ============================================================================================
template <int SP, bool FULL, bool DEP>
__global__ void __launch_bounds__(256) k_scatter(GridDev g, BucketDev b, StepParams sp, const opal_migrant* __restrict__ recs,
                                                  const int* __restrict__ seg_count, int nseg, int seg_cap, long long cell_shift,
                                                  opal_migrant* __restrict__ out_l, opal_migrant* __restrict__ out_r,
                                                  opal_migrant* __restrict__ overflow, StatusDev* __restrict__ st) {
    if (g.halt && *(volatile const int*)g.halt) return;   // burst mode: an earlier kernel needs the host first
    const double q = species_charge(SP, b.z_star);
    for (int seg = blockIdx.x; seg < nseg; seg += gridDim.x) {
        int n = seg_count[seg];
        if (n > seg_cap) n = seg_cap;
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            opal_migrant m = recs[(long long)seg * seg_cap + i];
            long long cell = m.cell + cell_shift;
            if (cell < 0) {
                atomicAdd(&st->gone[0], 1);
                if (sp.route_left == 1) cell += b.ncell;                       // mod.rs:313-315 + 366-368
                else if (sp.route_left == 2) {
                    int r = atomicAdd(&st->n_out[0], 1);
                    if (r < sp.out_cap) { m.cell = cell; out_l[r] = m; } else atomicOr(&st->err, ERR_CAPACITY | ERR_CAP_OUTBOX);
                    continue;
                } else continue;                                               // open boundary: deleted by the splice
            } else if (cell >= b.ncell) {
                atomicAdd(&st->gone[1], 1);
                if (sp.route_right == 1) cell -= b.ncell;                      // mod.rs:291-293
                else if (sp.route_right == 2) {
                    int r = atomicAdd(&st->n_out[1], 1);
                    if (r < sp.out_cap) { m.cell = cell - b.ncell; out_r[r] = m; } else atomicOr(&st->err, ERR_CAPACITY | ERR_CAP_OUTBOX);
                    continue;
                } else continue;
            }
            const int c = (int)cell;
            const int cap = (int)(b.start[c + 1] - b.start[c]);
            const int slot = atomicAdd(b.count + c, 1);
            if (slot >= cap) {
                // destination bucket full: park the record; the host re-lays the buckets out and
                // re-runs this kernel on the parked records (count[c] keeps the demand)
                int r = atomicAdd(&st->n_overflow, 1);
                if (r < sp.overflow_cap) { m.cell = cell; overflow[r] = m; } else atomicOr(&st->err, ERR_CAPACITY | ERR_CAP_PARKED);
                continue;
            }
            const long long s = b.start[c] + slot;
            b.f[F_X][s] = m.x; b.f[F_UX][s] = m.u[0]; b.f[F_UY][s] = m.u[1]; b.f[F_UZ][s] = m.u[2]; b.f[F_W][s] = m.weight;
            if (FULL) {
                b.f[F_PREVX][s] = m.prev_x; b.f[F_GAMMA][s] = SP == OPAL_ION ? m.pad[0] : m.gamma; b.f[F_CHI][s] = m.chi;
                b.f[F_TAU][s] = m.tau; b.f[F_WORK][s] = m.work; b.f[F_Y][s] = m.y; b.f[F_Z][s] = m.z;
                if (SP == OPAL_ELECTRON && b.eq_n && m.tau < 0.0) {        // an emitter of this step (see BucketDev::eq_slot)
                    const int k = atomicAdd(b.eq_n, 1);
                    if (k >= 0 && k < b.eq_cap) { b.eq_slot[k] = s; b.eq_cell[k] = c; }
                }
            }
            if (DEP && m.pad[1] == 0.0) {       // pad[1] == 1: the lean push that handed this record over deposited it already
                const double Q = m.weight * q;
                const double vy = OPAL_C * m.u[1] / m.gamma, vz = OPAL_C * m.u[2] / m.gamma;
                deposit_atomic(g, c + g.left, Q, m.prev_x, m.x, vy, vz);
            }
        }
    }
}

// ============================================================================================
// K4: stand-alone deposit of a whole species (yee.rs:551-641) for the call-by-call API and for
//     steps where emission sits between push and deposit.  Needs the full state (prev_x, gamma).
// ============================================================================================
template <int SP>
__global__ void __launch_bounds__(BLOCK) k_deposit(GridDev g, BucketDev b) {
    if (burst_blocked(g)) return;                         // burst mode: an earlier kernel needs the host first
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * WARPS_PER_BLOCK;
    const double q = species_charge(SP, b.z_star);
    for (int c = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5); c < b.ncell; c += warps_total) {
        const int n = b.count[c];
        if (n == 0) continue;
        const long long base = b.start[c];
        const int j = c + g.left;
        double acc[15];
#pragma unroll
        for (int k = 0; k < 15; k++) acc[k] = 0.0;
        for (int i = lane; i < n; i += 32) {
            const long long s = base + i;
            PState p;
            p.x = b.f[F_X][s]; p.prev_x = b.f[F_PREVX][s]; p.w = b.f[F_W][s];
            p.u = {b.f[F_UX][s], b.f[F_UY][s], b.f[F_UZ][s]}; p.gamma = b.f[F_GAMMA][s];
            const V3 v = SP == OPAL_ELECTRON ? electron_velocity(p) : ion_velocity(p);
            double t[15];
            deposit_terms(p.w * q, p.prev_x, p.x, v.y, v.z, g.dx, g.dt, t);
#pragma unroll
            for (int k = 0; k < 15; k++) acc[k] += t[k];
        }
#pragma unroll
        for (int k = 0; k < 15; k++) acc[k] = warp_sum(acc[k]);
        double mine = 0.0;
        double* dst = nullptr;
#pragma unroll
        for (int k = 0; k < 15; k++) {
            if (lane == k) {
                mine = acc[k];
                dst = k < 5 ? g.jx + j - 2 + k : (k < 8 ? g.jy + j - 1 + (k - 5) : (k < 11 ? g.jz + j - 1 + (k - 8) : g.rho + j - 2 + (k - 11)));
            }
        }
        if (dst && mine != 0.0) atomicAdd(dst, mine);
    }
}

// ============================================================================================
// EXTENSION (OPAL_FLAG_PAIR_CREATION; no counterpart in the reference -- parity unpinned): nonlinear Breit-Wheeler pair
// creation.  W = alpha_f F(chi) / (tau_C k0); dW/df and F from tools/gen_bw_tables.py (Baier-Katkov / Ritus).
// ============================================================================================
#include "bw_tables.inc"
__constant__ double c_bw_ln_f[BW_NCHI] = BW_LN_F_INIT;
__constant__ double c_bw_cdf[BW_NCHI][BW_NF] = BW_CDF_INIT;

// F(chi) = chi T(chi): tabulated for 0.05 <= chi <= 500, its asymptotes outside
__device__ double bw_big_f(double chi) {
    if (!(chi > 0.0)) return 0.0;
    const double lc = log(chi);
    if (lc < BW_LN_CHI0) return 0.22963966338592295 * chi * exp(-8.0 / (3.0 * chi));          // 3 sqrt(3) / (16 sqrt(2)) chi e^(-8/3chi)
    const double idx = (lc - BW_LN_CHI0) / BW_DLN_CHI;
    if (idx >= (double)(BW_NCHI - 1)) return exp(c_bw_ln_f[BW_NCHI - 1] + (2.0 / 3.0) * (lc - (BW_LN_CHI0 + (BW_NCHI - 1) * BW_DLN_CHI)));   // ~ chi^(2/3)
    const int i = (int)idx;
    const double w = idx - (double)i;
    return exp((1.0 - w) * c_bw_ln_f[i] + w * c_bw_ln_f[i + 1]);
}
// the electron's share f in (0, 1/2] of the photon energy with P(f' < f) = r at the table's chi node `row`
__device__ double bw_invert_row(int row, double r) {
    int j = 0;
    while (j < BW_NF - 2 && c_bw_cdf[row][j + 1] <= r) j++;
    const double lo = c_bw_cdf[row][j], hi = c_bw_cdf[row][j + 1];
    const double frac = hi > lo ? (r - lo) / (hi - lo) : 0.5;
    return ((double)j + fmin(fmax(frac, 0.0), 1.0)) / (2.0 * (BW_NF - 1));
}
// f in (0, 1): r1 picks the share below 1/2 (interpolated between the two neighbouring chi nodes), r2 which lepton gets it
__device__ double bw_sample_f(double chi, double r1, double r2) {
    double idx = (log(chi) - BW_LN_CHI0) / BW_DLN_CHI;
    idx = fmin(fmax(idx, 0.0), (double)(BW_NCHI - 1) - 1.0e-9);
    const int i = (int)idx;
    const double w = idx - (double)i;
    const double f = (1.0 - w) * bw_invert_row(i, r1) + w * bw_invert_row(i + 1, r1);
    return r2 < 0.5 ? f : 1.0 - f;
}
__global__ void k_bw_rate(int n, const double* chi, double* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = bw_big_f(chi[i]);
}
__global__ void k_bw_sample(int n, const double* chi, const double* r1, const double* r2, double* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = bw_sample_f(chi[i], r1[i], r2[i]);
}

// ============================================================================================
// Photons: flat explicit-cell store (they are never deposited, so no bucketing is needed);
// push = photon.rs:150-183, out-of-place with warp-aggregated compaction of survivors.
// ============================================================================================
__global__ void __launch_bounds__(256) k_photon_push(GridDev g, PhotonDev src, PhotonDev dst, StepParams sp,
                                                      opal_migrant* __restrict__ out_l, opal_migrant* __restrict__ out_r,
                                                      StatusDev* __restrict__ st, PairDev pr) {
    if (burst_blocked(g)) return;                         // burst mode: an earlier kernel needs the host first
    const int n = *src.n;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    bool cfl = false, bad_index = false;
    const int stride = gridDim.x * blockDim.x;
    const int n_round = (n + 31) & ~31;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        bool keep = false;
        PhState p;
        long long cell = 0;
        double tau = 0.0, birth = 0.0, tau_abs = 0.0, tau_st = 0.0;
        if (i < n) {
            cell = src.cell[i];
            p.x = src.f[P_X][i]; p.prev_x = src.f[P_PREVX][i]; p.k = {src.f[P_KX][i], src.f[P_KY][i], src.f[P_KZ][i]};
            p.w = src.f[P_W][i]; p.chi = src.f[P_CHI][i]; p.y = src.f[P_Y][i]; p.z = src.f[P_Z][i];
            tau = src.f[P_TAU][i]; birth = src.f[P_BIRTH][i]; tau_abs = src.f[P_TAU_ABS][i]; tau_st = src.f[P_TAU_ST][i];
            if (cell < 0 || cell >= g.nx) { bad_index = true; }
            else {
                CellFields cf;
                load_cell_fields(g, (int)cell + g.left, cf);
                V3 E, B;
                fields_at(cf, p.x, E, B);
                cell += photon_push(p, E, B, g.dx, g.dt, cfl);
                keep = true;
                int route = -1;
                if (cell < 0) { atomicAdd(&st->gone[0], 1); route = sp.route_left; if (route == 1) cell += g.nx; }
                else if (cell >= g.nx) { atomicAdd(&st->gone[1], 1); route = sp.route_right; if (route == 1) cell -= g.nx; }
                if (route == 0) keep = false;
                if (route == 2) {
                    keep = false;
                    const int side = cell < 0 ? 0 : 1;
                    int r = atomicAdd(&st->n_out[side], 1);
                    if (r < sp.out_cap) {
                        opal_migrant m;
                        m.cell = side == 0 ? cell : cell - g.nx;
                        m.x = p.x; m.prev_x = p.prev_x; m.u[0] = p.k.x; m.u[1] = p.k.y; m.u[2] = p.k.z; m.weight = p.w;
                        m.gamma = sqrt(dot(p.k, p.k)); m.chi = p.chi; m.tau = tau; m.work = 0.0; m.y = p.y; m.z = p.z;
                        m.birth_time = birth; m.pad[0] = tau_abs; m.pad[1] = tau_st;
                        (side == 0 ? out_l : out_r)[r] = m;
                    } else atomicOr(&st->err, ERR_CAPACITY);
                }
            }
        }
        if (pr.n && keep) {
            // ---- Breit-Wheeler (extension): the optical depth against pair creation runs down; at zero the photon becomes a pair ----
            const double k0 = sqrt(dot(p.k, p.k));
            tau -= (OPAL_ALPHA_FINE / OPAL_COMPTON_TIME) * bw_big_f(p.chi) / k0 * g.dt;
            if (tau < 0.0) {
                const int slot = atomicAdd(pr.n, 1);
                if (slot < pr.cap) {
                    double r1, r2, r3, r4;
                    philox_uniform2(pr.seed, (uint64_t)i, pr.call, 20u, r1, r2);
                    philox_uniform2(pr.seed, (uint64_t)i, pr.call, 21u, r3, r4);
                    const double f = bw_sample_f(p.chi, r1, r2);
                    const V3 ue = f * p.k, up = p.k - ue;                       // momentum is conserved exactly: u_e + u_p = k
                    opal_migrant m;
                    m.cell = cell; m.x = p.x; m.prev_x = p.x; m.weight = p.w; m.chi = 0.0; m.work = 0.0; m.y = p.y; m.z = p.z;
                    m.birth_time = 0.0; m.pad[1] = 0.0;
                    m.u[0] = ue.x; m.u[1] = ue.y; m.u[2] = ue.z; m.gamma = sqrt(1.0 + dot(ue, ue)); m.tau = -log1p(-r3); m.pad[0] = 0.0;
                    pr.e_recs[slot] = m;
                    const double up2 = dot(up, up), gm1 = up2 / (1.0 + sqrt(1.0 + up2));                 // ion.rs:83-85
                    m.u[0] = up.x; m.u[1] = up.y; m.u[2] = up.z; m.gamma = 1.0 + gm1; m.pad[0] = gm1; m.tau = CUDART_INF; (void)r4;
                    pr.p_recs[slot] = m;
                    keep = false;
                }                                                               // (no room left this call: the photon converts next step)
            }
        }
        const unsigned mk = __ballot_sync(FULLMASK, keep);
        int base = 0;
        if (mk) {
            const int leader = __ffs(mk) - 1;
            if (lane == leader) base = atomicAdd(dst.n, __popc(mk));
            base = __shfl_sync(FULLMASK, base, leader);
        }
        if (sp.route_left == 2 || sp.route_right == 2) {      // a neighbour slab: who sits in the edge cells now can leave next step
            const unsigned e0 = __ballot_sync(FULLMASK, keep && cell == 0), e1 = __ballot_sync(FULLMASK, keep && cell == g.nx - 1);
            if (lane == 0) { if (e0) atomicAdd(&st->edge[0], __popc(e0)); if (e1) atomicAdd(&st->edge[1], __popc(e1)); }
        }
        if (keep) {
            const int o = base + __popc(mk & lt_mask);
            dst.cell[o] = (int)cell;
            dst.f[P_X][o] = p.x; dst.f[P_PREVX][o] = p.prev_x; dst.f[P_KX][o] = p.k.x; dst.f[P_KY][o] = p.k.y; dst.f[P_KZ][o] = p.k.z;
            dst.f[P_W][o] = p.w; dst.f[P_CHI][o] = p.chi; dst.f[P_TAU][o] = tau; dst.f[P_Y][o] = p.y; dst.f[P_Z][o] = p.z;
            dst.f[P_BIRTH][o] = birth; dst.f[P_TAU_ABS][o] = tau_abs; dst.f[P_TAU_ST][o] = tau_st;
        }
    }
    if (cfl) atomicOr(&st->err, ERR_CFL);
    if (bad_index) atomicOr(&st->err, ERR_INDEX);
}

__global__ void k_photon_insert(PhotonDev dst, const opal_migrant* __restrict__ recs, const int* __restrict__ n_recs, int n_max,
                                long long cell_shift, StatusDev* __restrict__ st) {
    int n = *n_recs;
    if (n > n_max) n = n_max;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const opal_migrant m = recs[i];
        const int o = atomicAdd(dst.n, 1);
        if (o >= dst.cap) { atomicOr(&st->err, ERR_CAPACITY | ERR_CAP_PHOTONS); continue; }
        dst.cell[o] = (int)(m.cell + cell_shift);
        dst.f[P_X][o] = m.x; dst.f[P_PREVX][o] = m.prev_x; dst.f[P_KX][o] = m.u[0]; dst.f[P_KY][o] = m.u[1]; dst.f[P_KZ][o] = m.u[2];
        dst.f[P_W][o] = m.weight; dst.f[P_CHI][o] = m.chi; dst.f[P_TAU][o] = m.tau; dst.f[P_Y][o] = m.y; dst.f[P_Z][o] = m.z;
        dst.f[P_BIRTH][o] = m.birth_time; dst.f[P_TAU_ABS][o] = m.pad[0]; dst.f[P_TAU_ST][o] = m.pad[1];
    }
}

// ============================================================================================
// QED: quantum-synchrotron sampling (qed/photon_emission.rs:87-292, qed/pwmci.rs) and
// Electron::radiate + emit_radiation (electron.rs:208-251, interactions.rs:45-107)
// ============================================================================================
typedef const double (*Table2)[2];
struct Fit { double x0, x1, f0, f1, m0, m1; };

__device__ Fit fit_construct(int i, Table2 t, int len) {   // pwmci.rs:14-68
    double s0, s1, s2;
    if (i == 1) {
        s0 = (t[i][1] - t[i - 1][1]) / (t[i][0] - t[i - 1][0]); s1 = s0;
        s2 = (t[i + 1][1] - t[i][1]) / (t[i + 1][0] - t[i][0]);
    } else if (i == len - 1) {
        s0 = (t[i - 1][1] - t[i - 2][1]) / (t[i - 1][0] - t[i - 2][0]);
        s1 = (t[i][1] - t[i - 1][1]) / (t[i][0] - t[i - 1][0]); s2 = s1;
    } else {
        s0 = (t[i - 1][1] - t[i - 2][1]) / (t[i - 1][0] - t[i - 2][0]);
        s1 = (t[i][1] - t[i - 1][1]) / (t[i][0] - t[i - 1][0]);
        s2 = (t[i + 1][1] - t[i][1]) / (t[i + 1][0] - t[i][0]);
    }
    double m0 = (s0 * s1 > 0.0) ? 0.5 * (s0 + s1) : 0.0;
    double m1 = (s1 * s2 > 0.0) ? 0.5 * (s1 + s2) : 0.0;
    if (s1 != 0.0) { double a = m0 / s1; if (a > 3.0) m0 = 3.0 * s1; }
    if (s2 != 0.0) { double bb = m1 / s2; if (bb > 3.0) m1 = 3.0 * s2; }
    return Fit{t[i - 1][0], t[i][0], t[i - 1][1], t[i][1], m0, m1};
}
__device__ double fit_eval(const Fit& p, double x) {        // pwmci.rs:70-77
    double t = (x - p.x0) / (p.x1 - p.x0);
    double h00 = (1.0 + 2.0 * t) * ((1.0 - t) * (1.0 - t));
    double h10 = t * ((1.0 - t) * (1.0 - t));
    double h01 = (t * t) * (3.0 - 2.0 * t);
    double h11 = (t * t) * (t - 1.0);
    return p.f0 * h00 + p.f1 * h01 + (p.x1 - p.x0) * (p.m0 * h10 + p.m1 * h11);
}
__device__ bool is_between(double x, double mn, double mx) {   // pwmci.rs:86-94
    if (mn < mx) return x > mn && x < mx;
    else if (mx < mn) return x > mx && x < mn;
    return true;
}
__device__ bool pwmci_evaluate(double x, Table2 t, int len, double* out) {   // pwmci.rs:96-111
    int i = 0;
    while (i < len && x > t[i][0]) i++;
    if (i >= len - 1 && x > t[len - 1][0]) return false;
    if (i == 0) return false;
    Fit p = fit_construct(i, t, len);
    *out = fit_eval(p, x);
    return true;
}
__device__ bool pwmci_invert(double f, Table2 t, int len, double* out) {      // pwmci.rs:113-228
    int i = 0;
    while (i < len && f > t[i][1]) i++;
    if (i >= len - 1 && f > t[len - 1][1]) return false;
    if (i == 0) return false;
    Fit fit = fit_construct(i, t, len);
    double r0x = t[i - 1][0], r0f = t[i - 1][1] - f, r1x = t[i][0], r1f = t[i][1] - f;
    if (fabs(r1f) < fabs(r0f)) { double a = r0x, bq = r0f; r0x = r1x; r0f = r1f; r1x = a; r1f = bq; }
    double r2x = r1x, r2f = r1f, r3x = 0.0;
    bool prev_bisect = false;
    for (int it = 0; it < 32; it++) {
        double s;
        if (r0f != r2f && r1f != r2f) {
            double r = r0f / r2f, s_ = r0f / r1f, tt = r1f / r2f;
            double p = s_ * (tt * (r - tt) * (r2x - r0x) - (1.0 - r) * (r0x - r1x));
            double q = (tt - 1.0) * (r - 1.0) * (s_ - 1.0);
            s = r0x + p / q;
        } else {
            s = r0x - (r0f * (r0x - r1x) / (r0f - r1f));
        }
        if (!is_between(s, 0.25 * (3.0 * r1x + r0x), r0x) || (prev_bisect && fabs(s - r0x) >= 0.5 * fabs(r0x - r2x))
            || (!prev_bisect && fabs(s - r0x) >= 0.5 * fabs(r2x - r3x))) {
            prev_bisect = true;
            s = 0.5 * (r0x + r1x);
        } else prev_bisect = false;
        r3x = r2x;
        r2x = r0x; r2f = r0f;
        r0x = s; r0f = fit_eval(fit, s) - f;
        if (r1f * r0f >= 0.0) { r1x = r0x; r1f = r0f; r0x = r2x; r0f = r2f; }
        if (fabs(r1f) < fabs(r0f)) { double a = r0x, bq = r0f; r0x = r1x; r0f = r1f; r1x = a; r1f = bq; }
        double error = fabs(r1x - r0x) / fabs(r0x);
        if (error < 1.0e-6 || r0f == 0.0) break;
    }
    *out = r0x;
    return true;
}

// photon_emission.rs:207-249
__constant__ double c_classical_spectrum[41][2] = {
    {-3.91202301e+0, -6.60517238e+0}, {-3.73932912e+0, -6.32380043e+0}, {-3.56663524e+0, -6.04325741e+0},
    {-3.39394136e+0, -5.76365479e+0}, {-3.22124748e+0, -5.48512019e+0}, {-3.04855360e+0, -5.20779983e+0},
    {-2.87585971e+0, -4.93186151e+0}, {-2.70316583e+0, -4.65749813e+0}, {-2.53047195e+0, -4.38493162e+0},
    {-2.35777807e+0, -4.11441770e+0}, {-2.18508419e+0, -3.84625132e+0}, {-2.01239030e+0, -3.58077293e+0},
    {-1.83969642e+0, -3.31837570e+0}, {-1.66700254e+0, -3.05951374e+0}, {-1.49430866e+0, -2.80471134e+0},
    {-1.32161478e+0, -2.55457314e+0}, {-1.14892089e+0, -2.30979520e+0}, {-9.76227012e-1, -2.07117645e+0},
    {-8.03533130e-1, -1.83963003e+0}, {-6.30839248e-1, -1.61619332e+0}, {-4.58145366e-1, -1.40203511e+0},
    {-2.85451484e-1, -1.19845722e+0}, {-1.12757602e-1, -1.00688674e+0}, {5.99362800e-2, -8.28853548e-1},
    {2.32630162e-1, -6.65945768e-1}, {4.05324044e-1, -5.19734006e-1}, {5.78017926e-1, -3.91654410e-1},
    {7.50711808e-1, -2.82842583e-1}, {9.23405690e-1, -1.93918722e-1}, {1.09609957e+0, -1.24743702e-1},
    {1.26879345e+0, -7.41989867e-2}, {1.44148734e+0, -4.00839889e-2}, {1.61418122e+0, -1.92451749e-2},
    {1.78687510e+0, -8.00296501e-3}, {1.95956898e+0, -2.79652691e-3}, {2.13226286e+0, -7.92727241e-4},
    {2.30495675e+0, -1.74957252e-4}, {2.47765063e+0, -2.86508630e-5}, {2.65034451e+0, -3.28906586e-6},
    {2.82303839e+0, -2.47459791e-7}, {2.99573227e+0, -1.12649296e-8}};

// a 64-double CDF row: {coeff, power, table[31][2]}  (photon_emission_tables.rs:9-13)
__device__ __forceinline__ Table2 cdf_table(const double* row) { return reinterpret_cast<Table2>(row + 2); }

__device__ double from_linear_cdf_table(double global_zero, double local_zero, double rnd, const double* cdf) {   // :87-121
    const double coeff = cdf[0], power = cdf[1];
    Table2 tb = cdf_table(cdf);
    double r_zero;
    if (local_zero < tb[0][0]) r_zero = coeff * pow(local_zero - global_zero, power);
    else if (!pwmci_evaluate(local_zero, tb, 31, &r_zero)) return local_zero;
    double r = r_zero + (1.0 - r_zero) * rnd;
    double y;
    if (r <= tb[0][1]) y = exp((log(r) - log(coeff)) / power) + global_zero;
    else if (!pwmci_invert(r, tb, 31, &y)) y = local_zero;
    return y;
}

__device__ double clamp_acos(double cos_theta) {   // f64::min(1.0).max(-1.0).acos(), NaN -> 1.0
    double ct = isnan(cos_theta) ? 1.0 : fmin(cos_theta, 1.0);
    return acos(fmax(ct, -1.0));
}

__device__ void classical_sample(double chi, double gamma, double rand1, double rand2, double rand3, double out[3]) {   // :264-292
    double delta = acos((-9.0 + 50.0 * rand2 - 25.0 * (rand2 * rand2)) / 16.0);
    double zb = (2.0 + 4.0 * cos(delta / 3.0)) / (5.0 * (1.0 - rand2));
    double z = zb * zb * zb;
    double ln_rand = log(rand1);
    double x;
    if (ln_rand < c_classical_spectrum[0][1]) x = 1.020377255 * pow(rand1, 0.6);
    else {
        double ln_x;
        if (!pwmci_invert(ln_rand, c_classical_spectrum, 41, &ln_x)) ln_x = c_classical_spectrum[40][0];
        x = exp(ln_x);
    }
    double u = 3.0 * chi * x / (2.0 * z);
    out[0] = u * gamma;
    out[1] = clamp_acos((gamma - pow(z, 2.0 / 3.0) / (2.0 * gamma)) / sqrt(gamma * gamma - 1.0));
    out[2] = 2.0 * OPAL_PI * rand3;
}

#define OPAL_LN_CHI_MIN (-4.60517018599)
#define OPAL_LN_STEP 0.230258509299
#define OPAL_LN_DELTA_MIN (-2.30258509299)

// tables: 73 rows x 64 doubles (rows 0..40 QUANTUM_CDF, 41 Y_INFINITE_DELTA_CDF, 42..72 Y_CDF)
__device__ bool quantum_sample(const double* tables, double chi, double gamma, double rand1, double rand2, double rand3, double out[3]) {   // :129-203
    double index_f = (log(chi) - OPAL_LN_CHI_MIN) / OPAL_LN_STEP;
    double w = index_f - trunc(index_f);
    double fl = floor(index_f);
    int index = fl <= 0.0 ? 0 : (fl > 1.0e6 ? 1000000 : (int)fl);
    if (log(chi) <= OPAL_LN_CHI_MIN) {
        double c[3];
        classical_sample(chi, gamma, rand1, rand2, rand3, c);
        out[0] = c[0] * gamma / (gamma + c[0]); out[1] = c[1]; out[2] = c[2];
        return true;
    } else if (index >= 40) {
        out[0] = out[1] = out[2] = CUDART_NAN;
        return false;   // unimplemented!() photon_emission.rs:144
    }
    double ln_u[2];
#pragma unroll 1
    for (int k = 0; k < 2; k++) {
        const double* row = tables + (size_t)(index + k) * 64;
        Table2 tb = cdf_table(row);
        double lr = log(rand1), inv;
        if (lr <= tb[0][1]) ln_u[k] = (lr - log(row[0])) / row[1];
        else if (pwmci_invert(lr, tb, 31, &inv)) ln_u[k] = inv;
        else ln_u[k] = tb[30][0];
    }
    double u = exp((1.0 - w) * ln_u[0] + w * ln_u[1]);
    double beta = 2.0 * u / (3.0 * chi);
    double delta = (1.0 + (1.0 + u) * (1.0 + u)) * pow(beta, -2.0 / 3.0) / (1.0 + u);
    double index2_f = (log(delta) - OPAL_LN_DELTA_MIN) / OPAL_LN_STEP;
    double w2 = index2_f - trunc(index2_f);
    double fl2 = floor(index2_f);
    int index2 = fl2 <= 0.0 ? 0 : (fl2 > 1.0e6 ? 1000000 : (int)fl2);
    double y;
    if (index2 >= 30) {
        y = from_linear_cdf_table(0.0, beta, rand2, tables + 41 * 64);
    } else {
        double gz = pow(delta, -1.5);
        double y_lower = from_linear_cdf_table(gz, beta, rand2, tables + (size_t)(42 + index2) * 64);
        double y_upper = from_linear_cdf_table(gz, beta, rand2, tables + (size_t)(43 + index2) * 64);
        y = (1.0 - w2) * y_lower + w2 * y_upper;
    }
    double z = fmax(y / beta, 1.0);
    out[0] = gamma * u / (1.0 + u);
    out[1] = clamp_acos((gamma - pow(z, 2.0 / 3.0) / (2.0 * gamma)) / sqrt(gamma * gamma - 1.0));
    out[2] = 2.0 * OPAL_PI * rand3;
    return true;
}

__device__ V3 v_normalize(V3 a) { double mag = sqrt(dot(a, a)); return a / mag; }              // vec3.rs:114
__device__ V3 v_orthogonal(V3 a) {                                                              // vec3.rs:120
    V3 perp = fabs(a.x) > fabs(a.z) ? V3{-a.y, a.x, 0.0} : V3{0.0, -a.z, a.y};
    return v_normalize(perp);
}
__device__ V3 v_rotate_around(V3 v, V3 ax, double theta) {                                      // vec3.rs:129
    double s, c;
    sincos(theta, &s, &c);
    return {(c + ax.x * ax.x * (1.0 - c)) * v.x + (ax.x * ax.y * (1.0 - c) - ax.z * s) * v.y + (ax.x * ax.z * (1.0 - c) + ax.y * s) * v.z,
            (ax.y * ax.x * (1.0 - c) + ax.z * s) * v.x + (c + ax.y * ax.y * (1.0 - c)) * v.y + (ax.y * ax.z * (1.0 - c) - ax.x * s) * v.z,
            (ax.z * ax.x * (1.0 - c) - ax.y * s) * v.x + (ax.z * ax.y * (1.0 - c) + ax.x * s) * v.y + (c + ax.z * ax.z * (1.0 - c)) * v.z};
}

constexpr int QED_TABLE_DOUBLES = 73 * 64;

// One emission (electron.rs:209-240 + the filters of interactions.rs:74-95) of the electron in bucket slot s of cell c,
// whose optical depth has run out.  Random numbers: Philox4x32-10 keyed by (seed, call index), counter = bucket slot,
// so a run is reproducible for a given bucket layout irrespective of launch geometry and of which thread does the work.
// Returns true when the photon passes the filters (m is then filled in).
__device__ __forceinline__ bool emit_one(const BucketDev& b, const StepParams& sp, const double* s_tab, long long s, int c, double t,
                                         double min_energy, double max_angle, double max_fl, uint64_t seed, uint64_t call_index,
                                         opal_migrant& m, bool& chi_err) {
    double r_tau, r1, r2, r3, r4, r5, r6, r7;
    philox_uniform2(seed, (uint64_t)s, (uint32_t)call_index, 0u, r_tau, r1);
    philox_uniform2(seed, (uint64_t)s, (uint32_t)call_index, 1u, r2, r3);
    philox_uniform2(seed, (uint64_t)s, (uint32_t)call_index, 2u, r4, r5);
    philox_uniform2(seed, (uint64_t)s, (uint32_t)call_index, 3u, r6, r7);
    b.f[F_TAU][s] = -log1p(-r_tau);                                    // Exp1 by inversion
    double chi = b.f[F_CHI][s], gamma = b.f[F_GAMMA][s];
    V3 u = {b.f[F_UX][s], b.f[F_UY][s], b.f[F_UZ][s]};
    double smp[3];
    bool ok = true;
    if (sp.flags & OPAL_FLAG_NO_RADIATION_REACTION) classical_sample(chi, gamma, r1, r2, r3, smp);
    else ok = quantum_sample(s_tab, chi, gamma, r1, r2, r3, smp);
    if (!ok) { chi_err = true; return false; }
    const double omega = smp[0], theta = smp[1], cphi = smp[2];
    V3 par = v_normalize(u);                                           // electron.rs:221-228
    V3 perp = v_rotate_around(v_orthogonal(par), par, cphi);
    V3 k;
    if (sp.flags & OPAL_FLAG_NO_BEAMING) k = omega * par;
    else k = omega * (cos(theta) * par + sin(theta) * perp);
    const double fl = 2.0 * (gamma * gamma) * theta * OPAL_C * OPAL_COMPTON_TIME / chi;   // :231
    if (!(sp.flags & OPAL_FLAG_NO_RADIATION_REACTION)) {                // :234-239
        u = u - k;
        double ng = sqrt(1.0 + dot(u, u));
        b.f[F_UX][s] = u.x; b.f[F_UY][s] = u.y; b.f[F_UZ][s] = u.z;
        b.f[F_CHI][s] = chi * ng / gamma;
        b.f[F_GAMMA][s] = ng;
    }
    // filters of interactions.rs:74-95 (NaN == None)
    bool e_ok = isnan(min_energy) || sqrt(dot(k, k)) * OPAL_ME_MEV >= min_energy;
    bool a_ok = true;
    if (!isnan(max_angle)) {
        V3 pm = OPAL_ME_MEV * k;
        double mag = sqrt(pm.x * pm.x + pm.y * pm.y + pm.z * pm.z);
        a_ok = acos(-pm.x / mag) <= max_angle;
    }
    bool f_ok = isnan(max_fl) || fl < max_fl;
    if (!(e_ok && a_ok && f_ok)) return false;
    m.cell = c; m.x = b.f[F_X][s]; m.prev_x = CUDART_NAN;              // Photon::create with dx = dt = 0
    m.u[0] = k.x; m.u[1] = k.y; m.u[2] = k.z; m.weight = b.f[F_W][s];
    m.gamma = sqrt(dot(k, k)); m.chi = 0.0; m.tau = -log1p(-r4); m.work = 0.0; m.y = 0.0; m.z = 0.0;
    m.birth_time = t; m.pad[0] = -log1p(-r5); m.pad[1] = -log1p(-r6); (void)r7;   // tau_abs, tau_st (photon.rs:130-131)
    return true;
}

// append the photons of a (converged) warp to the flat photon store with one atomic per warp
__device__ __forceinline__ void emit_append(const PhotonDev& ph, bool store_photon, const opal_migrant& m, StatusDev* __restrict__ st) {
    const int lane = threadIdx.x & 31;
    const unsigned mk = __ballot_sync(FULLMASK, store_photon);
    if (!mk) return;
    int o = 0;
    const int leader = __ffs(mk) - 1;
    if (lane == leader) { o = atomicAdd(ph.n, __popc(mk)); atomicAdd(&st->n_new_photons, __popc(mk)); }
    o = __shfl_sync(FULLMASK, o, leader);
    if (store_photon) {
        o += __popc(mk & ((1u << lane) - 1u));
        if (o < ph.cap) {
            ph.cell[o] = (int)m.cell;
            ph.f[P_X][o] = m.x; ph.f[P_PREVX][o] = m.prev_x; ph.f[P_KX][o] = m.u[0]; ph.f[P_KY][o] = m.u[1]; ph.f[P_KZ][o] = m.u[2];
            ph.f[P_W][o] = m.weight; ph.f[P_CHI][o] = 0.0; ph.f[P_TAU][o] = m.tau; ph.f[P_Y][o] = 0.0; ph.f[P_Z][o] = 0.0;
            ph.f[P_BIRTH][o] = m.birth_time; ph.f[P_TAU_ABS][o] = m.pad[0]; ph.f[P_TAU_ST][o] = m.pad[1];
        } else atomicOr(&st->err, ERR_CAPACITY);
    }
}

#ifndef OPAL_EMIT_MB
#define OPAL_EMIT_MB 4              // blocks per SM.  Measured (tools/bench_qed.py, 2e7 electrons, 10 % radiating per step), whole
                                    // step: 1 -> 14.2 ms, 2 (128 registers, no spills) -> 11.2, 3 -> 12.5, 4 (64 registers, 0.7 KB
                                    // of spills) -> 10.2: the optical-depth scan wants the resident warps more than the sampling
                                    // wants the registers
#endif
#ifndef OPAL_EMIT_QCAP
#define OPAL_EMIT_QCAP 1024         // queued emitters per block: 12 KB beside the 37.4 KB of tables
#endif
constexpr int EMIT_QCAP = OPAL_EMIT_QCAP;

// K7: emission tables staged in shared memory (37.4 KB); new photons appended with a warp-aggregated atomic.
// Only a few per cent of the electrons radiate in a step, so sampling where the optical depth is found spent would
// leave ~3 lanes of 32 busy through the whole table inversion.  Instead every block works in rounds: (1) one warp per
// cell reads the optical depths and queues the (slot, cell) of the spent ones in shared memory, (2) once a block's
// worth has collected, all its threads sample one queued electron each.  A warp that finds the queue full samples its
// electrons in place (every electron radiating: nothing to compact).  Which thread samples an electron does not
// change its random numbers (keyed by slot); the order of the photon store is arbitrary as before.
__global__ void __launch_bounds__(BLOCK, OPAL_EMIT_MB) k_emit(GridDev g, BucketDev b, PhotonDev ph, StepParams sp, const double* __restrict__ tables_g,
                                                    double t, double min_energy, double max_angle, double max_fl,
                                                    uint64_t seed, uint64_t call_index, StatusDev* __restrict__ st, long long room_need, int room_phase) {
    if (burst_blocked(g)) return;                         // burst mode: an earlier kernel needs the host first
    if (room_need > 0) {
        // room for one photon per emitter (electron.rs:208-251)?  The emitter list, when valid, says exactly how many.
        long long need = room_need;
        if (b.eq_n) { const int q = *b.eq_n; if (q >= 0 && q <= b.eq_cap) need = q; }
        if ((long long)*ph.n + need > (long long)ph.cap) {                                  // the same in every thread: k_emit has not appended yet
            if (blockIdx.x == 0 && threadIdx.x == 0) { st->halt_species = room_phase; __threadfence(); *(volatile int*)&st->halt = 1; }
            return;
        }
    }
    if (g.step) {                                         // burst mode: step k of the burst -> its time and its random-number call index
        const int k = *(volatile const int*)g.step;
        t = g.laser_arr[3 * k + 2];
        call_index += (uint64_t)k;
    }
    const int qn = b.eq_n ? *b.eq_n : -1;
    const bool listed = qn >= 0 && qn <= b.eq_cap;        // the push and the scatter of this step listed the emitters
    if (listed && (long long)blockIdx.x * blockDim.x >= qn) return;      // nothing for this block (before the beam meets the pulse: for any block)
    extern __shared__ double s_tab[];
    long long* s_qs = reinterpret_cast<long long*>(s_tab + QED_TABLE_DOUBLES);
    int* s_qc = reinterpret_cast<int*>(s_qs + EMIT_QCAP);
    __shared__ int s_qn;
    for (int i = threadIdx.x; i < QED_TABLE_DOUBLES; i += blockDim.x) s_tab[i] = tables_g[i];
    if (threadIdx.x == 0) s_qn = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int warps_total = gridDim.x * WARPS_PER_BLOCK;
    bool chi_err = false;
    if (listed) {
        // sample exactly the listed electrons, every lane busy
        const int n_round = (qn + 31) & ~31;
        for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n_round; j += gridDim.x * blockDim.x) {      // whole warps: emit_append votes
            bool store_photon = false;
            opal_migrant m;
            if (j < qn) store_photon = emit_one(b, sp, s_tab, b.eq_slot[j], b.eq_cell[j], t, min_energy, max_angle, max_fl, seed, call_index, m, chi_err);
            emit_append(ph, store_photon, m, st);
        }
        if (chi_err) atomicOr(&st->err, ERR_CHI);
        return;
    }
    for (int c0 = blockIdx.x * WARPS_PER_BLOCK; c0 < b.ncell; c0 += warps_total) {     // block-uniform trip count
        const int c = c0 + (threadIdx.x >> 5);
        const int n = c < b.ncell ? b.count[c] : 0;
        const long long base = c < b.ncell ? b.start[c] : 0;
        for (int k0 = 0; k0 < n; k0 += 32) {
            const int i = k0 + lane;
            const long long s = base + i;
            const bool spent = i < n && b.f[F_TAU][s] < 0.0;                          // electron.rs:209
            const unsigned mk = __ballot_sync(FULLMASK, spent);
            if (!mk) continue;
            int o = 0;
            if (lane == 0) o = atomicAdd(&s_qn, __popc(mk));
            o = __shfl_sync(FULLMASK, o, 0) + __popc(mk & lt_mask);
            bool store_photon = false;
            opal_migrant m;
            if (spent) {
                if (o < EMIT_QCAP) { s_qs[o] = s; s_qc[o] = c; }
                else store_photon = emit_one(b, sp, s_tab, s, c, t, min_energy, max_angle, max_fl, seed, call_index, m, chi_err);
            }
            if (__shfl_sync(FULLMASK, o, 31 - __clz(mk)) >= EMIT_QCAP) emit_append(ph, store_photon, m, st);   // warp-uniform
        }
        __syncthreads();
        const int nq = min(s_qn, EMIT_QCAP);
        const bool last = c0 + warps_total >= b.ncell;
        const int take = last ? nq : (nq / BLOCK) * BLOCK;                             // whole blocks' worth off the tail
        for (int j0 = nq - take; j0 < nq; j0 += BLOCK) {
            const int j = j0 + threadIdx.x;
            bool store_photon = false;
            opal_migrant m;
            if (j < nq) store_photon = emit_one(b, sp, s_tab, s_qs[j], s_qc[j], t, min_energy, max_angle, max_fl, seed, call_index, m, chi_err);
            emit_append(ph, store_photon, m, st);
        }
        __syncthreads();
        if (threadIdx.x == 0) s_qn = nq - take;
        __syncthreads();
    }
    if (chi_err) atomicOr(&st->err, ERR_CHI);
}

__global__ void k_qed_rate(int n, const double* chi, const double* gamma, double* out) {
    __shared__ double s_ln_h[41];
    if (threadIdx.x < 41) s_ln_h[threadIdx.x] = c_ln_h[threadIdx.x];
    __syncthreads();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = qed_rate(chi[i], gamma[i], s_ln_h);
}
__global__ void k_qed_sample(int n, const double* __restrict__ tables_g, const double* chi, const double* gamma, const double* r1,
                             const double* r2, const double* r3, int classical, double* out3, StatusDev* st) {
    extern __shared__ double s_tab[];
    for (int i = threadIdx.x; i < QED_TABLE_DOUBLES; i += blockDim.x) s_tab[i] = tables_g[i];
    __syncthreads();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double o[3];
    bool ok = true;
    if (classical) classical_sample(chi[i], gamma[i], r1[i], r2[i], r3[i], o);
    else ok = quantum_sample(s_tab, chi[i], gamma[i], r1[i], r2[i], r3[i], o);
    if (!ok) atomicOr(&st->err, ERR_CHI);
    out3[3 * i] = o[0]; out3[3 * i + 1] = o[1]; out3[3 * i + 2] = o[2];
}
__global__ void k_fields_at(GridDev g, int n, const long long* cell, const double* xi, double* out6, StatusDev* st) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    long long c = cell[i];
    if (c < 0 || c >= g.nx) { atomicOr(&st->err, ERR_INDEX); return; }     // yee.rs:501-502
    CellFields cf;
    load_cell_fields(g, (int)c + g.left, cf);
    V3 E, B;
    fields_at(cf, xi[i], E, B);
    out6[6 * i] = E.x; out6[6 * i + 1] = E.y; out6[6 * i + 2] = E.z; out6[6 * i + 3] = B.x; out6[6 * i + 4] = B.y; out6[6 * i + 5] = B.z;
}

// ============================================================================================
// Grid: halo overlay + boundary conditions (yee.rs:351-496) and the fused FDTD update
// (yee.rs:345-349, 839-867)
// ============================================================================================
__device__ __forceinline__ double cell_x(const GridDev& g, int c) { return g.xmin + ((double)c - (double)g.left) * g.dx; }   // yee.rs:296

// one block.  halo_recv: [2][8][11] cells received from the left / right neighbour (remote case).
__device__ __forceinline__ void sync_bc_body(const GridDev& g, const double* __restrict__ halo_recv, double laser_y, double laser_z,
                                             const double* __restrict__ laser_dev) {       // one block of 256 threads
    __shared__ double s_send[2][OPAL_HALO_CELLS][10];   // copies of my first / last 8 cells (yee.rs:353-354)
    const int G = OPAL_GHOST_SIZE;
    double* arr[10] = {g.rho, g.jx, g.jy, g.jz, g.ex, g.ey, g.ez, g.bx, g.by, g.bz};
    const int tid = threadIdx.x;
    if (tid < 160) {
        const int side = tid / 80, i = (tid % 80) / 10, k = tid % 10;
        const int src = side == 0 ? i : g.len - 2 * G + i;
        s_send[side][i][k] = arr[k][src];
    }
    __syncthreads();
    if (tid < 160) {
        const int side = tid / 80, i = (tid % 80) / 10, k = tid % 10;   // side: which of MY edges is updated
        const bool internal = side == 0 ? g.left_bdy == OPAL_BDY_INTERNAL : g.right_bdy == OPAL_BDY_INTERNAL;
        if (internal) {
            // what the neighbour on `side` sent: self-neighbour => my opposite edge (yee.rs:369,380)
            double v = g.self_nbr ? s_send[1 - side][i][k] : halo_recv[(side * OPAL_HALO_CELLS + i) * OPAL_CELL_DOUBLES + 1 + k];
            const int dst = side == 0 ? i : g.len - 2 * G + i;
            // left edge: cells [0,4) overlay_ghost, [4,8) overlay; right edge: [len-8,len-4) overlay, [len-4,len) overlay_ghost
            const bool ghost = side == 0 ? i < G : i >= G;
            if (k < 4) arr[k][dst] += v;                  // rho, j: += (yee.rs:97-113)
            else if (ghost) arr[k][dst] = v;              // E, B: overwritten in ghost cells only
        }
    }
    __syncthreads();
    if (g.left_bdy == OPAL_BDY_LASER && tid == 0) {       // yee.rs:456-462
        if (laser_dev) { laser_y = laser_dev[0]; laser_z = laser_dev[1]; }
        if (g.laser_arr) { const int k = *(volatile const int*)g.step; laser_y = g.laser_arr[3 * k]; laser_z = g.laser_arr[3 * k + 1]; }
        const int start = g.left - 2;
        const double r = OPAL_C * g.dt / g.dx;
        g.ey[start] += 2.0 * r * laser_y;
        g.ez[start] += 2.0 * r * laser_z;
    }
    if (g.right_bdy == OPAL_BDY_ABSORBING) {              // yee.rs:464-479
        const double xmin = cell_x(g, g.nx + g.left), xmax = cell_x(g, g.nx + g.left + g.right - 1);
        const double sigma_max = 10.0 / (double)g.right;
        for (int i = g.len - g.right + 1 + tid; i < g.len; i += blockDim.x) {
            const double sigma = sigma_max * (cell_x(g, i) - xmin) / (xmax - xmin);
            const bool zero = i >= g.len - 2;
#pragma unroll
            for (int k = 4; k < 10; k++) arr[k][i] = zero ? 0.0 : (1.0 - sigma) * arr[k][i];
        }
    } else if (g.right_bdy == OPAL_BDY_CONDUCTING) {      // yee.rs:480-495
        const int start = g.len - g.right;
        if (tid == 0) { g.ex[start] = 0.0; g.by[start] = 0.0; g.bz[start] = 0.0; }
        if (tid >= 1 && tid < g.right) {
            const int i = tid;
            g.ex[start + i] = -g.ex[start - i];
            g.ey[start + i] = g.ey[start + 1 - i];
            g.ez[start + i] = g.ez[start + 1 - i];
            g.bx[start + i] = g.bx[start + 1 - i];
            g.by[start + i] = -g.by[start - i];
            g.bz[start + i] = -g.bz[start - i];
        }
    }
}

// Conducting mirror reads cell[start].{ey,ez,bx} (i = 1) which thread 0 does not modify, and
// cell[start - i] for i >= 1 which nobody writes: no intra-kernel hazard.
__global__ void __launch_bounds__(256) k_sync_bc(GridDev g, const double* __restrict__ halo_recv, double laser_y, double laser_z,
                                                  const double* __restrict__ laser_dev) {
    if (burst_blocked(g)) return;                         // burst mode: an earlier kernel needs the host first
    sync_bc_body(g, halo_recv, laser_y, laser_z, laser_dev);
}

constexpr int FT = 256;   // field tile
// B += dt/2, E += dt (Silver-Mueller on cell 0), B += dt/2 in ONE pass: each block recomputes the
// one-cell halo it needs (identical arithmetic to the owning block) and writes its tile to the
// other buffer of the E/B ping-pong, so no block ever reads a value another block has updated.
__device__ __forceinline__ void field_tile(const GridDev& g, const GridDev& o, const int t0) {     // FT threads
    __shared__ double s_ey[FT + 3], s_ez[FT + 3], s_by[FT + 2], s_bz[FT + 2];
    const int tid = threadIdx.x;
    __syncthreads();                                      // (a block that walks several tiles re-uses the staging arrays)
    const int len = g.len;
    const double hdt = 0.5 * g.dt;
    // the sources and E_x of the E step below: asked for now, so that their latency overlaps the staging (a small grid is a few
    // blocks, and this kernel is then a chain of memory round trips).  Thread 0 also does the one cell beyond the tile.
    const int ia = t0 + tid, ib = t0 + FT;
    double jx_a = 0.0, jy_a = 0.0, jz_a = 0.0, ex_a = 0.0, jy_b = 0.0, jz_b = 0.0;
    if (ia < len) { jx_a = g.jx[ia]; jy_a = g.jy[ia]; jz_a = g.jz[ia]; ex_a = g.ex[ia]; }
    if (tid == 0 && ib < len) { jy_b = g.jy[ib]; jz_b = g.jz[ib]; }
    // stage E (t0-1 .. t0+FT+1) and B (t0-1 .. t0+FT)
    for (int k = tid; k < FT + 3; k += FT) {
        const int i = t0 - 1 + k;
        const bool in = i >= 0 && i < len;
        s_ey[k] = in ? g.ey[i] : 0.0;
        s_ez[k] = in ? g.ez[i] : 0.0;
        if (k < FT + 2) { s_by[k] = in ? g.by[i] : 0.0; s_bz[k] = in ? g.bz[i] : 0.0; }
    }
    __syncthreads();
    // first half B step for i in [t0-1, t0+FT], valid while i < len-1 (yee.rs:839-848)
    for (int k = tid; k < FT + 2; k += FT) {
        const int i = t0 - 1 + k;
        if (i >= 0 && i < len - 1) {
            s_by[k] += hdt * (s_ez[k + 1] - s_ez[k]) / g.dx;
            s_bz[k] += hdt * (s_ey[k] - s_ey[k + 1]) / g.dx;
        }
    }
    __syncthreads();
    // E step for i in [t0, t0+FT] (yee.rs:852-867); E index k = i - (t0-1)
    double ex_new = 0.0;
    for (int k = 1 + tid; k < FT + 2; k += FT) {
        const int i = t0 - 1 + k;
        if (i < len) {
            if (i == 0) {
                const double kappa = 2.0 * OPAL_C * g.dt / (OPAL_C * g.dt + g.dx);
                const double sigma = 1.0 - kappa;
                if (k <= FT) ex_new = 0.0;
                const double ey = sigma * s_ey[k] - OPAL_C * kappa * s_bz[k];
                const double ez = sigma * s_ez[k] + OPAL_C * kappa * s_by[k];
                s_ey[k] = ey; s_ez[k] = ez;
            } else {
                const bool own = k <= FT;                  // i == ia (this thread's cell of the tile); else i == ib (thread 0)
                if (own) ex_new = ex_a + -g.dt * jx_a / OPAL_EPS0;
                const double ey = s_ey[k] + (g.dt * OPAL_C_SQD * (s_bz[k - 1] - s_bz[k]) / g.dx - g.dt * (own ? jy_a : jy_b) / OPAL_EPS0);
                const double ez = s_ez[k] + (g.dt * OPAL_C_SQD * (s_by[k] - s_by[k - 1]) / g.dx - g.dt * (own ? jz_a : jz_b) / OPAL_EPS0);
                s_ey[k] = ey; s_ez[k] = ez;
            }
        }
    }
    __syncthreads();
    // second half B step for own tile and write-out
    {
        const int k = 1 + tid;
        const int i = t0 + tid;
        if (i < len) {
            double by = s_by[k], bz = s_bz[k];
            if (i < len - 1) {
                by += hdt * (s_ez[k + 1] - s_ez[k]) / g.dx;
                bz += hdt * (s_ey[k] - s_ey[k + 1]) / g.dx;
            }
            o.ex[i] = ex_new; o.ey[i] = s_ey[k]; o.ez[i] = s_ez[k];
            o.bx[i] = g.bx[i]; o.by[i] = by; o.bz[i] = bz;
        }
    }
}
__global__ void __launch_bounds__(FT) k_field(GridDev g, GridDev o) {
    if (burst_blocked(g)) return;                         // burst mode: an earlier kernel needs the host first
    field_tile(g, o, blockIdx.x * FT);
    if (g.bst && g.count_step && blockIdx.x == 0 && threadIdx.x == 0) g.bst->step = g.bst->step + 1;   // (read by the NEXT step's kernels only)
}

// bursts: is a mover segment more than half full?  (block-wide; every thread of the block must call it)
__device__ __forceinline__ bool segments_filling(const StatusDev* st, int mover_cap, int tid, int nthreads) {
    const int* seg = reinterpret_cast<const int*>(st + 1);
    bool any = false;
    for (int i = tid; i < OPAL_MOVER_SEGMENTS; i += nthreads) any |= 2 * seg[i] > mover_cap;
    return __syncthreads_or(any);
}

// The grid's share of a step in ONE launch (graph-replayed bursts on a single slab): the check of the
// last particle advance (bucket overflow / assert -> halt), halo overlay + boundary conditions (yee.rs:351-496, block 0),
// a grid-wide barrier, B/2 - E - B/2 (yee.rs:345-349) over the tiles, and the step counter.  At most 148 blocks of FT
// threads with < 9 KB of shared memory each: all resident at once, so the spin barrier cannot starve (it still gives up
// after ~1 s and raises ERR_INDEX rather than hang the device).
// clear_next: o.rho / o.j* are the OTHER set of source arrays (the host alternates two, like E and B): the next step deposits
// there, so this launch also does that step's Grid::clear (yee.rs:543) and begin-mark -- nobody reads that set now -- and a
// replayed step of one lean species is three launches (push, scatter, this).
__global__ void __launch_bounds__(FT) k_grid_step(GridDev g, GridDev o, StatusDev* st, int last_species, unsigned* bar /* count, generation */,
                                                  int mover_cap, int clear_next) {
    // (all the status words in one round trip: the branch on `halt` comes after the loads of the others have been issued)
    const int halt = *(volatile const int*)&st->halt, n_overflow = *(volatile const int*)&st->n_overflow;
    const unsigned err = *(volatile const unsigned*)&st->err;
    const bool filling = segments_filling(st, mover_cap, threadIdx.x, blockDim.x);
    if (halt) return;
    if (n_overflow > 0 || err || filling) {               // the same for every block: nobody advances the fields
        if (blockIdx.x == 0 && threadIdx.x == 0) { st->halt_species = last_species; __threadfence(); st->halt = 1; }
        return;
    }
    if (blockIdx.x == 0) { sync_bc_body(g, nullptr, 0.0, 0.0, nullptr); __threadfence(); }
    __syncthreads();
    if (gridDim.x > 1) {
        if (threadIdx.x == 0) {
            const unsigned gen = *(volatile unsigned*)(bar + 1);
            if (atomicAdd(bar, 1u) == gridDim.x - 1) { *(volatile unsigned*)bar = 0u; __threadfence(); atomicAdd(bar + 1, 1u); }
            else {
                long long spins = 0;
                while (*(volatile unsigned*)(bar + 1) == gen) if (++spins > (1ll << 28)) { atomicOr(&st->err, ERR_INDEX); break; }
            }
            __threadfence();
        }
        __syncthreads();
    }
    for (int t0 = blockIdx.x * FT; t0 < g.len; t0 += gridDim.x * FT) field_tile(g, o, t0);
    if (clear_next) {
        for (int i = blockIdx.x * FT + threadIdx.x; i < g.len; i += gridDim.x * FT) { o.rho[i] = 0.0; o.jx[i] = 0.0; o.jy[i] = 0.0; o.jz[i] = 0.0; }
        if (blockIdx.x == 0) {                                  // (what k_clear does with st != null; the check above has looked at them)
            __syncthreads();                                    // every thread of this block is past segments_filling
            int* seg = reinterpret_cast<int*>(st + 1);
            for (int i = threadIdx.x; i < OPAL_MOVER_SEGMENTS; i += blockDim.x) seg[i] = 0;
            if (threadIdx.x == 0) { st->n_movers = 0; st->n_overflow = 0; st->n_out[0] = st->n_out[1] = 0; st->gone[0] = st->gone[1] = 0; st->n_new_photons = 0; }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) st->step = st->step + 1;     // (read only by the kernels of the next step)
}

// st != null (bursts): also the begin-mark of the species advance that follows (per-advance counters, mover segment counters)
__global__ void k_clear(GridDev g, StatusDev* st) {   // yee.rs:543-549
    if (g.halt && *(volatile const int*)g.halt) return;   // burst mode: an earlier kernel needs the host first
    if (st && blockIdx.x == 0) {
        int* seg = reinterpret_cast<int*>(st + 1);
        for (int i = threadIdx.x; i < OPAL_MOVER_SEGMENTS; i += blockDim.x) seg[i] = 0;
        if (threadIdx.x == 0) { st->n_movers = 0; st->n_overflow = 0; st->n_out[0] = st->n_out[1] = 0; st->gone[0] = st->gone[1] = 0; st->n_new_photons = 0; }
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < g.len; i += gridDim.x * blockDim.x) {
        g.rho[i] = 0.0; g.jx[i] = 0.0; g.jy[i] = 0.0; g.jz[i] = 0.0;
    }
}

__global__ void k_halo_pack(GridDev g, double* __restrict__ out) {   // [2][8][11]
    const int tid = threadIdx.x;
    if (tid >= 2 * OPAL_HALO_CELLS * OPAL_CELL_DOUBLES) return;
    const int side = tid / (OPAL_HALO_CELLS * OPAL_CELL_DOUBLES), i = (tid / OPAL_CELL_DOUBLES) % OPAL_HALO_CELLS, k = tid % OPAL_CELL_DOUBLES;
    const int src = side == 0 ? i : g.len - OPAL_HALO_CELLS + i;
    const double* arr[10] = {g.rho, g.jx, g.jy, g.jz, g.ex, g.ey, g.ez, g.bx, g.by, g.bz};
    out[tid] = k == 0 ? cell_x(g, src) : arr[k - 1][src];
}

// ============================================================================================
// Diagnostics: em_field_energy (yee.rs:787-799) and total_kinetic_energy (mod.rs:227-240)
// ============================================================================================
__device__ __forceinline__ void block_atomic_sum(double v, double* dst) {
    __shared__ double s_part[32];
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) s_part[w] = v;
    __syncthreads();
    if (w == 0) {
        v = lane < (blockDim.x >> 5) ? s_part[lane] : 0.0;
        v = warp_sum(v);
        if (lane == 0 && v != 0.0) atomicAdd(dst, v);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256) k_em_energy(GridDev g, double* out) {
    double e2 = 0.0, b2 = 0.0;
    for (int i = g.left + blockIdx.x * blockDim.x + threadIdx.x; i < g.len - g.right; i += gridDim.x * blockDim.x) {
        e2 = e2 + g.ex[i] * g.ex[i] + g.ey[i] * g.ey[i] + g.ez[i] * g.ez[i];
        b2 = b2 + g.bx[i] * g.bx[i] + g.by[i] * g.by[i] + g.bz[i] * g.bz[i];
    }
    // 0.5 * (eps0 * sum_e2 + sum_b2 / mu0) * dx is linear in the partial sums
    block_atomic_sum(0.5 * (OPAL_EPS0 * e2 + b2 / OPAL_MU0) * g.dx, out);
}

template <int SP, bool FULL>
__global__ void __launch_bounds__(BLOCK) k_kinetic(BucketDev b, double* out) {
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * WARPS_PER_BLOCK;
    double acc = 0.0;
    for (int c = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5); c < b.ncell; c += warps_total) {
        const int n = b.count[c];
        const long long base = b.start[c];
        for (int i = lane; i < n; i += 32) {
            const long long s = base + i;
            const V3 u = {b.f[F_UX][s], b.f[F_UY][s], b.f[F_UZ][s]};
            const double w = b.f[F_W][s];
            const double u2 = dot(u, u);
            if (SP == OPAL_ELECTRON) {                                    // electron.rs:122-126
                const double gamma = FULL ? b.f[F_GAMMA][s] : sqrt(1.0 + u2);
                const double to_joules = 1.0e6 * OPAL_ME_MEV * OPAL_E;
                acc += w * u2 * to_joules / (gamma + 1.0);
            } else {                                                      // ion.rs:128-134
                const double gm1 = FULL ? b.f[F_GAMMA][s] : u2 / (1.0 + sqrt(1.0 + u2));
                const double energy = gm1 * (b.mass_number * OPAL_MP / OPAL_ME) * OPAL_ME_MEV;
                acc += w * energy * 1.0e6 * OPAL_E;
            }
        }
    }
    block_atomic_sum(acc, out);
}

// How many particles will change cell in the next push, to first order (velocity before the push): the mover buffer is
// sized from this (ensure_movers in opal_host.cu).  v_x = c u_x / sqrt(1 + u^2) for electrons and ions alike.
__global__ void __launch_bounds__(BLOCK) k_crossing_estimate(BucketDev b, double dt_over_dx, double* out2) {
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * WARPS_PER_BLOCK;
    double leaving = 0.0, expected = 0.0;
    for (int c = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5); c < b.ncell; c += warps_total) {
        const int n = b.count[c];
        const long long base = b.start[c];
        for (int i = lane; i < n; i += 32) {
            const long long s = base + i;
            const V3 u = {b.f[F_UX][s], b.f[F_UY][s], b.f[F_UZ][s]};
            const double step = OPAL_C * u.x * rsqrt(1.0 + dot(u, u)) * dt_over_dx;
            const double x = b.f[F_X][s] + step;
            if (!(x >= 0.0 && x < 1.0)) leaving += 1.0;
            expected += fmin(1.0, fabs(step));
        }
    }
    block_atomic_sum(leaving, out2);
    block_atomic_sum(expected, out2 + 1);
}

__global__ void __launch_bounds__(256) k_photon_energy(PhotonDev ph, double* out) {   // photon.rs:220-226
    const int n = *ph.n;
    double acc = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const V3 k = {ph.f[P_KX][i], ph.f[P_KY][i], ph.f[P_KZ][i]};
        const double energy = sqrt(dot(k, k)) * OPAL_ME_MEV;
        acc += ph.f[P_W][i] * energy * 1.0e6 * OPAL_E;
    }
    block_atomic_sum(acc, out);
}

// ============================================================================================
// Synthetic plasma generator, download gather, re-layout copy
// ============================================================================================
template <bool FULL>
__global__ void __launch_bounds__(BLOCK) k_generate(BucketDev b, long long npc, double weight_, double uthx, double uthy, double uthz,
                                                     double dx_, double dy_, double dz_, int two_stream, long long cell0, uint64_t seed,
                                                     int species, double gdx, double gdt) {
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * WARPS_PER_BLOCK;
    for (int c = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5); c < b.ncell; c += warps_total) {
        const long long base = b.start[c];
        for (int i = lane; i < npc; i += 32) {
            const uint64_t gid = (uint64_t)(cell0 + c) * (uint64_t)npc + (uint64_t)i;
            double x, urand, a0, a1, a2, a3;
            philox_uniform2(seed, gid, (uint32_t)species, 0u, x, urand);
            philox_uniform2(seed, gid, (uint32_t)species, 1u, a0, a1);
            philox_uniform2(seed, gid, (uint32_t)species, 2u, a2, a3);
            // Box-Muller on (0,1]
            const double r0 = sqrt(-2.0 * log(1.0 - a0)), r1 = sqrt(-2.0 * log(1.0 - a2));
            double s0, c0, s1, c1;
            sincospi(2.0 * a1, &s0, &c0);
            sincospi(2.0 * a3, &s1, &c1);
            const double n0 = r0 * c0, n1 = r0 * s0, n2 = r1 * c1;
            double ux;
            if (two_stream) ux = dx_ * (1.0 + 0.001 * n0) * (2.0 * floor(urand - 0.5) + 1.0);   // two_stream.yaml:18
            else ux = dx_ + uthx * n0;
            const double uy = dy_ + uthy * n1, uz = dz_ + uthz * n2;
            const long long s = base + i;
            b.f[F_X][s] = x; b.f[F_UX][s] = ux; b.f[F_UY][s] = uy; b.f[F_UZ][s] = uz; b.f[F_W][s] = weight_;
            if (FULL) {
                const double u2 = ux * ux + uy * uy + uz * uz;
                double gamma, prev_x;
                if (species == OPAL_ION) {                                                      // ion.rs:81-101
                    gamma = u2 / (1.0 + sqrt(1.0 + u2));
                    prev_x = x - (OPAL_C * ux / (1.0 + gamma)) * gdt / gdx;
                } else {                                                                        // electron.rs:85-103
                    gamma = sqrt(1.0 + u2);
                    prev_x = x - OPAL_C * ux * gdt / (gamma * gdx);
                }
                b.f[F_PREVX][s] = prev_x; b.f[F_GAMMA][s] = gamma; b.f[F_CHI][s] = 0.0; b.f[F_TAU][s] = CUDART_INF;
                b.f[F_WORK][s] = 0.0; b.f[F_Y][s] = 0.0; b.f[F_Z][s] = 0.0;
            }
        }
        if (lane == 0) b.count[c] = (int)npc;
    }
}

struct OutPtrs { double* f[NF_FULL]; };
template <bool FULL>
__global__ void __launch_bounds__(BLOCK) k_gather_out(BucketDev b, const long long* __restrict__ off, long long* __restrict__ cell, OutPtrs out) {
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * WARPS_PER_BLOCK;
    for (int c = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5); c < b.ncell; c += warps_total) {
        const int n = b.count[c];
        const long long base = b.start[c], o = off[c];
        for (int i = lane; i < n; i += 32) {
            if (cell) cell[o + i] = c;
#pragma unroll
            for (int k = 0; k < (FULL ? NF_FULL : NF_LEAN); k++)
                if (out.f[k]) out.f[k][o + i] = b.f[k][base + i];
        }
    }
}

template <bool FULL>
__global__ void __launch_bounds__(BLOCK) k_copy_buckets(BucketDev src, BucketDev dst) {
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * WARPS_PER_BLOCK;
    for (int c = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5); c < src.ncell; c += warps_total) {
        const int cap_old = (int)(src.start[c + 1] - src.start[c]);
        int n = src.count[c];
        if (n > cap_old) n = cap_old;     // count may carry the demand of parked (overflowed) movers
        const long long bs = src.start[c], bd = dst.start[c];
        for (int i = lane; i < n; i += 32) {
#pragma unroll
            for (int k = 0; k < (FULL ? NF_FULL : NF_LEAN); k++) dst.f[k][bd + i] = src.f[k][bs + i];
        }
        if (lane == 0) dst.count[c] = n;
    }
}

// ============================================================================================
// launchers
// ============================================================================================
static inline int cell_grid(int ncell) {
    long long blocks = ((long long)ncell + WARPS_PER_BLOCK - 1) / WARPS_PER_BLOCK;
    if (blocks < 1) blocks = 1;
    if (blocks > (1 << 22)) blocks = 1 << 22;
    return (int)blocks;
}

static bool l_push(cudaStream_t s, const GridDev& g, const BucketDev& b, const StepParams& sp, bool full, bool dep,
                   opal_migrant* movers, const LeanStrips* ls, opal_migrant* out_l, opal_migrant* out_r, opal_migrant* overflow,
                   StatusDev* st, double* ke_out) {
    // the per-segment counters live directly behind the StatusDev block (see host: status + seg counters)
    int* seg = reinterpret_cast<int*>(st + 1);
    const int grid = cell_grid(b.ncell);
#ifdef OPAL_FAST
    if ((sp.species == OPAL_ELECTRON || sp.species == OPAL_ION) && !full && dep && ls && sp.cell_mover_cap > 0) {
        long long blocks = ((long long)(b.c_end - b.c_begin) + OPAL_PUSH_WPB - 1) / OPAL_PUSH_WPB;
        const int gr = (int)(blocks < 1 ? 1 : blocks);
        const int smem = OPAL_PUSH_WPB * (int)sizeof(PushScratch);
        // (the attribute is per device and contexts may live on several devices: set on every launch, a cheap host-side call)
        // cells above sp.heavy_min: skipped by the first launch, worked on block by block by the second (two blocks per SM)
        const bool heavy = sp.heavy_min < INT_MAX;
        const int smem_h = OPAL_HEAVY_WARPS * (int)sizeof(PushScratch) + (int)sizeof(HeavyShared);
        const int gr_h = (int)std::min<long long>(2 * 148, ((long long)b.ncell + 7) / 8);
#define LAUNCH2(B0, K, SPC, KE_OUT) do { \
        cudaFuncSetAttribute(k_push_lean_fast<B0, K, SPC, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); \
        k_push_lean_fast<B0, K, SPC, false><<<gr, OPAL_PUSH_WPB * 32, smem, s>>>(g, b, sp, movers, seg, ls->in, ls->scount_in, ls->out, ls->scount_out, ls->hcount, out_l, out_r, overflow, st, KE_OUT); \
        if (heavy) { \
            cudaFuncSetAttribute(k_push_lean_fast<B0, K, SPC, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_h); \
            k_push_lean_fast<B0, K, SPC, true><<<gr_h, OPAL_HEAVY_WARPS * 32, smem_h, s>>>(g, b, sp, movers, seg, ls->in, ls->scount_in, ls->out, ls->scount_out, ls->hcount, out_l, out_r, overflow, st, KE_OUT); \
        } } while (0)
#define LAUNCHF(B0, K) LAUNCH2(B0, K, OPAL_ELECTRON, ke_out)
#define LAUNCHI(B0) LAUNCH2(B0, false, OPAL_ION, nullptr)
        if (sp.species == OPAL_ION) { if (sp.flags & OPAL_STEP_BX_ZERO) LAUNCHI(true); else LAUNCHI(false); }
        else if (sp.flags & OPAL_STEP_BX_ZERO) { if (ke_out) LAUNCHF(true, true); else LAUNCHF(true, false); }
        else { if (ke_out) LAUNCHF(false, true); else LAUNCHF(false, false); }
#undef LAUNCHF
#undef LAUNCHI
#undef LAUNCH2
        return true;
    }
    if (sp.species == OPAL_ELECTRON && full) {
        // the reference's whole per-particle state in the reciprocal arithmetic (opal_push_full.cuh)
        long long blocks = ((long long)b.ncell + OPAL_FULL_WPB - 1) / OPAL_FULL_WPB;
        const int gr = (int)(blocks < 1 ? 1 : blocks);
        const int smem = OPAL_FULL_WPB * (int)sizeof(FullScratch);
#define LAUNCHQ(B0, D) do { cudaFuncSetAttribute(k_push_full_fast<B0, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); \
        k_push_full_fast<B0, D><<<gr, OPAL_FULL_WPB * 32, smem, s>>>(g, b, sp, movers, seg, st); } while (0)
        if (sp.flags & OPAL_STEP_BX_ZERO) { if (dep) LAUNCHQ(true, true); else LAUNCHQ(true, false); }
        else { if (dep) LAUNCHQ(false, true); else LAUNCHQ(false, false); }
#undef LAUNCHQ
        return false;
    }
#endif
    (void)ls; (void)out_l; (void)out_r; (void)overflow; (void)ke_out;
#define LAUNCH(SP, F, D) k_push<SP, F, D><<<grid, BLOCK, 0, s>>>(g, b, sp, movers, seg, st)
    if (sp.species == OPAL_ELECTRON) {
        if (full) { if (dep) LAUNCH(OPAL_ELECTRON, true, true); else LAUNCH(OPAL_ELECTRON, true, false); }
        else { if (dep) LAUNCH(OPAL_ELECTRON, false, true); else LAUNCH(OPAL_ELECTRON, false, false); }
    } else {
        if (full) { if (dep) LAUNCH(OPAL_ION, true, true); else LAUNCH(OPAL_ION, true, false); }
        else { if (dep) LAUNCH(OPAL_ION, false, true); else LAUNCH(OPAL_ION, false, false); }
    }
#undef LAUNCH
    return false;
}

// append the arrivals waiting in the strips to their buckets (lean layout; see opal_push_lean.cuh)
static void l_flush(cudaStream_t s, const BucketDev& b, const StepParams& sp, const StripRec* strips, const int* scount,
                    opal_migrant* overflow, StatusDev* st) {
#ifdef OPAL_FAST
    long long blocks = ((long long)b.ncell + 15) / 16;
    k_strip_flush<<<(int)(blocks < 1 ? 1 : blocks), 256, 0, s>>>(b, sp, strips, scount, overflow, st);
#endif
}

// close the free slots the lean push leaves in the buckets (count[] := live particles)
static void l_compact_cells(cudaStream_t s, const BucketDev& b, int* hcount) {
#ifdef OPAL_FAST
    long long blocks = ((long long)b.ncell + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_compact_cells<<<(int)(blocks < 1 ? 1 : blocks), 256, 0, s>>>(b, hcount);
#endif
}

static void l_scatter(cudaStream_t s, const GridDev& g, const BucketDev& b, const StepParams& sp, bool full, bool dep,
                      const opal_migrant* recs, const int* n_recs, int n_max, long long cell_shift, const int* idx_list,
                      opal_migrant* out_l, opal_migrant* out_r, int* overflow_idx, StatusDev* st) {
    // n_recs == nullptr: the segmented mover buffer of l_push (counters behind st); else one segment of n_max records
    (void)idx_list;
    const int* seg = n_recs ? n_recs : reinterpret_cast<const int*>(st + 1);
    const int nseg = n_recs ? 1 : OPAL_MOVER_SEGMENTS;
    const int seg_cap = n_recs ? n_max : sp.mover_cap;
    opal_migrant* ovf = reinterpret_cast<opal_migrant*>(overflow_idx);
    const int grid = n_recs ? 1 : OPAL_MOVER_SEGMENTS;
#define LAUNCH(SP, F, D) k_scatter<SP, F, D><<<grid, 256, 0, s>>>(g, b, sp, recs, seg, nseg, seg_cap, cell_shift, out_l, out_r, ovf, st)
    if (sp.species == OPAL_ELECTRON) {
        if (full) { if (dep) LAUNCH(OPAL_ELECTRON, true, true); else LAUNCH(OPAL_ELECTRON, true, false); }
        else { if (dep) LAUNCH(OPAL_ELECTRON, false, true); else LAUNCH(OPAL_ELECTRON, false, false); }
    } else {
        if (full) { if (dep) LAUNCH(OPAL_ION, true, true); else LAUNCH(OPAL_ION, true, false); }
        else { if (dep) LAUNCH(OPAL_ION, false, true); else LAUNCH(OPAL_ION, false, false); }
    }
#undef LAUNCH
}

static void l_deposit(cudaStream_t s, const GridDev& g, const BucketDev& b, const StepParams& sp, StatusDev*) {
    const int grid = cell_grid(b.ncell);
    if (sp.species == OPAL_ELECTRON) k_deposit<OPAL_ELECTRON><<<grid, BLOCK, 0, s>>>(g, b);
    else k_deposit<OPAL_ION><<<grid, BLOCK, 0, s>>>(g, b);
}

static void l_photon_push(cudaStream_t s, const GridDev& g, const PhotonDev& src, const PhotonDev& dst, const StepParams& sp,
                          opal_migrant* out_l, opal_migrant* out_r, StatusDev* st, const PairDev& pairs) {
    int grid = (src.cap + 255) / 256;
    if (grid > 148 * 8) grid = 148 * 8;
    if (grid < 1) grid = 1;
    k_photon_push<<<grid, 256, 0, s>>>(g, src, dst, sp, out_l, out_r, st, pairs);
}
static void l_photon_insert(cudaStream_t s, const PhotonDev& dst, const opal_migrant* recs, const int* n_recs, int n_max,
                            long long cell_shift, StatusDev* st) {
    int grid = (n_max + 255) / 256;
    if (grid > 148) grid = 148;
    if (grid < 1) grid = 1;
    k_photon_insert<<<grid, 256, 0, s>>>(dst, recs, n_recs, n_max, cell_shift, st);
}

static void l_emit(cudaStream_t s, const GridDev& g, const BucketDev& e, const PhotonDev& ph, const StepParams& sp,
                   const double* tables, double t, double min_energy, double max_angle, double max_fl, uint64_t seed,
                   uint64_t call_index, StatusDev* st, long long burst_room_need, int room_phase) {
    const int smem = QED_TABLE_DOUBLES * (int)sizeof(double) + EMIT_QCAP * (int)(sizeof(long long) + sizeof(int));
    // per DEVICE, and contexts may live on several devices of one process: set it on every launch (a cheap host-side call)
    cudaFuncSetAttribute(k_emit, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int grid = cell_grid(e.ncell);
    if (grid > 148 * OPAL_EMIT_MB) grid = 148 * OPAL_EMIT_MB;
    k_emit<<<grid, BLOCK, smem, s>>>(g, e, ph, sp, tables, t, min_energy, max_angle, max_fl, seed, call_index, st, burst_room_need, room_phase);
}

// burst-mode bookkeeping (one block): what == 0 zero the per-advance counters and the mover segment counters,
// 1 raise halt when the advance of `species` parked movers or tripped an assert, 2 count the completed step
// (mover_cap: room per mover segment.  A segment more than half full also halts the burst -- without an error -- so that the
//  host doubles the buffer before a later step overflows it, as the ordinary path does after every advance)
// what == 3: the check of `species` followed by the begin-mark of the next one, in one launch
__global__ void __launch_bounds__(1024) k_burst_mark(StatusDev* st, int what, int species, int mover_cap, int* zero_a, int* zero_b) {
    if (*(volatile int*)&st->halt) return;
    if (what == 3) {
        const bool filling = segments_filling(st, mover_cap, threadIdx.x, blockDim.x);
        if (st->n_overflow > 0 || st->err || filling) {
            if (threadIdx.x == 0) { st->halt_species = species; __threadfence(); st->halt = 1; }
            return;
        }
        __syncthreads();
        what = 0;
    }
    if (what == 0) {
        if (threadIdx.x == 0) { if (zero_a) *zero_a = 0; if (zero_b) *zero_b = 0; }
        int* seg = reinterpret_cast<int*>(st + 1);
        for (int i = threadIdx.x; i < OPAL_MOVER_SEGMENTS; i += blockDim.x) seg[i] = 0;
        if (threadIdx.x == 0) { st->n_movers = 0; st->n_overflow = 0; st->n_out[0] = st->n_out[1] = 0; st->gone[0] = st->gone[1] = 0; st->n_new_photons = 0; }
    } else if (what == 1) {
        const bool filling = segments_filling(st, mover_cap, threadIdx.x, blockDim.x);
        if (threadIdx.x == 0 && (st->n_overflow > 0 || st->err || filling)) { st->halt_species = species; __threadfence(); st->halt = 1; }
    } else if (threadIdx.x == 0) {
        st->step = st->step + 1;
    }
}
static void l_burst_mark(cudaStream_t s, StatusDev* st, int what, int species, int mover_cap, int* zero_a, int* zero_b) {
    k_burst_mark<<<1, what == 2 ? 32 : 1024, 0, s>>>(st, what, species, mover_cap, zero_a, zero_b);
}

static void l_sync_bc(cudaStream_t s, const GridDev& g, const double* halo_recv, double ly, double lz, const double* laser_dev) {
    k_sync_bc<<<1, 256, 0, s>>>(g, halo_recv, ly, lz, laser_dev);
}
static void l_field(cudaStream_t s, const GridDev& cur, const GridDev& nxt, bool) {
    k_field<<<(cur.len + FT - 1) / FT, FT, 0, s>>>(cur, nxt);
}
static void l_clear(cudaStream_t s, const GridDev& g, StatusDev* st_begin_mark) {
    int grid = (g.len + 255) / 256;
    if (grid > 148 * 8) grid = 148 * 8;
    k_clear<<<grid, 256, 0, s>>>(g, st_begin_mark);
}
static void l_grid_step(cudaStream_t s, const GridDev& cur, const GridDev& nxt, StatusDev* st, int last_species, unsigned* bar, int mover_cap, bool clear_next) {
    int grid = (cur.len + FT - 1) / FT;
    if (grid > 148) grid = 148;
    k_grid_step<<<grid, FT, 0, s>>>(cur, nxt, st, last_species, bar, mover_cap, clear_next ? 1 : 0);
}
static void l_halo_pack(cudaStream_t s, const GridDev& g, double* out) { k_halo_pack<<<1, 192, 0, s>>>(g, out); }

static void l_energy(cudaStream_t s, const GridDev& g, const BucketDev* e, const BucketDev* ions, const PhotonDev* ph, bool full, double* out4) {
    cudaMemsetAsync(out4, 0, 4 * sizeof(double), s);
    int grid = (g.nx + 255) / 256;
    if (grid > 148 * 4) grid = 148 * 4;
    k_em_energy<<<grid, 256, 0, s>>>(g, out4);
    if (e) {
        int gr = cell_grid(e->ncell); if (gr > 148 * 8) gr = 148 * 8;
        if (full) k_kinetic<OPAL_ELECTRON, true><<<gr, BLOCK, 0, s>>>(*e, out4 + 1); else k_kinetic<OPAL_ELECTRON, false><<<gr, BLOCK, 0, s>>>(*e, out4 + 1);
    }
    if (ions) {
        int gr = cell_grid(ions->ncell); if (gr > 148 * 8) gr = 148 * 8;
        if (full) k_kinetic<OPAL_ION, true><<<gr, BLOCK, 0, s>>>(*ions, out4 + 2); else k_kinetic<OPAL_ION, false><<<gr, BLOCK, 0, s>>>(*ions, out4 + 2);
    }
    if (ph) {
        int gr = (ph->cap + 255) / 256; if (gr > 148 * 4) gr = 148 * 4; if (gr < 1) gr = 1;
        k_photon_energy<<<gr, 256, 0, s>>>(*ph, out4 + 3);
    }
}

static void l_generate(cudaStream_t s, const BucketDev& b, int species, long long npc, double weight_, const double* uth, const double* drift,
                       int two_stream, long long cell0, uint64_t seed, bool full, double dx, double dt) {
    const int grid = cell_grid(b.ncell);
    if (full) k_generate<true><<<grid, BLOCK, 0, s>>>(b, npc, weight_, uth[0], uth[1], uth[2], drift[0], drift[1], drift[2], two_stream, cell0, seed, species, dx, dt);
    else k_generate<false><<<grid, BLOCK, 0, s>>>(b, npc, weight_, uth[0], uth[1], uth[2], drift[0], drift[1], drift[2], two_stream, cell0, seed, species, dx, dt);
}

static void l_gather_out(cudaStream_t s, const BucketDev& b, bool full, int, const long long* off, long long* cell, double* const* outf) {
    OutPtrs o;
    for (int k = 0; k < NF_FULL; k++) o.f[k] = outf[k];
    const int grid = cell_grid(b.ncell);
    if (full) k_gather_out<true><<<grid, BLOCK, 0, s>>>(b, off, cell, o);
    else k_gather_out<false><<<grid, BLOCK, 0, s>>>(b, off, cell, o);
}

static void l_qed_rate(cudaStream_t s, int n, const double* chi, const double* gamma, double* out) {
    k_qed_rate<<<(n + 127) / 128, 128, 0, s>>>(n, chi, gamma, out);
}
static void l_qed_sample(cudaStream_t s, int n, const double* tables, const double* chi, const double* gamma, const double* r1,
                         const double* r2, const double* r3, int classical, double* out3, StatusDev* st) {
    const int smem = QED_TABLE_DOUBLES * (int)sizeof(double);
    cudaFuncSetAttribute(k_qed_sample, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);      // per device: see l_emit
    k_qed_sample<<<(n + 127) / 128, 128, smem, s>>>(n, tables, chi, gamma, r1, r2, r3, classical, out3, st);
}
static void l_bw_rate(cudaStream_t s, int n, const double* chi, double* out) { k_bw_rate<<<(n + 127) / 128, 128, 0, s>>>(n, chi, out); }
static void l_bw_sample(cudaStream_t s, int n, const double* chi, const double* r1, const double* r2, double* out) {
    k_bw_sample<<<(n + 127) / 128, 128, 0, s>>>(n, chi, r1, r2, out);
}
static void l_fields_at(cudaStream_t s, const GridDev& g, int n, const long long* cell, const double* xi, double* out6, StatusDev* st) {
    k_fields_at<<<(n + 127) / 128, 128, 0, s>>>(g, n, cell, xi, out6, st);
}
static void l_copy_buckets(cudaStream_t s, const BucketDev& src, const BucketDev& dst, bool full) {
    const int grid = cell_grid(src.ncell);
    if (full) k_copy_buckets<true><<<grid, BLOCK, 0, s>>>(src, dst);
    else k_copy_buckets<false><<<grid, BLOCK, 0, s>>>(src, dst);
}

static void l_crossing_estimate(cudaStream_t s, const BucketDev& b, double dt_over_dx, double* out2) {
    cudaMemsetAsync(out2, 0, 2 * sizeof(double), s);
    int gr = cell_grid(b.ncell); if (gr > 148 * 8) gr = 148 * 8;
    k_crossing_estimate<<<gr, BLOCK, 0, s>>>(b, dt_over_dx, out2);
}

KernelTable table() {
    KernelTable t;
    t.crossing_estimate = l_crossing_estimate;
    t.push = l_push; t.flush = l_flush; t.scatter = l_scatter; t.deposit = l_deposit; t.photon_push = l_photon_push; t.photon_insert = l_photon_insert;
    t.emit = l_emit; t.sync_bc = l_sync_bc; t.field = l_field; t.grid_step = l_grid_step; t.clear = l_clear; t.halo_pack = l_halo_pack; t.energy = l_energy;
    t.generate = l_generate; t.gather_out = l_gather_out; t.qed_rate = l_qed_rate; t.qed_sample = l_qed_sample;
    t.fields_at = l_fields_at; t.bw_rate = l_bw_rate; t.bw_sample = l_bw_sample; t.copy_buckets = l_copy_buckets; t.compact_cells = l_compact_cells; t.burst_mark = l_burst_mark;
    return t;
}

}  // namespace OPAL_KNS
}  // namespace opal
