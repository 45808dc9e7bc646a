// opal_push_lean.cuh -- the headline kernel: lean (x, u, w) layout, fused gather + push + deposit + cell hand-over.
// Included by opal_kernels.cu inside namespace opal::fast (fast build only).
//
// One warp per cell, every step:
//   (1) the cell's RESIDENTS stream through a shared-memory ring filled by 1-D bulk copies (cp.async.bulk, one elected
//       lane, mbarrier completion -- no per-lane copy instructions or address arithmetic), are pushed (Vay electron.rs:268-330 /
//       Boris ion.rs:168-214) and written back TO THE SLOT THEY CAME FROM: x, u only -- w is never rewritten
//       (72 B per particle-step: read 40, write 32; the cell index is implicit);
//   (2) a resident that leaves the cell (|dxi| < 1 => cell +-1, electron.rs:313) is handed to this step's left / right
//       mover strip of the cell (48-B records: new offset, old offset, u, w), its slot is marked free (x := -1) and
//       remembered in a free list;
//   (3) the ARRIVALS -- the movers the two neighbour cells left in LAST step's strips -- are pushed by the same code
//       and dropped into the free slots (the rest is appended behind the bucket): the cell hand-over of
//       Population::advance (mod.rs:288-373) costs no separate pass over the buckets, and no kernel of its own;
//   (4) deposit (yee.rs:551-641): stayers accumulate in registers (13 sums) as they are pushed; the movers -- a few per
//       cent -- are deposited at the end of the cell FROM THE STRIP JUST WRITTEN, every lane busy, in the frame of the
//       cell they leave (8 more sums); one transposing shuffle reduction and <= 21 global atomics per CELL.
// Strips are double-buffered by step parity (read last step's, write this step's).  Between steps a bucket holds its
// residents plus a few free slots (count[c] counts slots in use; hcount[c] the free ones among them); the arrivals of
// the last step wait in the strips.  Anything other than this kernel sees the buckets only after k_strip_flush +
// k_compact_cells (host: settle).  A cell whose free slots exceed OPAL_HOLE_MAX shifts its live particles down
// itself (rare: arrivals fill the slots departures open; the excess random-walks).
// A cell of thousands of particles is one warp's serial work (~11 us per 1000): above StepParams::heavy_min a cell is left to a
// second launch of this kernel (HEAVY) in which a whole block shares it -- see HeavyShared below.
// At a slab edge: a leaver through an open boundary is dropped (mod.rs:372-373); towards a neighbour slab it goes
// straight into the outbox (the receiver deposits it in its own frame); on a periodic single slab it is an ordinary mover.
//
// Arithmetic: the order-2 spline is evaluated through R(d), the fraction of the particle's triangle to the right of a
// face at displacement d: W(node j) = R(j-1/2-x) - R(j+1/2-x), Cell::flux(x_i, x_f) = R(x_f) - R(x_i) (3e-16 against
// yee.rs:140-204); divisions by gamma, dt, dx are reciprocal multiplies (MUFU seed + one third-order step), q/dt, qc/dx,
// q/dx are applied once per cell.  Results differ from the literal formulas by re-association only (<= 1e-15), far inside
// the 1e-10 parity tolerance; the strict build keeps the literal code (k_push).
//
// Measured history (B200, 1e9 electrons / 1e6 cells, per step; DESIGN.md section 4, profiles/): round 1 shifted the
// stayers down over the movers' slots (w rewritten behind a cell's first mover: 80.2 B/particle DRAM) and appended the
// arrivals in a second kernel (k_pull, 0.56-1.0 ms): 15.7 + 0.6 ms.  This kernel: 77.9 B/particle, 324 instructions per
// 32 particles, 15.6 ms under the 1 kW power cap (1.6 GHz), 13.1 ms for one launch alone at the boost clock.  On the way:
// movers deposited inside the divergent branch through atomicAdd(double) on shared memory (CAS spin loops in SASS): 25 ms;
// through private shared-memory columns: 19.5 ms; an `if (live)` around the arithmetic instead of finite free-slot marks
// and a masked weight: +40 instructions; 16 warps per SM at 124 registers instead of 18 at 94: 17.0 ms; bulk copies of
// 256 B / 512 B / 768 B / 1 KB per stream and 2 to 6 ring stages: equal within noise beyond 3 stages.
// Also equal within noise (15.4-15.5 ms, power-capped): the five sign / zero tests and floor() moved from the FP64 pipe to integer
// compares on the high words (DSETP 26 -> 3 in SASS); the two chunks of a ring stage pushed side by side for instruction-level
// parallelism at 18 / 16 / 14 / 12 blocks per SM (96 + spills / 124 / 128 / 148 registers: 17.1 / 15.7 / 15.7 / 16.7 ms).
#pragma once

#ifndef OPAL_PUSH_STAGES
#define OPAL_PUSH_STAGES 3             // ring stages in flight per warp ...
#endif
#ifndef OPAL_PUSH_SUP
#define OPAL_PUSH_SUP 2                // ... of this many 32-particle chunks each: one bulk copy per stream moves 256 B x SUP
#endif
// One warp per block and 18 blocks per SM: the compiler then fits the kernel into 96 registers without spilling (124 when
// given 128), and 18 resident warps hide the FP64 dependency chains of the Vay push far better than 16 (measured: 17.0 ->
// 16.2 ms per 1e9-particle step; 20-24 warps gain nothing more, the part then sits on its power cap).
#ifndef OPAL_PUSH_WPB
#define OPAL_PUSH_WPB 1
#endif
#ifndef OPAL_PUSH_MINBLK
#define OPAL_PUSH_MINBLK 18
#endif
#ifndef OPAL_HOLE_MAX
#define OPAL_HOLE_MAX 48               // free slots a bucket may carry between steps before the cell compacts itself
#endif
#ifndef OPAL_PUSH_ARR
#define OPAL_PUSH_ARR 32
#endif
constexpr int PUSH_ARR = OPAL_PUSH_ARR; // arrivals per side staged in shared memory by one bulk copy (more: read from global)
constexpr int PUSH_FREE_CAP = 64;      // free-slot indices remembered per cell and step (more free slots simply stay free)

// A free slot inside a bucket carries x = -1 (valid offsets are in [0, 1)): FINITE, like everything else a lane can find in a
// bucket slot (buckets are zero-filled when allocated), so the push runs its arithmetic on every lane without a branch and
// only masks the weight.
#define OPAL_HOLE_MARK (-1.0)
__device__ __forceinline__ bool is_hole(double x) { return x < 0.0; }
__device__ __forceinline__ unsigned lanemask_lt() { unsigned m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

// ---- async-proxy helpers (sm_90+ PTX; SASS: UBLKCP / SYNCS) ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(void* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, void* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(void* bar, unsigned parity) {
    asm volatile("{\n.reg .pred p;\nOPAL_WAIT:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra OPAL_DONE;\nbra OPAL_WAIT;\nOPAL_DONE:\n}"
                 ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// one lane of the (converged) warp: the form the compiler maps to ELECT + uniform-datapath code for the bulk copies
__device__ __forceinline__ bool elect_one() {
    unsigned pred;
    asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
    return pred != 0;
}

struct RFaces { double a, b, c; };   // R at the faces -1/2, +1/2, +3/2 of the particle's cell
// R(d) of the order-2 spline at the three faces that cut the triangle of a particle at offset x in [0, 1), h = x - 1/2:
//   a = 1 - min(h,0)^2 / 2,  b = 1/2 + h - h|h| / 2,  c = max(h,0)^2 / 2      (min(h,0) = (h - |h|)/2, max(h,0) = (h + |h|)/2)
// 9 FP64 instructions, no selects (|h| is an operand modifier).
__device__ __forceinline__ RFaces r_faces(double x) {
    const double h = x - 0.5;
    const double tm = h - fabs(h), tp = h + fabs(h);         // 2 min(h,0), 2 max(h,0): exact
    RFaces r;
    r.a = fma(-0.125 * tm, tm, 1.0);
    r.b = fma(-0.5 * h, fabs(h), x);
    r.c = (0.125 * tp) * tp;
    return r;
}

// MUFU.RSQ64H / MUFU.RCP64H seeds (~2^-20 relative: they see only the high word of the operand) refined
// by ONE third-order step, error ~ e^3 < 2^-58: 5 (rsqrt, +1 for sqrt) / 3 (rcp) FP64 instructions
// instead of the ~10 of the IEEE-rounded library sequences with their slow-path branches.  Operands here are
// gamma^2-like quantities in [1, 1e300): no zero / denormal / infinity handling is needed.
__device__ __forceinline__ double rsqrt_seed(double x) { double y; asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x)); return y; }
__device__ __forceinline__ double rcp_seed(double x) { double y; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x)); return y; }
__device__ __forceinline__ double fast_rsqrt(double x) {
    const double y0 = rsqrt_seed(x);
    const double t = x * y0;
    const double e = fma(-t, y0, 1.0);             // 1 - x y0^2
    const double pe = fma(0.375, e, 0.5) * e;      // rsqrt(x) = y0 (1 + e/2 + 3 e^2/8 + O(e^3))
    return fma(y0, pe, y0);
}
__device__ __forceinline__ void fast_rsqrt_sqrt(double x, double& rs, double& sq) {
    const double y0 = rsqrt_seed(x);
    const double t = x * y0;                       // ~ sqrt(x)
    const double e = fma(-t, y0, 1.0);
    const double pe = fma(0.375, e, 0.5) * e;
    rs = fma(y0, pe, y0);
    sq = fma(t, pe, t);
}
__device__ __forceinline__ double fast_sqrt(double x) { double rs, sq; fast_rsqrt_sqrt(x, rs, sq); return sq; }
__device__ __forceinline__ double fast_rcp(double x) {
    const double y0 = rcp_seed(x);
    const double e = fma(-x, y0, 1.0);             // 1/x = y0 (1 + e + e^2 + O(e^3))
    return fma(y0, fma(e, e, e), y0);
}

// Per-cell gather coefficients (one copy per warp in shared memory).  With the R-face values a, b, c of a particle at
// offset x and xm = 1 - x:
//   alpha E_x   = ex1 + xm^2 dex0 + x^2 dex2        (dex0 = alpha/2 (Ex[j-1]-Ex[j]), dex2 = alpha/2 (Ex[j+1]-Ex[j]))
//   alpha E_y   = ey0 + a dey0 + b dey1 + c dey2     (dey_k = alpha (Ey[j+k]-Ey[j-1+k]))
//   alpha c B_y = ty1 + xm^2 dty0 + x^2 dty2, ...    (the Vay rotation vector tau directly)
// i.e. the spline weights (yee.rs:499-529) with the partition of unity folded into per-cell differences,
// and alpha = q dt / 2 m c (electron.rs:273) applied once per cell instead of once per particle.
struct __align__(16) CellPre { double ex1, dex0, dex2, ey0, dey[3], ez0, dez[3], ty1, dty0, dty2, tz1, dtz0, dtz2, tx; };
// coefficient k = coef_k * (raw[ia] - raw[ib]); raw: Ex[3] 0-2, Ey[4] 3-6, Ez[4] 7-10, Bx 11, By[3] 12-14, Bz[3] 15-17, zero 18
__constant__ signed char c_pre_ia[18] = {1, 0, 2, 3, 4, 5, 6, 7, 8, 9, 10, 13, 12, 14, 16, 15, 17, 11};
__constant__ signed char c_pre_ib[18] = {18, 1, 1, 18, 3, 4, 5, 18, 7, 8, 9, 18, 13, 13, 18, 16, 16, 18};
__constant__ signed char c_pre_k[18] = {0, 1, 1, 0, 0, 0, 0, 0, 0, 0, 0, 2, 3, 3, 2, 3, 3, 2};   // alpha, alpha/2, alpha c, alpha c/2

// shared memory of one warp
struct __align__(128) PushScratch {
    double ring[OPAL_PUSH_STAGES][5][32 * OPAL_PUSH_SUP];   // residents in flight: x, ux, uy, uz, w (bulk-copy destinations)
    StripRec arr[2][PUSH_ARR];                   // arrivals from the left / right neighbour cell
    CellPre pre;
    double raw[20];
    int free_list[PUSH_FREE_CAP];
    unsigned long long bar[OPAL_PUSH_STAGES + 1];   // mbarriers: one per ring stage + the arrivals
};

// Heavy cells (k_push_lean_fast<.., HEAVY = true>): a cell holding more than StepParams::heavy_min particles is left out by the
// one-warp-per-cell launch (which marks it: HEAVY_MARK) and worked on by a whole block of the launch that follows: its
// OPAL_HEAVY_WARPS warps take the ring stages of the cell's residents in turn, hand out strip slots and free slots through
// counters in shared memory, share the arrivals, and each reduce and deposit their own partial sums.  One warp alone needs
// ~11 us per 1000 particles of a cell (all latency).  Measured (tools/bench_heavy_cell.py: ONE cold pile among 2048 cells of
// 100, whole step): 2000 particles 49 -> 41 us, 4000: 71 -> 43, 8000: 115 -> 53, 16 000: 204 -> 69, 32 000: 550 -> 106; at
// 1000 the second launch costs more (2-3 us) than it saves, and a few dozen cells of 1000-3000 each (the reference's
// hole_boring and two_stream decks) run as fast without it -- different warps have them anyway.  Hence the host's default:
// above 4096 particles AND above 8x the species' average per cell (opal_host.cu: sparams, note_heavy).
#ifndef OPAL_HEAVY_WARPS
#define OPAL_HEAVY_WARPS 8
#endif
constexpr int HEAVY_LIST_CAP = OPAL_HEAVY_WARPS * 32;   // candidate cells looked at per round: one per thread
constexpr int HEAVY_FREE_CAP = 1024;                    // free-slot indices remembered per heavy cell and step
constexpr int HEAVY_MARK = -1;                          // in scount_out[2c] between the two launches: cell c is the second one's
struct HeavyShared {
    int list[HEAVY_LIST_CAP];        // the heavy cells among this round's candidates
    int n_list;
    int mloc_l, mloc_r, n_free, a_stay;   // the cell's running counters (registers in the one-warp kernel)
    int free_list[HEAVY_FREE_CAP];
};

// shift the live particles of one lean bucket down over its free slots (one warp); returns the live count
__device__ __forceinline__ int compact_cell(const BucketDev& b, long long base, int n, int lane, unsigned lt_mask) {
    __syncwarp();
    int wr = 0;
    for (int k0 = 0; k0 < n; k0 += 32) {
        const int i = k0 + lane;
        double v0 = 0.0, v1 = 0.0, v2 = 0.0, v3 = 0.0, v4 = 0.0;
        bool live = false;
        if (i < n) {
            const long long s = base + i;
            v0 = b.f[F_X][s]; live = !is_hole(v0);
            v1 = b.f[F_UX][s]; v2 = b.f[F_UY][s]; v3 = b.f[F_UZ][s]; v4 = b.f[F_W][s];
        }
        const unsigned m_live = __ballot_sync(FULLMASK, live);
        __syncwarp();        // every lane has read its slot before any lane overwrites a lower one
        const int pos = wr + __popc(m_live & lt_mask);
        if (live && pos != i) {
            const long long s = base + pos;
            b.f[F_X][s] = v0; b.f[F_UX][s] = v1; b.f[F_UY][s] = v2; b.f[F_UZ][s] = v3; b.f[F_W][s] = v4;
        }
        wr += __popc(m_live);
    }
    __syncwarp();
    return wr;
}

// closes the free slots of every cell (host: ensure_settled, before anything but the step kernels reads the buckets)
__global__ void __launch_bounds__(256) k_compact_cells(BucketDev b, int* __restrict__ hcount) {
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    for (int c = blockIdx.x * 8 + (threadIdx.x >> 5); c < b.ncell; c += gridDim.x * 8) {
        if (hcount[c] == 0) continue;
        const int cap = (int)(b.start[c + 1] - b.start[c]);
        const int n = min(b.count[c], cap);
        const int live = compact_cell(b, b.start[c], n, lane, lt_mask);
        const int parked = b.count[c] - n;                        // demand beyond the bucket's room (see k_push_lean_fast)
        const int fill_end = min(live + parked, cap);
        for (int i = live + lane; i < fill_end; i += 32) b.f[F_X][b.start[c] + i] = OPAL_HOLE_MARK;
        __syncwarp();
        if (lane == 0) { b.count[c] = live + parked; hcount[c] = fill_end - live; }
    }
}

// Appends the arrivals waiting in the strips to their buckets (no deposit: the push that wrote the strips deposited
// them).  Host: ensure_settled -- before a download, a histogram, a re-layout, the end of a lean run.
__global__ void __launch_bounds__(256) k_strip_flush(BucketDev b, StepParams sp, const StripRec* __restrict__ strips, const int* __restrict__ scount,
                                                       opal_migrant* __restrict__ overflow, StatusDev* __restrict__ st) {
    constexpr int SUB = 16;                                         // lanes per cell: two cells per warp
    const int lane = threadIdx.x & (SUB - 1);
    const int MC = sp.cell_mover_cap;
    const int nc = b.ncell;
    for (int c = (blockIdx.x * 8 + (threadIdx.x >> 5)) * 2 + ((threadIdx.x & 31) / SUB); c < nc; c += gridDim.x * 16) {
        int src_l = c - 1, src_r = c + 1;
        if (src_l < 0) src_l = sp.route_left == 1 ? nc - 1 : -1;
        if (src_r >= nc) src_r = sp.route_right == 1 ? 0 : -1;
        const int n_from_l = src_l >= 0 ? min(scount[2 * src_l + 1], MC) : 0;
        const int n_from_r = src_r >= 0 ? min(scount[2 * src_r], MC) : 0;
        const int n_in = n_from_l + n_from_r;
        if (n_in == 0) continue;
        const int cnt = b.count[c];
        const long long bstart = b.start[c];
        const int cap = (int)(b.start[c + 1] - bstart);
        for (int i = lane; i < n_in; i += SUB) {
            const StripRec r = i < n_from_l ? strips[((long long)src_l * 2 + 1) * MC + i] : strips[((long long)src_r * 2) * MC + (i - n_from_l)];
            const int slot = cnt + i;
            if (slot < cap) {
                const long long s = bstart + slot;
                b.f[F_X][s] = r.x; b.f[F_UX][s] = r.ux; b.f[F_UY][s] = r.uy; b.f[F_UZ][s] = r.uz; b.f[F_W][s] = r.w;
            } else {
                // bucket full: park; the host re-lays the buckets out (count[c] keeps the demand)
                const int o = atomicAdd(&st->n_overflow, 1);
                if (o < sp.overflow_cap) {
                    opal_migrant m;
                    m.cell = c; m.x = r.x; m.prev_x = r.prev_x; m.u[0] = r.ux; m.u[1] = r.uy; m.u[2] = r.uz; m.weight = r.w;
                    m.gamma = sqrt(fma(r.ux, r.ux, fma(r.uy, r.uy, fma(r.uz, r.uz, 1.0))));
                    m.chi = 0.0; m.tau = 0.0; m.work = 0.0; m.y = 0.0; m.z = 0.0; m.birth_time = 0.0; m.pad[0] = 0.0; m.pad[1] = 1.0;   // deposited
                    overflow[o] = m;
                } else atomicOr(&st->err, ERR_CAPACITY | ERR_CAP_PARKED);
            }
        }
        if (lane == 0) b.count[c] = cnt + n_in;
    }
}

// BX0: B_x == 0 on the whole grid (the solver never updates it, yee.rs:839-848) => tau_x terms drop out.
// KE:  also accumulate total_kinetic_energy (electron.rs:122-126: w u^2 mc^2 / (gamma + 1), mod.rs:227-240) of the
//      pushed particles into *ke_out, so the energy diagnostic that follows a step costs no second pass over HBM.
// SP:  OPAL_ELECTRON -> Vay push; OPAL_ION -> Boris push in the same reciprocal arithmetic, alpha and the deposited
//      charge from Z* and A (ion.rs:170-172).
// HEAVY: the block-per-cell variant for the cells the one-warp launch skipped (see HeavyShared above).
template <bool BX0, bool KE, int SP, bool HEAVY>
__global__ void __launch_bounds__(HEAVY ? OPAL_HEAVY_WARPS * 32 : OPAL_PUSH_WPB * 32, HEAVY ? 2 : OPAL_PUSH_MINBLK)
k_push_lean_fast(GridDev g, BucketDev b, StepParams sp, opal_migrant* __restrict__ movers, int* __restrict__ seg_count,
                 const StripRec* __restrict__ strips_in, const int* __restrict__ scount_in, StripRec* __restrict__ strips_out,
                 int* __restrict__ scount_out, int* __restrict__ hcount, opal_migrant* __restrict__ out_l, opal_migrant* __restrict__ out_r,
                 opal_migrant* __restrict__ overflow, StatusDev* __restrict__ st, double* __restrict__ ke_out) {
    if (g.halt && *(volatile const int*)g.halt) return;   // burst mode: an earlier kernel needs the host first
    constexpr int NST = OPAL_PUSH_STAGES, SUP = OPAL_PUSH_SUP, SUPN = 32 * SUP;
    extern __shared__ __align__(128) unsigned char s_dyn[];
    PushScratch& S = reinterpret_cast<PushScratch*>(s_dyn)[threadIdx.x >> 5];
    constexpr int WPC = HEAVY ? OPAL_HEAVY_WARPS : 1;                 // warps that share a cell
    constexpr int FREE_CAP = HEAVY ? HEAVY_FREE_CAP : PUSH_FREE_CAP;
    const int wib = HEAVY ? (int)(threadIdx.x >> 5) : 0;              // this warp's place among them
    HeavyShared& H = *reinterpret_cast<HeavyShared*>(s_dyn + WPC * sizeof(PushScratch));     // (HEAVY only)
    int* const free_list = HEAVY ? H.free_list : S.free_list;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = lanemask_lt();
    const int warps_total = gridDim.x * OPAL_PUSH_WPB;
    const double qsp = SP == OPAL_ELECTRON ? OPAL_QE : b.z_star * OPAL_E;
    const double alpha = SP == OPAL_ELECTRON ? OPAL_QE * g.dt / (2.0 * OPAL_ME * OPAL_C) : qsp * g.dt / (2.0 * (b.mass_number * OPAL_MP) * OPAL_C);
    const double ac = alpha * OPAL_C;
    const double kdt = OPAL_C * g.dt / g.dx;
    bool cfl = false, cap_err = false;
    double ske = 0.0;
    double* const fx = b.f[F_X]; double* const fux = b.f[F_UX]; double* const fuy = b.f[F_UY]; double* const fuz = b.f[F_UZ];
    double* const fw = b.f[F_W];
    const int MC = sp.cell_mover_cap;
    const int nc = b.ncell;

    if (elect_one()) {
#pragma unroll
        for (int k = 0; k <= NST; k++) mbar_init(&S.bar[k], 1);
        mbar_fence_init();
    }
    __syncwarp();
    unsigned phase = 0;                          // bit k: the parity the next wait on bar[k] expects

    // HEAVY: the next cell of this block (every thread calling together); nc when there is none
    int h_round = 0, h_pos = 0, h_n = 0;
    auto next_heavy = [&]() -> int {
        if constexpr (!HEAVY) return nc;
        else {
            // rounds of blockDim.x candidates, cell = blockIdx.x + k gridDim.x (neighbouring heavy cells go to different blocks):
            // every thread looks at one, the heavy ones are listed in shared memory and worked through by the whole block
            for (;;) {
                if (h_pos < h_n) return H.list[h_pos++];
                const long long k0 = (long long)h_round * blockDim.x;
                if (blockIdx.x + k0 * gridDim.x >= nc) return nc;
                __syncthreads();                                   // everyone is done with the old list
                if (threadIdx.x == 0) H.n_list = 0;
                __syncthreads();
                // (the mark the one-warp launch left, not the population: that launch has changed the counts of the cells it did)
                const long long cand = blockIdx.x + (k0 + threadIdx.x) * gridDim.x;
                if (cand < nc && scount_out[2 * cand] == HEAVY_MARK) H.list[atomicAdd(&H.n_list, 1)] = (int)cand;
                __syncthreads();
                h_n = H.n_list; h_pos = 0; h_round++;
            }
        }
    };
    // (the one-warp launch covers the cells [c_begin, c_end) the host vouches for -- all of them outside a burst; next_heavy() ends
    //  with nc >= c_end)
    const int c_end = HEAVY ? nc : b.c_end;
    for (int c = HEAVY ? next_heavy() : b.c_begin + (int)(blockIdx.x * OPAL_PUSH_WPB + (threadIdx.x >> 5)); c < c_end; c = HEAVY ? next_heavy() : c + warps_total) {
        const long long base = b.start[c];
        const int cap = (int)(b.start[c + 1] - base);
        const int n = min(b.count[c], cap);
        if constexpr (HEAVY) {
            __syncthreads();                                       // every warp is done with the last cell's counters
            if (threadIdx.x == 0) { H.mloc_l = 0; H.mloc_r = 0; H.n_free = 0; H.a_stay = 0; }
            __syncthreads();
        } else if (n > sp.heavy_min) {                             // left to the block-per-cell launch that follows
            if (lane == 0) scount_out[2 * c] = HEAVY_MARK;
            continue;
        }
        // arrivals: last step's right-movers of the left neighbour cell and left-movers of the right neighbour cell
        int src_l = c - 1, src_r = c + 1;
        if (src_l < 0) src_l = sp.route_left == 1 ? nc - 1 : -1;         // periodic self-neighbour: the opposite edge
        if (src_r >= nc) src_r = sp.route_right == 1 ? 0 : -1;
        const int n_from_l = src_l >= 0 ? min(scount_in[2 * src_l + 1], MC) : 0;
        const int n_from_r = src_r >= 0 ? min(scount_in[2 * src_r], MC) : 0;
        const int n_arr = n_from_l + n_from_r;
        const int n_chunks = (n + 31) >> 5;
        const int n_sup = (n_chunks + SUP - 1) / SUP;                     // ring stages' worth of residents
        if (n_chunks == 0 && n_arr == 0) {
            if (lane < 2) scount_out[2 * c + lane] = 0;
            continue;
        }
        const int j = c + g.left;
        const StripRec* const arr_l = strips_in + ((long long)(src_l < 0 ? 0 : src_l) * 2 + 1) * MC;
        const StripRec* const arr_r = strips_in + ((long long)(src_r < 0 ? 0 : src_r) * 2) * MC;
        // staged through shared memory (HEAVY: the arrivals are shared out among the warps and read from global memory)
        const int st_l = HEAVY ? 0 : min(n_from_l, PUSH_ARR), st_r = HEAVY ? 0 : min(n_from_r, PUSH_ARR);
        __syncwarp();
        if (elect_one()) {
#pragma unroll
            for (int s_ = 0; s_ < NST - 1; s_++) {
                const int q = wib + s_ * WPC;                      // this warp's ring stages: wib, wib + WPC, ...
                if (q < n_sup) {
                    const unsigned bytes = 256u * (unsigned)min(SUP, n_chunks - q * SUP);
                    mbar_expect_tx(&S.bar[s_], 5 * bytes);
                    const long long o = base + (long long)q * SUPN;
                    bulk_g2s(S.ring[s_][0], fx + o, bytes, &S.bar[s_]); bulk_g2s(S.ring[s_][1], fux + o, bytes, &S.bar[s_]);
                    bulk_g2s(S.ring[s_][2], fuy + o, bytes, &S.bar[s_]); bulk_g2s(S.ring[s_][3], fuz + o, bytes, &S.bar[s_]);
                    bulk_g2s(S.ring[s_][4], fw + o, bytes, &S.bar[s_]);
                }
            }
            if (st_l + st_r) {
                const unsigned bl = (unsigned)st_l * (unsigned)sizeof(StripRec), br = (unsigned)st_r * (unsigned)sizeof(StripRec);
                mbar_expect_tx(&S.bar[NST], bl + br);
                if (bl) bulk_g2s(S.arr[0], arr_l, bl, &S.bar[NST]);
                if (br) bulk_g2s(S.arr[1], arr_r, br, &S.bar[NST]);
            }
        }
        // the 18 field values every lane of this warp gathers from -> 18 per-cell gather coefficients (CellPre)
        if (lane < 3) { S.raw[lane] = __ldg(g.ex + j - 1 + lane); S.raw[12 + lane] = __ldg(g.by + j - 1 + lane); S.raw[15 + lane] = __ldg(g.bz + j - 1 + lane); }
        if (lane < 4) { S.raw[3 + lane] = __ldg(g.ey + j - 1 + lane); S.raw[7 + lane] = __ldg(g.ez + j - 1 + lane); }
        if (lane == 0) { S.raw[11] = BX0 ? 0.0 : __ldg(g.bx + j); S.raw[18] = 0.0; }
        __syncwarp();
        if (lane < 18) {
            const int k = c_pre_k[lane];
            // the Vay push needs E only as u' = u + 2 alpha E + ...: its E coefficients carry the factor 2 (the Boris push adds alpha E twice)
            const double coef = (k & 2 ? ac : (SP == OPAL_ELECTRON ? 2.0 * alpha : alpha)) * (k & 1 ? 0.5 : 1.0);
            reinterpret_cast<double*>(&S.pre)[lane] = coef * (S.raw[c_pre_ia[lane]] - S.raw[c_pre_ib[lane]]);
        }
        __syncwarp();
        double sjx0 = 0.0, sjx1 = 0.0, sjx2 = 0.0, sjy0 = 0.0, sjy1 = 0.0, sjy2 = 0.0, sjz0 = 0.0, sjz1 = 0.0, sjz2 = 0.0;
        double sr0 = 0.0, sr1 = 0.0, sr2 = 0.0, sr3 = 0.0;
        int mloc_l = 0, mloc_r = 0;   // left / right movers of this cell so far (warp-uniform): slots in its two strips
        int n_free = 0;               // free slots seen so far (warp-uniform): departures + slots that were already free
        // edge cells: what happens to a particle that leaves the slab (mod.rs:290-373).  0 delete, 1 wrap (an ordinary
        // mover: the opposite edge cell reads this cell's strip), 2 hand to the neighbour slab through the outbox
        const int route_l = c == 0 ? sp.route_left : 1, route_r = c == nc - 1 ? sp.route_right : 1;

        // ---- gather + push of one particle per lane; every lane runs it (a lane without a live particle holds finite
        //      junk and weight 0).  x, u in; u, xn, fl (cell change), igam = 1/gamma, gam2 = gamma^2, ro = R(x) out ----
        auto push = [&](double x, V3& u, double W, bool live, double& xn, double& fl, double& igam, double& gam2, RFaces& ro) {
            // gather (yee.rs:499-529): alpha E and tau = alpha c B from the per-cell coefficients (16-byte shared loads)
            const double2* const P2 = reinterpret_cast<const double2*>(&S.pre);
            const double2 p0 = P2[0], p1 = P2[1], p2 = P2[2], p3 = P2[3], p4 = P2[4], p5 = P2[5], p6 = P2[6], p7 = P2[7], p8 = P2[8];
            // p0 = (ex1, dex0) p1 = (dex2, ey0) p2 = (dey0, dey1) p3 = (dey2, ez0) p4 = (dez0, dez1) p5 = (dez2, ty1) p6 = (dty0, dty2)
            // p7 = (tz1, dtz0) p8 = (dtz2, tx)
            ro = r_faces(x);
            const double xm = 1.0 - x, xm2 = xm * xm, x2 = x * x;
            V3 aE, tv;
            aE.x = fma(x2, p1.x, fma(xm2, p0.y, p0.x));
            aE.y = fma(ro.c, p3.x, fma(ro.b, p2.y, fma(ro.a, p2.x, p1.y)));
            aE.z = fma(ro.c, p5.x, fma(ro.b, p4.y, fma(ro.a, p4.x, p3.y)));
            tv.x = BX0 ? 0.0 : p8.y;
            tv.y = fma(x2, p6.y, fma(xm2, p6.x, p5.y));
            tv.z = fma(x2, p8.x, fma(xm2, p7.y, p7.x));
            double gam;
            if (SP == OPAL_ELECTRON) {
                // ---- Vay push (electron.rs:268-330) ----
                // u_half = u + alpha (E + v x B) with v = c u / gamma  ==  u + alpha E + (u x tau) / gamma;  u' = u_half + alpha E
                // (aE holds 2 alpha E here)
                const double ig = fast_rsqrt(fma(u.x, u.x, fma(u.y, u.y, fma(u.z, u.z, 1.0))));
                const V3 uxt = BX0 ? V3{u.y * tv.z - u.z * tv.y, -(u.x * tv.z), u.x * tv.y} : cross(u, tv);
                const V3 u_prime = {fma(ig, uxt.x, u.x + aE.x), fma(ig, uxt.y, u.y + aE.y), fma(ig, uxt.z, u.z + aE.z)};
                const double gp2 = fma(u_prime.x, u_prime.x, fma(u_prime.y, u_prime.y, fma(u_prime.z, u_prime.z, 1.0)));
                const double u_star = BX0 ? fma(u_prime.y, tv.y, u_prime.z * tv.z) : dot(u_prime, tv);
                const double tau2 = BX0 ? fma(tv.y, tv.y, tv.z * tv.z) : dot(tv, tv);
                const double hs = 0.5 * (gp2 - tau2);
                gam2 = hs + fast_sqrt(fma(hs, hs, fma(u_star, u_star, tau2)));
                fast_rsqrt_sqrt(gam2, igam, gam);
                // u = s (u' + (u'.t) t + u' x t), t = tau/gamma, s = 1/(1+t^2)   ==   (gamma^2 u' + (u'.tau) tau + gamma u' x tau) / (gamma^2 + tau^2)
                const double inv = fast_rcp(gam2 + tau2);
                const V3 pxt = BX0 ? V3{u_prime.y * tv.z - u_prime.z * tv.y, -(u_prime.x * tv.z), u_prime.x * tv.y} : cross(u_prime, tv);
                V3 w = {fma(gam, pxt.x, gam2 * u_prime.x), fma(gam, pxt.y, gam2 * u_prime.y), fma(gam, pxt.z, gam2 * u_prime.z)};
                if (!BX0) w.x = fma(u_star, tv.x, w.x);
                w.y = fma(u_star, tv.y, w.y); w.z = fma(u_star, tv.z, w.z);
                u = inv * w;
            } else {
                // ---- Boris push (ion.rs:168-214): u- = u + alpha E; u' = u- + (u- x tau)/gamma-;
                //      u+ = u- + (u' x tau) 2 gamma- / (gamma-^2 + tau^2); u = u+ + alpha E   (tau = alpha c B) ----
                const V3 um = u + aE;
                const double ig = fast_rsqrt(fma(um.x, um.x, fma(um.y, um.y, fma(um.z, um.z, 1.0))));
                const V3 mxt = BX0 ? V3{um.y * tv.z - um.z * tv.y, -(um.x * tv.z), um.x * tv.y} : cross(um, tv);
                const V3 u_prime = {fma(ig, mxt.x, um.x), fma(ig, mxt.y, um.y), fma(ig, mxt.z, um.z)};
                const double tau2 = BX0 ? fma(tv.y, tv.y, tv.z * tv.z) : dot(tv, tv);
                const double s2 = (2.0 * ig) * fast_rcp(fma(tau2 * ig, ig, 1.0));
                const V3 pxt = BX0 ? V3{u_prime.y * tv.z - u_prime.z * tv.y, -(u_prime.x * tv.z), u_prime.x * tv.y} : cross(u_prime, tv);
                u = {fma(s2, pxt.x, um.x) + aE.x, fma(s2, pxt.y, um.y) + aE.y, fma(s2, pxt.z, um.z) + aE.z};
                gam2 = fma(u.x, u.x, fma(u.y, u.y, fma(u.z, u.z, 1.0)));
                fast_rsqrt_sqrt(gam2, igam, gam);
            }
            const double dxi = (kdt * igam) * u.x;
            if (live && !(dxi < 1.0)) cfl = true;
            xn = x + dxi;
            fl = floor(xn);
            xn -= fl;
            if (KE) ske = fma(W * fma(u.x, u.x, fma(u.y, u.y, u.z * u.z)), fast_rcp(gam + 1.0), ske);   // W == 0 on a lane without a particle
        };
        // deposit of a stayer (yee.rs:584-609)
        auto deposit_stayer = [&](double W, const V3& u, double igam, const RFaces& ro, double xn) {
            const RFaces rn = r_faces(xn);
            sjx0 = fma(W, rn.a - ro.a, sjx0); sjx1 = fma(W, rn.b - ro.b, sjx1); sjx2 = fma(W, rn.c - ro.c, sjx2);
            const double w0 = 1.0 - rn.a, w1 = rn.a - rn.b, w2 = rn.b - rn.c, w3 = rn.c;
            sr0 = fma(W, w3, sr0); sr1 = fma(W, w0, sr1); sr2 = fma(W, w1, sr2); sr3 = fma(W, w2, sr3);
            const double tw = W * igam, ty = tw * u.y, tz = tw * u.z;
            sjy0 = fma(ty, w0, sjy0); sjy1 = fma(ty, w1, sjy1); sjy2 = fma(ty, w2, sjy2);
            sjz0 = fma(tz, w0, sjz0); sjz1 = fma(tz, w1, sjz1); sjz2 = fma(tz, w2, sjz2);
        };
        // a mover: into this step's strip of the cell (the neighbour cell's next push picks it up), or out of the slab
        // Movers that find their cell's strip full go to the shared segmented buffer (k_scatter, right after this kernel, puts
        // them into their buckets and deposits them).  ONE reservation there for all of a chunk's -- same cell, same chunk: same
        // segment; a relativistically hot cell sends half a chunk this way, and that many same-address atomics with a return
        // value cost microseconds.  Called in warp-uniform code; returns the lane's slot in the segment.  (HEAVY only: in the
        // one-warp kernel the test for "will this chunk spill" sits in the path of every chunk with a mover and costs the headline
        // case 2.6 %; there each spilling lane reserves its own slot inside hand_over, as it always did.)
        auto reserve_spill = [&](bool spills, int it) -> int {
            const unsigned m_ov = __ballot_sync(FULLMASK, spills);
            if (!m_ov) return 0;
            const int seg = (int)(((unsigned)c * 0x9E3779B1u + (unsigned)it * 0x85EBCA77u) >> 22);   // top 10 bits: OPAL_MOVER_SEGMENTS == 1024
            int rg = 0;
            if (lane == 0) {
                rg = atomicAdd(seg_count + seg, __popc(m_ov));
                atomicAdd(&st->strip_spill[SP == OPAL_ELECTRON ? 0 : 1], __popc(m_ov));           // the host widens the strips when this is the rule
            }
            return __shfl_sync(FULLMASK, rg, 0) + __popc(m_ov & lt_mask);
        };
        auto hand_over = [&](bool left, int r, int rg /* HEAVY: from reserve_spill */, double x_old, double xn, double fl, const V3& u, double W, double igam, double gam2, int it) {
            const int route = left ? route_l : route_r;
            if (route == 1) {
                if (r < MC) {
                    double2* const rec = reinterpret_cast<double2*>(strips_out + ((long long)c * 2 + (left ? 0 : 1)) * MC + r);   // 3 x 16-byte stores
                    rec[0] = make_double2(xn, x_old); rec[1] = make_double2(u.x, u.y); rec[2] = make_double2(u.z, W);
                } else {
                    // strip full (a cell far hotter or denser than the strips were sized for): slot rg of its segment (reserve_spill)
                    const int seg = (int)(((unsigned)c * 0x9E3779B1u + (unsigned)it * 0x85EBCA77u) >> 22);   // (as in reserve_spill)
                    if constexpr (!HEAVY) {
                        rg = atomicAdd(seg_count + seg, 1);
                        atomicAdd(&st->strip_spill[SP == OPAL_ELECTRON ? 0 : 1], 1);
                    }
                    if (rg < sp.mover_cap) {
                        opal_migrant m;
                        long long dst = (long long)c + (left ? -1 : 1);
                        if (dst < 0) dst += nc; else if (dst >= nc) dst -= nc;      // route 1 at a slab edge: periodic wrap
                        m.cell = dst; m.x = xn; m.prev_x = x_old - fl; m.u[0] = u.x; m.u[1] = u.y; m.u[2] = u.z;
                        m.weight = W; m.gamma = gam2 * igam; m.chi = 0.0; m.tau = 0.0; m.work = 0.0; m.y = 0.0; m.z = 0.0;
                        m.birth_time = 0.0; m.pad[0] = 0.0; m.pad[1] = 0.0;
                        movers[(long long)seg * sp.mover_cap + rg] = m;
                    } else cap_err = true;
                }
            } else if (route == 2) {
                // to the neighbour slab: it deposits the particle in its own frame (ghost cells included) on arrival
                const int o = atomicAdd(&st->n_out[left ? 0 : 1], 1);
                if (o < sp.out_cap) {
                    opal_migrant m;
                    m.cell = left ? -1 : 0;                    // already re-indexed for the receiver (mod.rs:291-293)
                    m.x = xn; m.prev_x = x_old - fl; m.u[0] = u.x; m.u[1] = u.y; m.u[2] = u.z; m.weight = W; m.gamma = gam2 * igam;
                    m.chi = 0.0; m.tau = 0.0; m.work = 0.0; m.y = 0.0; m.z = 0.0; m.birth_time = 0.0; m.pad[0] = 0.0; m.pad[1] = 0.0;
                    (left ? out_l : out_r)[o] = m;
                } else atomicOr(&st->err, ERR_CAPACITY | ERR_CAP_OUTBOX);
            } else if (KE) {
                // an open edge: the particle is deleted (mod.rs:372-373) and is not part of the population's energy any more
                const double gam = gam2 * igam;
                ske -= (W * fma(u.x, u.x, fma(u.y, u.y, u.z * u.z))) * fast_rcp(gam + 1.0);
            }
        };

        // ================= the residents: in place =================
        auto resident_chunk = [&](int k0, const double (*R)[32 * SUP], int e) {
            const int i = k0 + lane;
            const double x = R[0][e];
            V3 u = {R[1][e], R[2][e], R[3][e]};
            const bool live = i < n && !is_hole(x);
            const double W = live ? R[4][e] : 0.0;
            double xn, fl, igam, gam2;
            RFaces ro;
            push(x, u, W, live, xn, fl, igam, gam2, ro);
            const bool stay = live && fl == 0.0;
            const unsigned m_l = __ballot_sync(FULLMASK, live && fl < 0.0);
            const unsigned m_r = __ballot_sync(FULLMASK, live && fl > 0.0);
            // free slots of this chunk: just vacated, or empty since an earlier step
            const unsigned m_free = __ballot_sync(FULLMASK, i < n && !stay);
            const long long s = base + i;
            if (stay) {
                fx[s] = xn; fux[s] = u.x; fuy[s] = u.y; fuz[s] = u.z;
                deposit_stayer(W, u, igam, ro, xn);
            }
            if (m_free) {                                      // warp-uniform
                int bf = n_free, bl = mloc_l, br = mloc_r;     // where this chunk's free slots / movers go
                if constexpr (HEAVY) {
                    if (lane == 0) {
                        bf = atomicAdd(&H.n_free, __popc(m_free));
                        bl = m_l ? atomicAdd(&H.mloc_l, __popc(m_l)) : 0;
                        br = m_r ? atomicAdd(&H.mloc_r, __popc(m_r)) : 0;
                    }
                    bf = __shfl_sync(FULLMASK, bf, 0); bl = __shfl_sync(FULLMASK, bl, 0); br = __shfl_sync(FULLMASK, br, 0);
                }
                const bool left = fl < 0.0;
                const int r = left ? bl + __popc(m_l & lt_mask) : br + __popc(m_r & lt_mask);      // (a mover's) slot in its strip
                int rg = 0;
                if constexpr (HEAVY) {
                    if (bl + __popc(m_l) > MC || br + __popc(m_r) > MC)                            // warp-uniform
                        rg = reserve_spill(live && !stay && (left ? route_l : route_r) == 1 && r >= MC, k0 >> 5);
                }
                if ((m_free >> lane) & 1u) {
                    const int k = bf + __popc(m_free & lt_mask);
                    if (k < FREE_CAP) free_list[k] = i;
                    if (live) {                                // a mover
                        fx[s] = OPAL_HOLE_MARK;                // the vacated slot
                        hand_over(left, r, rg, x, xn, fl, u, W, igam, gam2, k0 >> 5);
                    }
                }
                if constexpr (!HEAVY) {
                    n_free += __popc(m_free);
                    mloc_l += __popc(m_l); mloc_r += __popc(m_r);
                }
            }
        };
        int stage = 0;
        for (int ks = wib; ks < n_sup; ks += WPC) {               // one ring stage = SUP chunks of 32
            mbar_wait(&S.bar[stage], (phase >> stage) & 1u);
            phase ^= 1u << stage;
            __syncwarp();                                          // every lane is done with the stage the next copy refills
            if (elect_one()) {
                const int nxt = ks + (NST - 1) * WPC;
                if (nxt < n_sup) {
                    const int ps = stage == 0 ? NST - 1 : stage - 1;      // consumed before this one
                    const unsigned bytes = 256u * (unsigned)min(SUP, n_chunks - nxt * SUP);
                    mbar_expect_tx(&S.bar[ps], 5 * bytes);
                    const long long o = base + (long long)nxt * SUPN;
                    bulk_g2s(S.ring[ps][0], fx + o, bytes, &S.bar[ps]); bulk_g2s(S.ring[ps][1], fux + o, bytes, &S.bar[ps]);
                    bulk_g2s(S.ring[ps][2], fuy + o, bytes, &S.bar[ps]); bulk_g2s(S.ring[ps][3], fuz + o, bytes, &S.bar[ps]);
                    bulk_g2s(S.ring[ps][4], fw + o, bytes, &S.bar[ps]);
                }
            }
            const int k_base = ks * SUPN;
#pragma unroll
            for (int sub = 0; sub < SUP; sub++)
                if (sub == 0 || k_base + sub * 32 < n) resident_chunk(k_base + sub * 32, S.ring[stage], sub * 32 + lane);
            stage = stage == NST - 1 ? 0 : stage + 1;
        }

        // ================= the arrivals: into the free slots, the rest behind the bucket =================
        int a_stay = 0;               // arrivals that stay, so far
        if (n_arr) {                                           // (block-uniform)
            if (st_l + st_r) { mbar_wait(&S.bar[NST], (phase >> NST) & 1u); phase ^= 1u << NST; }
            if constexpr (HEAVY) { __syncthreads(); n_free = H.n_free; }      // every warp has listed its free slots
            const int n_listed = min(n_free, FREE_CAP);
            __syncwarp();                                      // the free list is complete
            for (int k0 = wib * 32; k0 < n_arr; k0 += WPC * 32) {
                const int i = k0 + lane;
                const bool live = i < n_arr;
                double x = 0.5, W = 0.0;
                V3 u = {0.0, 0.0, 0.0};
                if (live) {
                    StripRec r;
                    if (i < n_from_l) r = i < st_l ? S.arr[0][i] : arr_l[i];
                    else { const int k = i - n_from_l; r = k < st_r ? S.arr[1][k] : arr_r[k]; }
                    x = r.x; u = {r.ux, r.uy, r.uz}; W = r.w;
                }
                double xn, fl, igam, gam2;
                RFaces ro;
                push(x, u, W, live, xn, fl, igam, gam2, ro);
                const bool stay = live && fl == 0.0;
                const unsigned m_stay = __ballot_sync(FULLMASK, stay);
                const unsigned m_l = __ballot_sync(FULLMASK, live && fl < 0.0);
                const unsigned m_r = __ballot_sync(FULLMASK, live && fl > 0.0);
                int ba = a_stay, bl = mloc_l, br = mloc_r;
                if constexpr (HEAVY) {
                    if (lane == 0) {
                        ba = m_stay ? atomicAdd(&H.a_stay, __popc(m_stay)) : 0;
                        bl = m_l ? atomicAdd(&H.mloc_l, __popc(m_l)) : 0;
                        br = m_r ? atomicAdd(&H.mloc_r, __popc(m_r)) : 0;
                    }
                    ba = __shfl_sync(FULLMASK, ba, 0); bl = __shfl_sync(FULLMASK, bl, 0); br = __shfl_sync(FULLMASK, br, 0);
                }
                const bool left = fl < 0.0;
                const int r = left ? bl + __popc(m_l & lt_mask) : br + __popc(m_r & lt_mask);
                int rg = 0;
                if constexpr (HEAVY) {
                    if (bl + __popc(m_l) > MC || br + __popc(m_r) > MC)
                        rg = reserve_spill(live && !stay && (left ? route_l : route_r) == 1 && r >= MC, n_chunks + (k0 >> 5));
                }
                if (stay) {
                    const int k = ba + __popc(m_stay & lt_mask);
                    const int slot = k < n_listed ? free_list[k] : n + (k - n_listed);     // a free slot, else appended
                    if (slot < cap) {
                        const long long s = base + slot;
                        fx[s] = xn; fux[s] = u.x; fuy[s] = u.y; fuz[s] = u.z; fw[s] = W;
                    } else {
                        // bucket full: park; the host re-lays the buckets out (count[c] keeps the demand)
                        const int o = atomicAdd(&st->n_overflow, 1);
                        if (o < sp.overflow_cap) {
                            opal_migrant m;
                            m.cell = c; m.x = xn; m.prev_x = x; m.u[0] = u.x; m.u[1] = u.y; m.u[2] = u.z; m.weight = W; m.gamma = gam2 * igam;
                            m.chi = 0.0; m.tau = 0.0; m.work = 0.0; m.y = 0.0; m.z = 0.0; m.birth_time = 0.0; m.pad[0] = 0.0; m.pad[1] = 1.0;   // deposited below
                            overflow[o] = m;
                        } else atomicOr(&st->err, ERR_CAPACITY | ERR_CAP_PARKED);
                    }
                    deposit_stayer(W, u, igam, ro, xn);
                } else if (live) hand_over(left, r, rg, x, xn, fl, u, W, igam, gam2, n_chunks + (k0 >> 5));
                if constexpr (!HEAVY) {
                    a_stay += __popc(m_stay);
                    mloc_l += __popc(m_l); mloc_r += __popc(m_r);
                }
            }
        }

        // ================= the cell's bookkeeping =================
        if constexpr (HEAVY) {
            __syncthreads();                                       // the cell's counters are final, its strip records visible to the block
            n_free = H.n_free; a_stay = H.a_stay; mloc_l = H.mloc_l; mloc_r = H.mloc_r;
        }
        const int n_listed = min(n_free, FREE_CAP);
        const int used = min(a_stay, n_listed);                   // free slots the arrivals took
        const int n_new = n + (a_stay - used);                    // slots in use now (parked demand included)
        const int holes = n_free - used;
        // (a heavy cell may carry an eighth of its slots free before its first warp closes them, one chunk at a time)
        if (holes > (HEAVY ? max(OPAL_HOLE_MAX, n_new >> 3) : OPAL_HOLE_MAX)) {
            if (wib == 0) {
                const int live_n = compact_cell(b, base, min(n_new, cap), lane, lt_mask);
                // arrivals parked because the bucket was full stay counted (count[c] is the demand the host re-lays the buckets out
                // for); the slots that count now covers behind the live particles hold what the shift left there: mark them free
                const int parked = n_new - min(n_new, cap);
                const int fill_end = min(live_n + parked, cap);
                for (int i = live_n + lane; i < fill_end; i += 32) fx[base + i] = OPAL_HOLE_MARK;
                if (lane == 0) { b.count[c] = live_n + parked; hcount[c] = fill_end - live_n; }
            }
        } else if (lane == 0 && wib == 0) {
            b.count[c] = n_new; hcount[c] = holes;
        }
        if (lane == 0 && wib == 0 && n_new > sp.heavy_hint) st->heavy_seen = 1;      // the host switches the block-per-cell launch on / off
        const int dep_l = route_l == 1 ? min(mloc_l, MC) : 0, dep_r = route_r == 1 ? min(mloc_r, MC) : 0;   // movers in this step's strips
        if (lane == 0 && wib == 0) {
            // leavers of the slab (mod.rs:284), whatever happens to them
            if (c == 0 && mloc_l) atomicAdd(&st->gone[0], mloc_l);
            if (c == nc - 1 && mloc_r) atomicAdd(&st->gone[1], mloc_r);
            scount_out[2 * c] = dep_l; scount_out[2 * c + 1] = dep_r;
        }

        // ================= the movers' deposit: from the strips just written, every lane busy =================
        // In the frame of the cell they left (faces j-2 .. j+2, nodes j-3 .. j+2): R of the new position one cell to the right
        // is (1, 1, a, b, c), one cell to the left (a, b, c, 0, 0); of the old position (1, a_o, b_o, c_o, 0);
        // flux = R_new - R_old (== Cell::flux, yee.rs:185).  e*: the targets beyond the stayers' stencil.
        double ejx_m = 0.0, ejy_m = 0.0, ejz_m = 0.0, er_m = 0.0, ejx_p = 0.0, ejy_p = 0.0, ejz_p = 0.0, er_p = 0.0;
        if (dep_l + dep_r) {
            __syncwarp();                                          // the records written above are visible to every lane
            const StripRec* const mine_l = strips_out + ((long long)c * 2) * MC;
            for (int k = wib * 32 + lane; k < dep_l; k += WPC * 32) {
                const StripRec r = mine_l[k];
                const RFaces ro = r_faces(r.prev_x), rn = r_faces(r.x);
                const double W = r.w, tw = W * fast_rsqrt(fma(r.ux, r.ux, fma(r.uy, r.uy, fma(r.uz, r.uz, 1.0)))), ty = tw * r.uy, tz = tw * r.uz;
                const double w0 = 1.0 - rn.a, w1 = rn.a - rn.b, w2 = rn.b - rn.c, w3 = rn.c;
                ejx_m = fma(W, rn.a - 1.0, ejx_m); sjx0 = fma(W, rn.b - ro.a, sjx0); sjx1 = fma(W, rn.c - ro.b, sjx1); sjx2 = fma(-W, ro.c, sjx2);
                ejy_m = fma(ty, w0, ejy_m); sjy0 = fma(ty, w1, sjy0); sjy1 = fma(ty, w2, sjy1);
                ejz_m = fma(tz, w0, ejz_m); sjz0 = fma(tz, w1, sjz0); sjz1 = fma(tz, w2, sjz1);
                er_m = fma(W, w3, er_m); sr0 = fma(W, w0, sr0); sr1 = fma(W, w1, sr1); sr2 = fma(W, w2, sr2);
            }
            const StripRec* const mine_r = mine_l + MC;
            for (int k = wib * 32 + lane; k < dep_r; k += WPC * 32) {
                const StripRec r = mine_r[k];
                const RFaces ro = r_faces(r.prev_x), rn = r_faces(r.x);
                const double W = r.w, tw = W * fast_rsqrt(fma(r.ux, r.ux, fma(r.uy, r.uy, fma(r.uz, r.uz, 1.0)))), ty = tw * r.uy, tz = tw * r.uz;
                const double w0 = 1.0 - rn.a, w1 = rn.a - rn.b, w2 = rn.b - rn.c, w3 = rn.c;
                sjx0 = fma(W, 1.0 - ro.a, sjx0); sjx1 = fma(W, rn.a - ro.b, sjx1); sjx2 = fma(W, rn.b - ro.c, sjx2); ejx_p = fma(W, rn.c, ejx_p);
                sjy1 = fma(ty, w0, sjy1); sjy2 = fma(ty, w1, sjy2); ejy_p = fma(ty, w2, ejy_p);
                sjz1 = fma(tz, w0, sjz1); sjz2 = fma(tz, w1, sjz2); ejz_p = fma(tz, w2, ejz_p);
                sr1 = fma(W, w3, sr1); sr2 = fma(W, w0, sr2); sr3 = fma(W, w1, sr3); er_p = fma(W, w2, er_p);
            }
        }

        // ONE transposing reduction (31 shuffles of doubles: lane k ends up with the warp total of sum k) and <= 21 atomics
        // per CELL; q/dt, q c/dx, q/dx applied here.
        // targets: 0-2 jx j-1..j+1 | 3-5 jy | 6-8 jz | 9-12 rho j-2..j+1 | 13-15 jx, jy, jz j-2 | 16 rho j-3 | 17-19 jx, jy, jz j+2 | 20 rho j+2
        double v[32] = {sjx0, sjx1, sjx2, sjy0, sjy1, sjy2, sjz0, sjz1, sjz2, sr0, sr1, sr2, sr3, ejx_m, ejy_m, ejz_m, er_m, ejx_p, ejy_p, ejz_p, er_p,
                        0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int width = 16; width >= 1; width >>= 1) {
            const bool upper = (lane & width) != 0;
#pragma unroll
            for (int k = 0; k < width; k++) {
                const double send = upper ? v[k] : v[k + width];
                const double keep = upper ? v[k + width] : v[k];
                v[k] = keep + __shfl_xor_sync(FULLMASK, send, width);
            }
        }
        if (lane < 21) {
            const double kx = qsp / g.dt, kr = qsp / g.dx, ky = qsp * OPAL_C / g.dx;
            const int k = lane;
            const double scale = (k < 3 || k == 13 || k == 17) ? kx : (((k >= 9 && k < 13) || k == 16 || k == 20) ? kr : ky);
            double* const dst = k < 3 ? g.jx + j - 1 + k : k < 6 ? g.jy + j - 1 + (k - 3) : k < 9 ? g.jz + j - 1 + (k - 6) : k < 13 ? g.rho + j - 2 + (k - 9)
                              : k == 13 ? g.jx + j - 2 : k == 14 ? g.jy + j - 2 : k == 15 ? g.jz + j - 2 : k == 16 ? g.rho + j - 3
                              : k == 17 ? g.jx + j + 2 : k == 18 ? g.jy + j + 2 : k == 19 ? g.jz + j + 2 : g.rho + j + 2;
            if (v[0] != 0.0) atomicAdd(dst, scale * v[0]);
        }
    }
    if (cfl) atomicOr(&st->err, ERR_CFL);
    if (cap_err) atomicOr(&st->err, ERR_CAPACITY | ERR_CAP_MOVERS);
    if (KE) {
        ske = warp_sum(ske) * (1.0e6 * OPAL_ME_MEV * OPAL_E);
        if (lane == 0 && ske != 0.0) atomicAdd(ke_out, ske);
    }
}
