// opal_push_full.cuh -- electrons with the reference's WHOLE per-particle state (QED on, or r / work / chi outputs asked for):
// the arithmetic of k_push_lean_fast (per-cell gather coefficients, R-face spline, reciprocal multiplies) extended by what
// Electron::vay_push maintains besides x and u (electron.rs:268-330): work (:278), chi (:282), tau -= rate dt (:287, executed
// whether or not QED is on: quirk Q4), gamma, prev_x, and y, z advanced with the PRE-push velocity (:315-316, quirk Q3).
// Included by opal_kernels.cu inside namespace opal::fast after opal_push_lean.cuh (fast build only); the strict build keeps
// the literal statement-by-statement code (k_push + electron_push), which this kernel matches to re-association (<= 1e-14).
//
// One warp per cell.  Ten input streams (x, u, w, gamma, tau, work, y, z) arrive through a shared-memory ring of 1-D bulk
// copies; stayers are shifted down over the movers' slots in place (eleven output streams: all but w change every step);
// movers go to the segmented buffer as full 128-B records for k_scatter.  The literal kernel spent its time in 24 IEEE
// divisions + 3 sqrt per particle at one resident block per SM (140 registers): profiles/r02a_qed_step.summary.md.
#pragma once

#ifndef OPAL_FULL_STAGES
#define OPAL_FULL_STAGES 3
#endif
#ifndef OPAL_FULL_WPB
#define OPAL_FULL_WPB 4
#endif
#ifndef OPAL_FULL_MINBLK
#define OPAL_FULL_MINBLK 4
#endif
constexpr int FULL_NIN = 10;          // streams read: x ux uy uz w gamma tau work y z
__constant__ signed char c_full_in[FULL_NIN] = {F_X, F_UX, F_UY, F_UZ, F_W, F_GAMMA, F_TAU, F_WORK, F_Y, F_Z};

struct __align__(128) FullScratch {
    double ring[OPAL_FULL_STAGES][FULL_NIN][32];
    CellPre pre;
    double raw[20];
    unsigned long long bar[OPAL_FULL_STAGES];
};

// photon_emission::rate (qed/photon_emission.rs:59-79) times gamma: sqrt(3) alpha chi h(chi) / (2 pi tau_C); the caller
// multiplies by 1/gamma.  Same three branches of h as qed_rate (opal_device.cuh).
__device__ __forceinline__ double qed_rate_times_gamma(double chi, uint32_t classical, const double* __restrict__ ln_h) {
    double h;
    if (classical || chi < 0.01) {
        h = classical ? 5.0 * OPAL_FRAC_PI_3 : 5.0 * OPAL_FRAC_PI_3 * fma(chi, -8.0 / (5.0 * 1.7320508075688772), 1.0);
    } else if (chi >= 100.0) {
        const double cb = cbrt(chi);
        h = -1019.4661473121777 + 1786.716527650374 * cb * cb;
        h = 1750.6263395722715 + cb * cb * h;
        h = -2260.1819695887225 + cb * h;
        h = 0.00296527643253334 * h / (chi * chi);
    } else {
        const double index = (log(chi) - OPAL_LN_H_X0) * (1.0 / OPAL_DELTA_LN_CHI);
        const double fi = floor(index);
        int i = (int)fi;
        i = i > 39 ? 39 : i;
        const double w = index - fi;
        h = exp(fma(w, ln_h[i + 1] - ln_h[i], ln_h[i]));
    }
    return (1.7320508075688772 * OPAL_ALPHA_FINE / (2.0 * OPAL_PI * OPAL_COMPTON_TIME)) * chi * h;
}

template <bool BX0, bool DEP>
__global__ void __launch_bounds__(OPAL_FULL_WPB * 32, OPAL_FULL_MINBLK)
k_push_full_fast(GridDev g, BucketDev b, StepParams sp, opal_migrant* __restrict__ movers, int* __restrict__ seg_count, StatusDev* __restrict__ st) {
    if (g.halt && *(volatile const int*)g.halt) return;   // burst mode: an earlier kernel needs the host first
    constexpr int NST = OPAL_FULL_STAGES;
    extern __shared__ __align__(128) unsigned char s_dyn[];
    FullScratch& S = reinterpret_cast<FullScratch*>(s_dyn)[threadIdx.x >> 5];
    __shared__ double s_ln_h[41];
    if (threadIdx.x < 41) s_ln_h[threadIdx.x] = c_ln_h[threadIdx.x];
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = lanemask_lt();
    if (elect_one()) {
#pragma unroll
        for (int k = 0; k < NST; k++) mbar_init(&S.bar[k], 1);
        mbar_fence_init();
    }
    __syncthreads();
    const int warps_total = gridDim.x * OPAL_FULL_WPB;
    const double alpha = OPAL_QE * g.dt / (2.0 * OPAL_ME * OPAL_C);
    const double ac = alpha * OPAL_C;
    const double kdt = OPAL_C * g.dt / g.dx;
    const double k_work = OPAL_QE * OPAL_C * g.dt / alpha;          // work += q c (u_half . E) dt / gamma_half, E = (alpha E) / alpha
    const double k_chi = 1.0 / (fabs(alpha) * OPAL_ECRIT);
    const double cdt = OPAL_C * g.dt;
    const uint32_t classical = sp.flags & 1u;                       // OPAL_FLAG_NO_RADIATION_REACTION: classical rate (electron.rs:284)
    bool cfl = false, cap_err = false;
    unsigned phase = 0;
    const double* in[FULL_NIN];
#pragma unroll
    for (int k = 0; k < FULL_NIN; k++) in[k] = b.f[c_full_in[k]];

    for (int c = blockIdx.x * OPAL_FULL_WPB + (threadIdx.x >> 5); c < b.ncell; c += warps_total) {
        const int n = b.count[c];
        if (n == 0) continue;
        const long long base = b.start[c];
        const int j = c + g.left;
        const int n_chunks = (n + 31) >> 5;
        __syncwarp();
        if (elect_one()) {
#pragma unroll
            for (int s_ = 0; s_ < NST - 1; s_++) {
                if (s_ < n_chunks) {
                    mbar_expect_tx(&S.bar[s_], FULL_NIN * 256);
#pragma unroll
                    for (int k = 0; k < FULL_NIN; k++) bulk_g2s(S.ring[s_][k], in[k] + base + s_ * 32, 256, &S.bar[s_]);
                }
            }
        }
        if (lane < 3) { S.raw[lane] = __ldg(g.ex + j - 1 + lane); S.raw[12 + lane] = __ldg(g.by + j - 1 + lane); S.raw[15 + lane] = __ldg(g.bz + j - 1 + lane); }
        if (lane < 4) { S.raw[3 + lane] = __ldg(g.ey + j - 1 + lane); S.raw[7 + lane] = __ldg(g.ez + j - 1 + lane); }
        if (lane == 0) { S.raw[11] = BX0 ? 0.0 : __ldg(g.bx + j); S.raw[18] = 0.0; }
        __syncwarp();
        if (lane < 18) {
            const int k = c_pre_k[lane];
            const double coef = (k & 2 ? ac : alpha) * (k & 1 ? 0.5 : 1.0);
            reinterpret_cast<double*>(&S.pre)[lane] = coef * (S.raw[c_pre_ia[lane]] - S.raw[c_pre_ib[lane]]);
        }
        __syncwarp();
        double sjx0 = 0.0, sjx1 = 0.0, sjx2 = 0.0, sjy0 = 0.0, sjy1 = 0.0, sjy2 = 0.0, sjz0 = 0.0, sjz1 = 0.0, sjz2 = 0.0;
        double sr0 = 0.0, sr1 = 0.0, sr2 = 0.0, sr3 = 0.0;
        int written = 0, stage = 0;
        for (int k0 = 0; k0 < n; k0 += 32) {
            mbar_wait(&S.bar[stage], (phase >> stage) & 1u);
            phase ^= 1u << stage;
            __syncwarp();                                      // every lane is done with the stage the next copy refills
            if (elect_one()) {
                const int nxt = (k0 >> 5) + NST - 1;
                if (nxt < n_chunks) {
                    const int ps = stage == 0 ? NST - 1 : stage - 1;
                    mbar_expect_tx(&S.bar[ps], FULL_NIN * 256);
#pragma unroll
                    for (int k = 0; k < FULL_NIN; k++) bulk_g2s(S.ring[ps][k], in[k] + base + (long long)nxt * 32, 256, &S.bar[ps]);
                }
            }
            const int i = k0 + lane;
            const bool valid = i < n;
            double (*R)[32] = S.ring[stage];
            const double x = valid ? R[0][lane] : 0.5;
            V3 u = {valid ? R[1][lane] : 0.0, valid ? R[2][lane] : 0.0, valid ? R[3][lane] : 0.0};
            const double W = valid ? R[4][lane] : 0.0;
            const double gamma_in = valid ? R[5][lane] : 1.0;
            double tau = R[6][lane], work = R[7][lane], y = R[8][lane], z = R[9][lane];
            stage = stage == NST - 1 ? 0 : stage + 1;

            // ---- gather (yee.rs:499-529) ----
            const double2* const P2 = reinterpret_cast<const double2*>(&S.pre);
            const double2 p0 = P2[0], p1 = P2[1], p2 = P2[2], p3 = P2[3], p4 = P2[4], p5 = P2[5], p6 = P2[6], p7 = P2[7], p8 = P2[8];
            const RFaces ro = r_faces(x);
            const double xm = 1.0 - x, xm2 = xm * xm, x2 = x * x;
            V3 aE, tv;
            aE.x = fma(x2, p1.x, fma(xm2, p0.y, p0.x));
            aE.y = fma(ro.c, p3.x, fma(ro.b, p2.y, fma(ro.a, p2.x, p1.y)));
            aE.z = fma(ro.c, p5.x, fma(ro.b, p4.y, fma(ro.a, p4.x, p3.y)));
            tv.x = BX0 ? 0.0 : p8.y;
            tv.y = fma(x2, p6.y, fma(xm2, p6.x, p5.y));
            tv.z = fma(x2, p8.x, fma(xm2, p7.y, p7.x));
            // ---- Vay push (electron.rs:268-330) ----
            const double ig = fast_rcp(gamma_in);                               // v = c u / gamma (:273)
            const V3 uxt = BX0 ? V3{u.y * tv.z - u.z * tv.y, -(u.x * tv.z), u.x * tv.y} : cross(u, tv);
            const V3 u_half = {fma(ig, uxt.x, u.x + aE.x), fma(ig, uxt.y, u.y + aE.y), fma(ig, uxt.z, u.z + aE.z)};
            double igh, gh;
            fast_rsqrt_sqrt(fma(u_half.x, u_half.x, fma(u_half.y, u_half.y, fma(u_half.z, u_half.z, 1.0))), igh, gh);
            const double eu = dot(u_half, aE);                                  // alpha (E . u_half)
            work = fma(k_work * eu, igh, work);                                 // :278
            {
                const V3 hxt = BX0 ? V3{u_half.y * tv.z - u_half.z * tv.y, -(u_half.x * tv.z), u_half.x * tv.y} : cross(u_half, tv);
                const V3 f = {fma(gh, aE.x, hxt.x), fma(gh, aE.y, hxt.y), fma(gh, aE.z, hxt.z)};      // alpha (gamma_half E + c u_half x B)
                const double chi2 = fmax(dot(f, f) - eu * eu, 0.0);
                const double chi = chi2 > 0.0 ? fast_sqrt(chi2) * k_chi : 0.0;                      // :282
                tau = fma(-(qed_rate_times_gamma(chi, classical, s_ln_h) * igh), g.dt, tau);         // :287
                // (chi is stored below)
                y = fma(cdt * ig, u.y, y);                                       // :315-316: the OLD velocity
                z = fma(cdt * ig, u.z, z);
                const V3 u_prime = u_half + aE;
                const double gp2 = fma(u_prime.x, u_prime.x, fma(u_prime.y, u_prime.y, fma(u_prime.z, u_prime.z, 1.0)));
                const double u_star = BX0 ? fma(u_prime.y, tv.y, u_prime.z * tv.z) : dot(u_prime, tv);
                const double tau2 = BX0 ? fma(tv.y, tv.y, tv.z * tv.z) : dot(tv, tv);
                const double hs = 0.5 * (gp2 - tau2);
                const double gam2 = hs + fast_sqrt(fma(hs, hs, fma(u_star, u_star, tau2)));
                double igam, gam;
                fast_rsqrt_sqrt(gam2, igam, gam);
                const double inv = fast_rcp(gam2 + tau2);
                const V3 pxt = BX0 ? V3{u_prime.y * tv.z - u_prime.z * tv.y, -(u_prime.x * tv.z), u_prime.x * tv.y} : cross(u_prime, tv);
                V3 w = {fma(gam, pxt.x, gam2 * u_prime.x), fma(gam, pxt.y, gam2 * u_prime.y), fma(gam, pxt.z, gam2 * u_prime.z)};
                if (!BX0) w.x = fma(u_star, tv.x, w.x);
                w.y = fma(u_star, tv.y, w.y); w.z = fma(u_star, tv.z, w.z);
                u = inv * w;
                const double dxi = (kdt * igam) * u.x;
                if (valid && !(dxi < 1.0)) cfl = true;
                double xn = x + dxi;
                const double fl = floor(xn);
                xn -= fl;
                const double prev_x = x - fl;                                    // :319-329
                const bool stay = valid && fl == 0.0, move = valid && fl != 0.0;
                const unsigned m_stay = __ballot_sync(FULLMASK, stay);
                const unsigned m_move = __ballot_sync(FULLMASK, move);
                // reserve mover slots (one atomic per chunk that has movers, on one of OPAL_MOVER_SEGMENTS counters)
                int mbase = 0;
                const int seg = (int)(((unsigned)c * 0x9E3779B1u + (unsigned)(k0 >> 5) * 0x85EBCA77u) >> 22);
                if (m_move) {
                    const int leader = __ffs(m_move) - 1;
                    if (lane == leader) mbase = atomicAdd(seg_count + seg, __popc(m_move));
                    mbase = __shfl_sync(FULLMASK, mbase, leader);
                }
                // electrons whose optical depth has run out (electron.rs:209) and that stay here: listed for k_emit
                const unsigned m_q = b.eq_n ? __ballot_sync(FULLMASK, stay && tau < 0.0) : 0u;
                int qbase = 0;
                if (m_q) {
                    const int leader = __ffs(m_q) - 1;
                    if (lane == leader) qbase = atomicAdd(b.eq_n, __popc(m_q));
                    qbase = __shfl_sync(FULLMASK, qbase, leader);
                }
                if (stay) {
                    const int pos = written + __popc(m_stay & lt_mask);
                    const long long s = base + pos;
                    if ((m_q >> lane) & 1u) {
                        const int k = qbase + __popc(m_q & lt_mask);
                        if (k >= 0 && k < b.eq_cap) { b.eq_slot[k] = s; b.eq_cell[k] = c; }
                    }
                    b.f[F_X][s] = xn; b.f[F_PREVX][s] = prev_x; b.f[F_UX][s] = u.x; b.f[F_UY][s] = u.y; b.f[F_UZ][s] = u.z;
                    b.f[F_GAMMA][s] = gam; b.f[F_CHI][s] = chi; b.f[F_TAU][s] = tau; b.f[F_WORK][s] = work; b.f[F_Y][s] = y; b.f[F_Z][s] = z;
                    if (pos != i) b.f[F_W][s] = W;
                    if (DEP) {
                        const RFaces rn = r_faces(xn);
                        sjx0 = fma(W, rn.a - ro.a, sjx0); sjx1 = fma(W, rn.b - ro.b, sjx1); sjx2 = fma(W, rn.c - ro.c, sjx2);
                        const double w0 = 1.0 - rn.a, w1 = rn.a - rn.b, w2 = rn.b - rn.c, w3 = rn.c;
                        sr0 = fma(W, w3, sr0); sr1 = fma(W, w0, sr1); sr2 = fma(W, w1, sr2); sr3 = fma(W, w2, sr3);
                        const double tw = W * igam, ty = tw * u.y, tz = tw * u.z;
                        sjy0 = fma(ty, w0, sjy0); sjy1 = fma(ty, w1, sjy1); sjy2 = fma(ty, w2, sjy2);
                        sjz0 = fma(tz, w0, sjz0); sjz1 = fma(tz, w1, sjz1); sjz2 = fma(tz, w2, sjz2);
                    }
                } else if (move) {
                    const int r = mbase + __popc(m_move & lt_mask);
                    if (r < sp.mover_cap) {
                        opal_migrant m;
                        m.cell = (long long)c + (fl < 0.0 ? -1 : 1); m.x = xn; m.prev_x = prev_x; m.u[0] = u.x; m.u[1] = u.y; m.u[2] = u.z;
                        m.weight = W; m.gamma = gam; m.chi = chi; m.tau = tau; m.work = work; m.y = y; m.z = z;
                        m.birth_time = 0.0; m.pad[0] = 0.0; m.pad[1] = 0.0;
                        movers[(long long)seg * sp.mover_cap + r] = m;
                    } else cap_err = true;
                }
                written += __popc(m_stay);
            }
        }
        if (lane == 0) b.count[c] = written;
        if (DEP) {
            const double qsp = OPAL_QE;
            const double kx = qsp / g.dt, kr = qsp / g.dx, ky = qsp * OPAL_C / g.dx;
            double v[13] = {kx * sjx0, kx * sjx1, kx * sjx2, ky * sjy0, ky * sjy1, ky * sjy2, ky * sjz0, ky * sjz1, ky * sjz2, kr * sr0, kr * sr1, kr * sr2, kr * sr3};
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
