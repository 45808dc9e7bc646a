/* opal_b200.h -- plain-C ABI of the B200-native replacement for opal's per-timestep PIC loop.
 *
 * The reference (tgblackburn/opal, a Rust binary crate) has no FFI of its own; the seam this
 * library replaces is the set of Rust calls main() makes inside the step loop
 * (src/main.rs:238-267).  One opal_ctx == one MPI rank of the reference == one x-slab of the
 * grid plus the particles inside it, resident on one GPU.  Every entry point below cites the
 * reference interface it stands in for.  C99-clean: a Rust `extern "C"` block (see
 * INTEGRATION.md) binds it unchanged.
 *
 * Conventions
 *   - every call returns opal_status; 0 == success.  Device-side assertion failures
 *     (the reference's assert!/panic sites) set a sticky status that the next call returns.
 *   - all pointers are HOST pointers owned by the caller unless the name says `dev`.
 *   - a context is not re-entrant; call from one host thread (reference: Threading::Funneled,
 *     src/main.rs:49).
 *   - there is no CPU fallback: without a CUDA device opal_create fails with OPAL_ECUDA.
 *
 * The CPU oracle (oracle/, test infrastructure only) exports the same symbols with the prefix
 * `opal_ref_` so one harness drives both.
 */
#ifndef OPAL_B200_H
#define OPAL_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef OPAL_PREFIX
#define OPAL_PREFIX opal_
#endif
#define OPAL_CAT2(a, b) a##b
#define OPAL_CAT(a, b) OPAL_CAT2(a, b)
#define OPAL_FN(name) OPAL_CAT(OPAL_PREFIX, name)

typedef struct opal_ctx opal_ctx;

typedef enum {
    OPAL_OK = 0,
    OPAL_EINVAL = 1,        /* bad argument / call not valid in this mode                         */
    OPAL_ECUDA = 2,         /* CUDA runtime error or no device                                     */
    OPAL_ENCCL = 3,         /* NCCL error / communicator not initialised                           */
    OPAL_ENOMEM = 4,
    OPAL_EASSERT_CFL = 5,   /* assert!(dxi < 1.0): electron.rs:313, ion.rs:199, photon.rs:164      */
    OPAL_EASSERT_INDEX = 6, /* index asserts: yee.rs:501-502, yee.rs:578-579                        */
    OPAL_ECHI_RANGE = 7,    /* unimplemented!() for chi beyond the table: photon_emission.rs:144   */
    OPAL_ECAPACITY = 8,     /* a fixed-capacity exchange buffer overflowed                         */
    OPAL_EASSERT_PROB = 9   /* assert!(partial_prob >= 0.0): photon.rs:325, photon.rs:336          */
} opal_status;

/* grid/mod.rs:22-28 */
typedef enum {
    OPAL_BDY_INTERNAL = 0, OPAL_BDY_LASER = 1, OPAL_BDY_ABSORBING = 2, OPAL_BDY_CONDUCTING = 3
} opal_boundary;

/* the three Population<T> instantiations of src/main.rs:119-170 */
typedef enum { OPAL_ELECTRON = 0, OPAL_ION = 1, OPAL_PHOTON = 2 } opal_species;

enum { OPAL_LEFT = 0, OPAL_RIGHT = 1 };

/* flags: the reference's cargo features (Cargo.toml:24-30) as runtime bits + layout choices */
enum {
    OPAL_FLAG_NO_RADIATION_REACTION = 1u << 0,
    OPAL_FLAG_NO_BEAMING            = 1u << 1,
    OPAL_FLAG_IMMOBILE_PHOTONS      = 1u << 2,
    OPAL_FLAG_NO_STIMULATED_EMISSION = 1u << 3,
    /* EXTENSION, off by default, no counterpart in the reference (north_star names it under src/qed; opal only carries the
       unused optical depth Photon.tau, photon.rs:27,126-132 -- parity unpinned): nonlinear Breit-Wheeler pair creation.
       Every photon advance lowers Photon.tau by W dt, W = alpha_f F(chi) / (tau_C k0) (Baier-Katkov / Ritus; tables from
       tools/gen_bw_tables.py); a photon whose depth runs out is replaced by an electron with momentum f k and a positron
       with (1 - f) k (f sampled from dW/df; momentum conserved exactly) at its position, both with its weight and fresh
       optical depths.  POSITRONS LIVE IN THE ION SLOT (Z* = +1, A = m_e / m_p: Boris push ion.rs:168-214, deposit, migration,
       outputs as `ions`; they do not radiate): the flag therefore needs a run without ions, and opal_set_ion is refused.
       With the flag off every result is bit-identical to a build without the extension.                                     */
    OPAL_FLAG_PAIR_CREATION         = 1u << 4,
    /* Charged species keep only (x, u, weight) per particle; prev_x, gamma and velocity are
       recomputed in registers by the fused push+deposit kernel and work/chi/tau/y/z (which are
       unobservable with QED and such outputs off) are not maintained.  Only opal_run_steps()
       is valid on such a context.  This is the 80 B/particle-step layout of SURVEY 8(d).      */
    OPAL_FLAG_LEAN_STATE            = 1u << 8,
    /* Use the kernels compiled with --fmad=false (no FMA contraction; mirrors rustc).          */
    OPAL_FLAG_STRICT_FP             = 1u << 9
};

#define OPAL_GHOST_SIZE 4            /* yee.rs:239 */
#define OPAL_LASER_BDY_SIZE 4        /* yee.rs:240 */
#define OPAL_ABSORBING_BDY_SIZE 200  /* yee.rs:241 */
#define OPAL_CONDUCTING_BDY_SIZE 4   /* yee.rs:242 */
#define OPAL_CELL_DOUBLES 11         /* struct Cell {x, rho, j[3], E[3], B[3]}  yee.rs:86-92 */
#define OPAL_HALO_CELLS 8            /* 2*GHOST_SIZE cells each way, yee.rs:353-354 */

/* One boundary-crossing particle on the wire (all species; unused fields are 0).  Replaces the
   MPI struct datatypes of electron.rs:35-69, ion.rs:34-66, photon.rs:40-80.                 */
typedef struct {
    int64_t cell;        /* already re-indexed for the receiver's left edge (mod.rs:291-293) */
    double x, prev_x;
    double u[3];         /* p/mc  (photons: k/mc)                                              */
    double weight;
    double gamma;        /* ions: 1 + gamma_m1                                                 */
    double chi, tau, work;
    double y, z;
    double birth_time;   /* photons only                                                       */
    double pad[2];       /* ions: gamma_m1 in pad[0]; photons: tau_abs, tau_st (photon.rs:28-29)  */
} opal_migrant;          /* 128 bytes */

/* ---- lifetime ------------------------------------------------------------------------- */

/* One slab = GridDesign entry `rank` (grid/mod.rs:126-206) + YeeGrid::build (yee.rs:272-319).
   `xmin` is the slab's own left edge (global xmin + offset[rank]); `left`/`right` are the
   global boundaries: as in yee.rs:278-292 only rank 0 takes the Laser boundary and only rank
   nranks-1 the right boundary, every other side is Internal with ring neighbours
   (rank +- 1) mod nranks (yee.rs:309-310).  nranks == 1 with Internal boundaries is the
   periodic self-neighbour case (yee.rs:369).                                                 */
opal_status OPAL_FN(create)(opal_ctx** out, int device, int64_t nx, double xmin, double dx, double dt,
                            int left, int right, int rank, int nranks, uint64_t seed, uint32_t flags);
opal_status OPAL_FN(destroy)(opal_ctx*);
const char* OPAL_FN(last_error)(opal_ctx*);
/* total cells incl. boundary zones, and the two zone sizes (yee.rs:316) */
opal_status OPAL_FN(grid_shape)(opal_ctx*, int64_t* ncells_total, int64_t* left_size, int64_t* right_size);

/* ---- state transfer ------------------------------------------------------------------- */

/* Population<T>::new's inner T::create(...) (electron.rs:85, ion.rs:81, photon.rs:95): gamma and
   prev_x are derived exactly as create() does; tau NULL => +inf (electron.rs:99).  `cell` is
   slab-local.  Any order; the library buckets by cell.  Replaces the species' previous content.
   n == 0: on a lone slab (nranks == 1) the species is absent and later calls skip it; on a slab of a
   multi-slab run it is "present but empty here": the slab still takes part in every per-step
   neighbour exchange (mod.rs:308-354).  A species must be uploaded (possibly with n == 0) on EVERY
   slab or on none -- mixing the two deadlocks the ring, exactly as a rank skipping
   Population::advance would in the reference.                                                  */
opal_status OPAL_FN(particles_upload)(opal_ctx*, int species, int64_t n, const int64_t* cell,
                                      const double* x, const double* ux, const double* uy, const double* uz,
                                      const double* weight, const double* tau);
/* Ion::with_charge_state (ion.rs:236): Z* and A, uniform over the species (main.rs:145-147) */
opal_status OPAL_FN(set_ion)(opal_ctx*, double z_star, double mass_number);
opal_status OPAL_FN(particles_count)(opal_ctx*, int species, int64_t* n);
/* Population<T>::all() after advance (sorted by cell, mod.rs:288; order inside a cell is
   unspecified as with the reference's unstable sort).  Any output pointer may be NULL.
   In lean mode gamma/prev_x/... are returned as recomputed/zero.                             */
opal_status OPAL_FN(particles_download)(opal_ctx*, int species, int64_t cap, int64_t* cell, double* x,
                                        double* prev_x, double* ux, double* uy, double* uz, double* weight,
                                        double* gamma, double* chi, double* tau, double* work,
                                        double* y, double* z, double* birth_time);
/* AoS Cell image, ncells_total x 11 doubles (yee.rs:86-92), including ghost/boundary cells */
opal_status OPAL_FN(fields_upload)(opal_ctx*, const double* cells11);
opal_status OPAL_FN(fields_download)(opal_ctx*, double* cells11);

/* ---- the loop body, one call per reference call (main.rs:238-267) ------------------------ */

/* Population<T>::advance (particle/mod.rs:247): gather (yee.rs:499) + push (electron.rs:268 Vay,
   ion.rs:168 Boris, photon.rs:150) + re-bucket by cell (mod.rs:288) + boundary migration
   (mod.rs:290-373).  With a self neighbour or open boundaries this completes the migration.
   With remote neighbours it uses the NCCL communicator if one was attached (comm_init), else
   leavers wait in the outbox for migrants_take()/migrants_give().                             */
opal_status OPAL_FN(advance)(opal_ctx*, int species);
/* (gone_left, gone_right) of the last advance -- mod.rs:284, part of the bit-exact contract   */
opal_status OPAL_FN(last_leavers)(opal_ctx*, int species, int64_t* gone_left, int64_t* gone_right);
/* host-staged migration: the send_left/send_right slices of mod.rs:296-297 ...                */
opal_status OPAL_FN(migrants_take)(opal_ctx*, int species, int side, int64_t cap, int64_t* n, opal_migrant* out);
/* ... and the recv_left/recv_right splices of mod.rs:366-373 (side = where they came FROM)    */
opal_status OPAL_FN(migrants_give)(opal_ctx*, int species, int side, int64_t n, const opal_migrant* in);

/* emit_radiation (particle/interactions.rs:45) incl. Electron::radiate (electron.rs:208);
   NULL pointer == None for the three optional filters.                                         */
opal_status OPAL_FN(emit_radiation)(opal_ctx*, double t, const double* min_energy_mev,
                                    const double* max_angle, const double* max_formation_length);

/* absorb (particle/interactions.rs:145) incl. Photon::interacts_with (photon.rs:312) and
   Electron::kick (electron.rs:256): every photon walks the electrons of its cell, spending its two optical
   depths; an absorbed photon is deleted and its momentum given to the electron, a stimulated emission
   adds a copy of the photon (weight of the electron) and takes its momentum from the electron.
   NULL == None for the two optional limits.  Call it where main.rs:246 does: after both advances.
   Reference behaviour kept: the electrons of the slab's LAST occupied cell never interact
   (interactions.rs:216-222 needs a following electron to close the range).                    */
opal_status OPAL_FN(absorb)(opal_ctx*, double t, const double* max_displacement, const double* stop_time);
/* makes opal_run_steps call absorb after the three advances of every step, where main.rs:246-248 does
   (`if photon_absorption { absorb(...) }`), with these two limits (NULL == None)                  */
opal_status OPAL_FN(set_absorption)(opal_ctx*, int enabled, const double* max_displacement, const double* stop_time);
/* Photon::tau_abs / tau_st (photon.rs:28-29) in the order particles_download(OPAL_PHOTON) uses;
   set != 0 writes them (n must equal the photon count), set == 0 reads them                     */
opal_status OPAL_FN(photon_pair_depths)(opal_ctx*, int set, int64_t n, double* tau_abs, double* tau_st);

opal_status OPAL_FN(clear)(opal_ctx*);                 /* Grid::clear    yee.rs:543 */
opal_status OPAL_FN(deposit)(opal_ctx*, int species);  /* Grid::deposit  yee.rs:551 */

/* host-staged halo: the send_left/send_right copies of yee.rs:353-354 (8 cells x 11 doubles) */
opal_status OPAL_FN(halo_take)(opal_ctx*, int side, double* cells8x11);
/* ... and what the neighbour on `side` sent us (recv_left/recv_right, yee.rs:355-409)         */
opal_status OPAL_FN(halo_give)(opal_ctx*, int side, const double* cells8x11);
/* Grid::synchronize (yee.rs:351): ghost overlay (self-neighbour: local; remote: NCCL if
   attached, else the halo_give()n strips) then laser / absorbing / conducting boundary
   conditions; the host evaluates the two laser closures at (t, cell[2].x) (yee.rs:456-462).    */
opal_status OPAL_FN(synchronize)(opal_ctx*, double t, double laser_y, double laser_z);
/* x of the laser injection cell, cell[left_bdy_size-2].x (yee.rs:457-458)                     */
opal_status OPAL_FN(laser_x)(opal_ctx*, double* x);
opal_status OPAL_FN(field_advance)(opal_ctx*);         /* Grid::advance  yee.rs:345 */

/* The whole loop body main.rs:238-267 repeated nsteps times without returning to the host:
   advance(e), advance(i), advance(ph), [absorb], [emit], [clear, deposit(e), deposit(i)], synchronize,
   field advance, t += dt.  laser_y/laser_z: nsteps precomputed values (NULL => no laser).
   emit_filters3: {min_energy_mev, max_angle, max_formation_length}, NaN == None; NULL == all None.
   When deposition is on and emission off the push and deposit of each charged species run as
   ONE fused kernel.                                                                          */
opal_status OPAL_FN(run_steps)(opal_ctx*, int64_t nsteps, double t0, const double* laser_y, const double* laser_z,
                               int current_deposition, int photon_emission, const double* emit_filters3);

/* Local (this slab) contributions to write_energies (main.rs:23-42): em_field_energy
   (yee.rs:787) and total_kinetic_energy per species (mod.rs:227; electron.rs:122, ion.rs:132,
   photon.rs:224), joules.  The caller sums over slabs (the reference: reduce_into_root).      */
opal_status OPAL_FN(energy)(opal_ctx*, double* em, double* electrons, double* ions, double* photons);

/* ---- output: distribution functions (particle/hgram.rs, particle/mod.rs:383-568) ---- */
/* the per-particle outputs of Population::write_data (mod.rs:388-449), in the reference's units */
typedef enum { OPAL_Q_X = 0, OPAL_Q_R, OPAL_Q_ENERGY, OPAL_Q_PX, OPAL_Q_PY, OPAL_Q_PZ, OPAL_Q_P_PERP, OPAL_Q_THETA,
               OPAL_Q_PHI, OPAL_Q_LONGITUDE, OPAL_Q_LATITUDE, OPAL_Q_WORK, OPAL_Q_CHI, OPAL_Q_COUNT_ } opal_quantity;
typedef enum { OPAL_W_WEIGHT = 0, OPAL_W_ENERGY = 1 } opal_hgram_weight;   /* mod.rs:496-500 */
/* Pass 1 of Histogram::generate_1d/_2d (hgram.rs:103-124,186-204,290-316): this slab's min and max of
   each of the ndim (1 or 2) axis quantities over its FINITE values -- ln() applied first on a log-scaled
   axis -- and its particle count.  An empty slab returns (DBL_MAX, -DBL_MAX, 0) like hgram.rs:189.
   The caller reduces (min, max, sum) over slabs as the reference does with all_reduce.            */
opal_status OPAL_FN(hgram_minmax)(opal_ctx*, int species, int ndim, const int* quantity, const int* log_scaled,
                                  double* min, double* max, int64_t* n);
/* Pass 2 (hgram.rs:222-258, 333-366): bin this slab's particles on the GLOBAL grid (gmin, bin_sz, nbins per
   axis; bin_sz == 0 => bin 0).  counts[] (nbins[0]*nbins[1] doubles, host) receives the weighted counts of
   this slab BEFORE the division by the bin volume; flat index = bin[1]*nbins[1] + bin[0] as hgram.rs:357
   (1-D: bin[0]); indices >= the array length are dropped (a value equal to the maximum, hgram.rs:253).
   *total = sum of all weights, binned or not (hgram.rs:232).  log_density_adjust != 0 applies the
   re-weighting of log-scaled axes, hgram.rs:236-238 / 349-355.                                     */
opal_status OPAL_FN(hgram_bin)(opal_ctx*, int species, int ndim, const int* quantity, const int* log_scaled,
                               const double* gmin, const double* bin_sz, const int64_t* nbins, int weight_kind,
                               int log_density_adjust, double* counts, double* total);
/* opal_hgram_bin for callers that may not allocate device memory (code under CUDA graph capture, tight
   serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
   Same arguments and results as opal_hgram_bin, but the device-side bin counters live
   in a static device array of OPAL_HGRAM_STATIC_MAX_BINS + 1 doubles (512 KiB, held by every GPU that loads
   the library until exit) instead of the scratch memory opal_hgram_bin allocates (and grows) on demand; the
   call makes no cudaMalloc / cudaFree.
   nbins[0]*nbins[1] > OPAL_HGRAM_STATIC_MAX_BINS returns OPAL_ECAPACITY (message in opal_last_error; the
   context stays usable).  A lean-layout species whose cell hand-over is still pending (the step just run
   left movers in the strips; settling them may re-lay the buckets out, which allocates) returns
   OPAL_EINVAL: call any reader (opal_particles_count) once after the steps, then this.  All calls on one
   GPU share the array: it is cleared at the start of every call, and calls that overlap on the GPU (two
   contexts, two host threads) race on it.                                                          */
#define OPAL_HGRAM_STATIC_MAX_BINS 65536
opal_status OPAL_FN(hgram_bin_static)(opal_ctx*, int species, int ndim, const int* quantity, const int* log_scaled,
                                      const double* gmin, const double* bin_sz, const int64_t* nbins, int weight_kind,
                                      int log_density_adjust, double* counts, double* total);

/* ---- multi-GPU: NCCL ring of slabs (replaces rsmpi p2p) -------------------- */
opal_status OPAL_FN(comm_unique_id)(void* id128);                 /* ncclGetUniqueId; 128 bytes */
opal_status OPAL_FN(comm_init)(opal_ctx*, const void* id128);     /* ncclCommInitRank(rank,nranks of create) */

/* ---- instrumentation ----------------------------------------------------------------- */
/* the CUDA stream every kernel of this context is launched on (for cudaEvent timing) */
opal_status OPAL_FN(stream)(opal_ctx*, void** cuda_stream);
/* number of kernels this context has launched so far */
opal_status OPAL_FN(launch_count)(opal_ctx*, int64_t* n);
/* per-kernel device timing: timing(1) makes the context bracket every launch of the three step
   kernels with cudaEvents on its stream; kernel_times() synchronises and returns the summed
   durations [ms] and launch counts for {0: gather+push(+deposit), 1: mover scatter, 2: halo/BC +
   field update} since timing was enabled.                                                      */
opal_status OPAL_FN(timing)(opal_ctx*, int enable);
opal_status OPAL_FN(kernel_times)(opal_ctx*, double ms3[3], int64_t launches3[3]);
/* Steps of opal_run_steps that ran as replayed CUDA graphs ("burst mode": launch-bound decks with deposition on,
   emission off, no remote neighbour, timing off; two steps per graph, status read back once per 128 steps, a
   device-side halt flag freezes the batch where the host must re-lay buckets out).  OPAL_BURST=0 disables it.   */
opal_status OPAL_FN(burst_steps)(opal_ctx*, int64_t* n);
/* device-generated synthetic plasma (BASELINE configs 4/5): npc particles in every cell,
   x uniform in [0,1), u = drift[k]*sign + u_th[k]*N(0,1) from Philox4x32-10 keyed by
   (seed, global particle index) so any slab count yields the same global particle set;
   two_stream != 0 applies two_stream.yaml:18's  u_x = drift*(1+0.001 n)*(2 floor(r-0.5)+1).   */
opal_status OPAL_FN(particles_generate)(opal_ctx*, int species, int64_t npc, double weight,
                                        const double u_th[3], const double drift[3], int two_stream,
                                        int64_t global_cell_offset, uint64_t seed);

/* ---- pure functions of the path, exposed for known-answer tests --------------------- */
/* photon_emission::rate (qed/photon_emission.rs:59), n values */
opal_status OPAL_FN(qed_rate)(int64_t n, const double* chi, const double* gamma, double* out);
/* photon_emission::sample (:129) -> out[3n] = (omega/mc^2, theta, phi) */
opal_status OPAL_FN(qed_sample)(int64_t n, const double* chi, const double* gamma, const double* r1,
                                const double* r2, const double* r3, int classical, double* out3);
/* EXTENSION (OPAL_FLAG_PAIR_CREATION; nothing in the reference): F(chi) = chi T(chi) of nonlinear Breit-Wheeler pair creation --
   the rate is alpha_f F(chi) / (tau_C k0) -- and the electron's share f of the photon energy for uniforms r1, r2 */
opal_status OPAL_FN(bw_rate)(int64_t n, const double* chi, double* out);
opal_status OPAL_FN(bw_sample)(int64_t n, const double* chi, const double* r1, const double* r2, double* out);
/* pairs created so far on this slab (0 unless OPAL_FLAG_PAIR_CREATION) */
opal_status OPAL_FN(pairs_created)(opal_ctx*, int64_t* n);
/* airy_ai_for_positive (qed/special_functions/airy.rs:19): ai[i] = Ai(x[i]); defined[i] = 0 where the
   reference returns None (x < 0, x >= 50 or NaN), and ai[i] = 0 there */
opal_status OPAL_FN(qed_airy)(int64_t n, const double* x, double* ai, int* defined);
/* photon_absorption::scaled_cross_section (qed/photon_absorption.rs:17) and
   stimulated_emission::scaled_cross_section (qed/stimulated_emission.rs:17) for n photon/electron pairs:
   k4, p4 = four-momenta / mc, 4 doubles per pair; defined[i] bit 0: absorption is Some, bit 1: stimulated
   emission is Some (the sigma of a None is written as 0) */
opal_status OPAL_FN(qed_pair_cross_sections)(int64_t n, const double* k4, const double* p4,
                                             const double* chi_gamma, const double* chi_e,
                                             double* sigma_abs, double* sigma_st, int* defined);
/* YeeGrid::fields_at (yee.rs:499) for n (cell, xi) pairs -> out[6n] = E[3],B[3] */
opal_status OPAL_FN(fields_at)(opal_ctx*, int64_t n, const int64_t* cell, const double* xi, double* out6);

#ifdef __cplusplus
}
#endif
#endif /* OPAL_B200_H */
