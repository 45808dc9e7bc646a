// opal_types.h -- plain structs shared by the host library and the kernels.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/opal_b200.h"

namespace opal {

// The mover buffer is split into this many independently counted segments (cell c appends to
// segment c mod N) so the per-chunk slot reservations do not serialise on one L2 atomic address.
#define OPAL_MOVER_SEGMENTS 1024

// particle field indices of a BucketStore (charged species).  The first NF_LEAN are always
// allocated; the rest only for full-state contexts.
enum { F_X = 0, F_UX, F_UY, F_UZ, F_W, NF_LEAN, F_PREVX = NF_LEAN, F_GAMMA, F_CHI, F_TAU, F_WORK, F_Y, F_Z, NF_FULL };
// flat photon store
enum { P_X = 0, P_PREVX, P_KX, P_KY, P_KZ, P_W, P_CHI, P_TAU, P_Y, P_Z, P_BIRTH,
       P_TAU_ABS, P_TAU_ST,      // optical depths against one-photon absorption / stimulated emission (photon.rs:28-29)
       NF_PHOTON };

// device-side view of the Yee grid: SoA over `len` cells (ghost/boundary zones included).
struct StatusDev;
struct GridDev {
    int len, left, right, nx;       // total cells, zone sizes, interior cells
    int left_bdy, right_bdy;        // resolved opal_boundary of this slab
    int self_nbr;                   // nranks == 1: periodic self-neighbour (yee.rs:369)
    double xmin, dx, dt;
    double *rho, *jx, *jy, *jz;     // current set (deposit target / field kernel source)
    double *ex, *ey, *ez, *bx, *by, *bz;
    // graph-replayed bursts of steps (host: run_burst): every step kernel returns at once when *halt is set (a bucket
    // overflowed or an assert fired in an earlier kernel of the burst); the laser amplitudes of step *step are
    // laser_arr[3 * *step .. +1], its time laser_arr[3 * *step + 2].  All three are null outside a burst.
    const int* halt;
    const int* step;
    const double* laser_arr;
    // ... and, to keep a replayed step at a handful of launches, two chores ride on the step kernels themselves:
    StatusDev* bst;                 // the status block (null outside a burst)
    int prev_phase;                 // >= 0: before doing anything, look at what the advance of that phase left behind (parked movers, an
                                    // assert) and raise the halt in its name -- the check a separate k_burst_mark launch would do
    int count_step;                 // k_field: the step is complete, count it
    int seg_cap;                    // room per mover segment: a segment more than half full is a reason to halt, too (the host doubles it)
};

// Cell-major gapped buckets: cell c owns slots [start[c], start[c+1]) of which the first
// count[c] are live.  start[] is 32-slot aligned so a warp's 32-particle chunk is one 256-byte
// aligned segment per field.
struct BucketDev {
    int ncell;
    const long long* start;
    int* count;
    double* f[NF_FULL];
    double z_star, mass_number;     // ions (ion.rs:236)
    // lean strip kernel inside a burst: the cells the one-warp-per-cell launch covers, [c_begin, c_end).  Everything outside is
    // empty, has empty strips in both buffers, and cannot be reached before the burst ends (a particle moves less than a
    // cell per step): host, run_burst.  Every other kernel, and every launch outside a burst, covers [0, ncell).
    int c_begin = 0, c_end = 0;
    // Emitter queue (electrons, full state, emission inside opal_run_steps; null otherwise): the push and the scatter that
    // precede emit_radiation in a step know which electrons' optical depth has run out (electron.rs:209) and where they ended
    // up -- they list (slot, cell), and k_emit samples exactly those with every lane busy instead of scanning tau of all.
    // *eq_n < 0 or > eq_cap: the list is not valid (a re-layout moved the slots / it overflowed) -> k_emit scans.
    long long* eq_slot = nullptr; int* eq_cell = nullptr; int* eq_n = nullptr; int eq_cap = 0;
};

struct PhotonDev {
    int cap;
    int* n;                         // live count (device)
    int* cell;
    double* f[NF_PHOTON];
};

// Breit-Wheeler pair creation inside the photon push (OPAL_FLAG_PAIR_CREATION): pair k of a call is e_recs[k] + p_recs[k]
struct PairDev {
    opal_migrant* e_recs = nullptr; opal_migrant* p_recs = nullptr;
    int* n = nullptr;            // pairs created by this call (null: pair creation off); may run past cap -- then the photon waits a step
    int cap = 0;
    uint64_t seed = 0; unsigned call = 0;
};

// device-side counters / sticky error bits, mirrored to pinned host memory after each step
struct StatusDev {
    int n_movers;        // records in the mover buffer
    int n_overflow;      // movers that did not fit their destination bucket
    int n_out[2];        // records in the left/right outbox
    int gone[2];         // leaver counts of the last advance (mod.rs:284)
    unsigned err;        // bit 0: CFL assert, 1: index assert, 2: chi range, 3: buffer capacity
    int n_new_photons;
    int pad;
    // burst mode (see GridDev::halt)
    int halt;            // set by k_burst_check: the host has to intervene before the next kernel may run
    int step;            // steps of the burst completed so far
    int halt_species;    // the species (phase) whose advance raised the halt
    int heavy_seen;      // lean strip kernel: some cell holds more than StepParams::heavy_hint particles
    int strip_spill[2];  // lean strip kernel, per charged species: movers that found their cell's strip full since the block was last zeroed
    int edge[2];         // multi-slab: particles of the species just advanced that sit in the slab's first / last cell -- all that can
                         // leave through that edge in the NEXT step (|dxi| < 1): the sender sizes its outbox from it
};
enum { ERR_CFL = 1u, ERR_INDEX = 2u, ERR_CHI = 4u, ERR_CAPACITY = 8u,
       // which fixed-capacity buffer overflowed (set together with ERR_CAPACITY)
       ERR_CAP_MOVERS = 16u, ERR_CAP_PARKED = 32u, ERR_CAP_OUTBOX = 64u, ERR_CAP_PHOTONS = 128u,
       ERR_PROB = 256u };     // assert!(partial_prob >= 0.0) (photon.rs:325)

// compact record of one cell-crossing particle in a per-cell, per-direction mover strip (lean layout); the destination
// (c-1 / c+1) is implied by the strip it sits in.  x: offset in the destination cell; prev_x: offset it had in the cell it left
// (the writing warp deposits its movers from the strip at the end of the cell, all lanes busy; the next step's push ignores it).
struct __align__(16) StripRec { double x, prev_x, ux, uy, uz, w; };   // 48 bytes

// the strips of one lean species as one push sees them (double-buffered by step parity): it reads the movers its neighbour
// cells left in `in` during the previous step and writes its own movers to `out`
struct LeanStrips {
    const StripRec* in; const int* scount_in;    // [2 * ncell * MC] records, [2 * ncell] counts: cell c's left strip is 2c, its right strip 2c + 1
    StripRec* out; int* scount_out;
    int* hcount;                                 // [ncell] free slots inside count[c]
};

// internal StepParams.flags bit (above the public OPAL_FLAG_* bits): B_x is zero on the whole grid
#define OPAL_STEP_BX_ZERO (1u << 31)

struct StepParams {
    uint32_t flags;
    int species;             // opal_species
    int route_left, route_right;   // what happens to a leaver: 0 = delete (open boundary), 1 = wrap (self), 2 = outbox
    int mover_cap, out_cap;
    int overflow_cap;        // records in the parked-mover list (destination bucket full)
    int cell_mover_cap;      // records per cell and direction in the mover strips (0: strips not in use)
    // lean strip kernel: cells holding more than heavy_min particles are left to the block-per-cell launch that follows the
    // one-warp-per-cell launch (INT_MAX: there is none); a cell ending the step with more than heavy_hint raises
    // StatusDev::heavy_seen, from which the host decides whether the next steps need that second launch
    int heavy_min, heavy_hint;
};

// one set of kernel launchers; two instances exist (fast: FMA contraction on, strict: off)
struct KernelTable {
    // returns true when the lean strip kernel ran (`ls` given, lean layout, fused deposit, fast build): the movers then
    // wait in ls->out for the NEXT push (or for flush), and the buckets carry free slots (ls->hcount)
    bool (*push)(cudaStream_t, const GridDev&, const BucketDev&, const StepParams&, bool full, bool fused_deposit,
                 opal_migrant* movers, const LeanStrips* ls, opal_migrant* outbox_l, opal_migrant* outbox_r, opal_migrant* overflow,
                 StatusDev* st, double* ke_out /* null, or accumulator for the pushed particles' kinetic energy (strip kernel only) */);
    // append the movers waiting in `strips` to their destination buckets (no deposit: the push did it)
    void (*flush)(cudaStream_t, const BucketDev&, const StepParams&, const StripRec* strips, const int* strip_count,
                  opal_migrant* overflow, StatusDev* st);
    void (*scatter)(cudaStream_t, const GridDev&, const BucketDev&, const StepParams&, bool full, bool deposit,
                    const opal_migrant* recs, const int* n_recs, int n_max, long long cell_shift, const int* idx_list,
                    opal_migrant* outbox_l, opal_migrant* outbox_r, int* overflow_idx, StatusDev* st);
    void (*deposit)(cudaStream_t, const GridDev&, const BucketDev&, const StepParams&, StatusDev* st);
    void (*photon_push)(cudaStream_t, const GridDev&, const PhotonDev& src, const PhotonDev& dst, const StepParams&,
                        opal_migrant* outbox_l, opal_migrant* outbox_r, StatusDev* st, const PairDev& pairs);
    void (*photon_insert)(cudaStream_t, const PhotonDev& dst, const opal_migrant* recs, const int* n_recs, int n_max,
                          long long cell_shift, StatusDev* st);
    // burst_room_need > 0 (bursts): first make sure the photon store has room for that many more photons -- or, when the
    // emitter list is valid, for as many as it lists -- else raise the halt in the name of `room_phase` and do nothing
    void (*emit)(cudaStream_t, const GridDev&, const BucketDev& electrons, const PhotonDev& photons, const StepParams&,
                 const double* qed_tables, double t, double min_energy, double max_angle, double max_fl,
                 uint64_t seed, uint64_t call_index, StatusDev* st, long long burst_room_need, int room_phase);
    void (*sync_bc)(cudaStream_t, const GridDev&, const double* halo_recv /*2 x 8 x 11 or null*/, double laser_y, double laser_z,
                    const double* laser_dev /*2 doubles or null*/);
    void (*field)(cudaStream_t, const GridDev& cur, const GridDev& nxt, bool zero_next_j);
    void (*clear)(cudaStream_t, const GridDev&, StatusDev* begin_mark /* bursts: also zero the per-advance counters; else null */);
    // bursts on one slab: check of the last advance + halo/BC + field update + step counter in one launch (k_grid_step)
    // clear_next: nxt's rho / j arrays are the other source set; clear them and zero the per-advance counters for the next step
    void (*grid_step)(cudaStream_t, const GridDev& cur, const GridDev& nxt, StatusDev* st, int last_species, unsigned* barrier2, int mover_cap, bool clear_next);
    void (*halo_pack)(cudaStream_t, const GridDev&, double* out2x8x11);
    void (*energy)(cudaStream_t, const GridDev&, const BucketDev* electrons, const BucketDev* ions, const PhotonDev* photons,
                   bool full, double* out4);
    void (*generate)(cudaStream_t, const BucketDev&, int species, long long npc, double weight, const double* uth3, const double* drift3,
                     int two_stream, long long global_cell_offset, uint64_t seed, bool full, double dx, double dt);
    void (*gather_out)(cudaStream_t, const BucketDev&, bool full, int species, const long long* out_offset,
                       long long* cell, double* const* out_fields /*NF_FULL device ptrs*/);
    void (*qed_rate)(cudaStream_t, int n, const double* chi, const double* gamma, double* out);
    void (*qed_sample)(cudaStream_t, int n, const double* qed_tables, const double* chi, const double* gamma, const double* r1,
                       const double* r2, const double* r3, int classical, double* out3, StatusDev* st);
    void (*bw_rate)(cudaStream_t, int n, const double* chi, double* out);                 // F(chi) of the pair-creation extension
    void (*bw_sample)(cudaStream_t, int n, const double* chi, const double* r1, const double* r2, double* out);
    void (*fields_at)(cudaStream_t, const GridDev&, int n, const long long* cell, const double* xi, double* out6, StatusDev* st);
    void (*copy_buckets)(cudaStream_t, const BucketDev& src, const BucketDev& dst, bool full);
    // sizing of the mover buffer: out2[0] += particles whose ballistic step x + v_x dt/dx leaves [0,1),
    // out2[1] += sum of min(1, |v_x| dt/dx) (the same count for positions uniform inside their cells)
    void (*crossing_estimate)(cudaStream_t, const BucketDev&, double dt_over_dx, double* out2);
    // lean layout: close the free slots the strip kernel leaves inside the buckets
    void (*compact_cells)(cudaStream_t, const BucketDev&, int* hcount);
    // burst mode bookkeeping kernels: 0 begin (zero the per-advance counters), 1 check (raise halt on overflow / error), 2 end (step++)
    // 3 the check of `species` + the begin-mark of the next advance; zero_a / zero_b: ints the begin-mark also zeroes (may be null)
    void (*burst_mark)(cudaStream_t, StatusDev*, int what, int species, int mover_cap, int* zero_a, int* zero_b);
};

namespace fast { KernelTable table(); }
namespace strict { KernelTable table(); }

}  // namespace opal
