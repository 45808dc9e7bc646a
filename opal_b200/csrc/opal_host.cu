// opal_host.cu -- context, memory layout and the C ABI of include/opal_b200.h.
//
// One opal_ctx = one x-slab (one reference MPI rank) on one GPU.  Everything between
// main.rs:240 and main.rs:266 stays on the device; the host contributes the two laser scalars
// per step and reads back a 44-byte status block.  There is no CPU compute path in this file.
#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <limits>
#include <string>
#include <utility>
#include <vector>

#include "opal_types.h"

#define OPAL_QED_TABLE_DATA
static const double H_QED_TABLES[73 * 64] = {
#include "qed_tables.inc"
};
#undef OPAL_QED_TABLE_DATA

using namespace opal;

namespace {

constexpr double C_LIGHT = 2.997925e8;   // constants.rs:4

// ---- NCCL, resolved at run time (torch's bundled libnccl.so.2 is already in the process) -----
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
struct NcclApi {
    void* h = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    bool load() {
        if (h) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
        if (!h) return false;
#define SYM(f) *(void**)(&f) = dlsym(h, "nccl" #f); if (!f) return false
        SYM(GetUniqueId); SYM(CommInitRank); SYM(CommDestroy); SYM(Send); SYM(Recv); SYM(GroupStart); SYM(GroupEnd);
#undef SYM
        return true;
    }
} g_nccl;
constexpr int NCCL_CHAR = 0;   // ncclInt8 / ncclChar

struct Buckets {
    bool allocated = false;
    bool full = false;
    long long slots = 0;              // slots the layout uses (h_start[ncell]) ...
    long long cap_slots = 0;          // ... of the cap_slots each field array holds
    std::vector<long long> h_start;   // ncell + 1
    long long* d_start = nullptr;
    int* d_count = nullptr;
    double* d_f[NF_FULL] = {nullptr};
    long long n_hint = 0;             // particle count at the last layout (sizing of the mover buffer)
    bool movers_sized = false;        // the mover buffer has been sized for this population's crossing rate (size_movers_for)
};

struct Photons {
    int cap = 0;
    int* d_n = nullptr;
    int* d_cell = nullptr;
    double* d_f[NF_PHOTON] = {nullptr};
};

}  // namespace

namespace opal {   // opal_absorb.cu
int absorb_launch(cudaStream_t stream, const GridDev& g, const BucketDev& e, const PhotonDev& src, const PhotonDev& dst, int n0, double t,
                  double max_displacement, double stop_time, unsigned long long seed, unsigned call_index, int no_stimulated, int* d_last,
                  int* d_kind, long long* d_slot, StatusDev* d_status);
}
namespace opal {   // opal_hgram.cu
void hgram_minmax_charged(cudaStream_t, const GridDev&, const BucketDev&, int species, bool full, int ndim, const int* q, const int* lg,
                          unsigned long long* d_mm, double* d_out4);
void hgram_minmax_photon(cudaStream_t, const GridDev&, const PhotonDev&, int ndim, const int* q, const int* lg, unsigned long long* d_mm, double* d_out4);
void hgram_bin_charged(cudaStream_t, const GridDev&, const BucketDev&, int species, bool full, int ndim, const int* q, const int* lg,
                       const double* gmin, const double* bin_sz, const long long* nbins, int weight_kind, int log_adjust, double* d_counts, double* d_total);
void hgram_bin_photon(cudaStream_t, const GridDev&, const PhotonDev&, int ndim, const int* q, const int* lg, const double* gmin,
                      const double* bin_sz, const long long* nbins, int weight_kind, int log_adjust, double* d_counts, double* d_total);
}

struct opal_ctx {
    int device = 0, rank = 0, nranks = 1;
    uint32_t flags = 0;
    uint64_t seed = 0;
    bool full = true;
    bool bx_zero = true;     // no non-zero B_x was ever uploaded (the solver never changes B_x): selects the BX0 push kernel
    cudaStream_t stream = nullptr;
    KernelTable k;
    GridDev g;              // current E/B buffers
    double* eb_alt[6] = {nullptr};   // the other half of the E/B ping-pong
    double* j_alt[4] = {nullptr};    // a second set of rho, jx, jy, jz: in a burst the grid launch of step k clears the set step k + 1 deposits into
    double* grid_block = nullptr;    // one allocation for all grid arrays
    Buckets bk[2];          // electrons, ions
    Buckets spare[2];       // the arrays the last re-layout of each moved out of: the next one moves into them (alloc_buckets)
    double z_star = 1.0, mass_number = 1.0;
    Photons ph[2];          // ping-pong
    int ph_cur = 0;
    // exchange / mover buffers
    opal_migrant* d_movers = nullptr; int mover_seg_cap = 0;
    opal_migrant* d_overflow = nullptr;
    // absorb: per-photon event records (opal_absorb.cu)
    int* d_absorb_kind = nullptr; long long* d_absorb_slot = nullptr; int* d_absorb_last = nullptr; int absorb_cap = 0; long long absorb_calls = 0;
    bool absorb_on = false; double absorb_max_displacement = 0.0, absorb_stop_time = 0.0;   // opal_set_absorption (NaN == None)
    int overflow_cap = 0;
    // lean layout, per charged species: the per-cell left / right mover strips, double-buffered by step parity (the push of
    // step n reads what step n-1 wrote: opal_push_lean.cuh), the strip counts and the per-cell free-slot counts
    struct Lean {
        StripRec* d_strips[2] = {nullptr, nullptr}; int* d_scount[2] = {nullptr, nullptr}; int* d_hcount = nullptr;
        int mc = 0;               // records per cell and direction
        int parity = 0;           // the buffer the NEXT push reads
        bool pending = false;     // movers of the last push wait in d_strips[parity] (not yet in the buckets)
        bool holes = false;       // the buckets carry free slots (count[] counts slots in use)
        bool range_on = false; int c_begin = 0, c_end = 0;   // inside a burst: the cells the push launch covers (BucketDev::c_begin)
        int mc_floor = 0;         // strips at least this wide: raised when movers keep finding them full (maybe_widen_strips)
        long long spill = 0, spill_adv = 0;   // movers that did, and the advances they were counted over
    } lean[2];
    opal_migrant* d_out[2] = {nullptr, nullptr};   // each: [header record][out_cap records]
    opal_migrant* d_in[2] = {nullptr, nullptr};
    int out_cap = 0;           // records the outboxes hold: grown before a push to the edge-cell populations (ensure_outbox)
    int in_cap = 0;            // records the inboxes hold: grown to what the neighbours announce (exchange_migrants)
    long long edge_bound[3][2] = {{-1, -1}, {-1, -1}, {-1, -1}};   // per species and side: most particles that can leave next step; -1 unknown
    int64_t pending_in[3][2] = {{0, 0}, {0, 0}, {0, 0}};   // host-staged migrants given, per species/side (for bookkeeping)
    double* d_halo_send = nullptr;   // [2][8][11]
    double* d_halo_recv = nullptr;   // [2][8][11]
    bool halo_given[2] = {false, false};
    StatusDev* d_status = nullptr;   // followed by OPAL_MOVER_SEGMENTS ints
    StatusDev* h_status = nullptr;   // pinned
    double* d_energy = nullptr;      // 4 doubles + [4]: electron kinetic energy accumulated by the last fused push
    bool ke_fused = false;
    double* d_laser = nullptr; long long laser_cap = 0;   // burst mode: per-step laser amplitudes on the device
    int64_t burst_steps = 0;         // steps executed through replayed CUDA graphs (instrumentation)
                                     // d_energy[4] is the electrons' current total_kinetic_energy
    double* h_energy = nullptr;      // pinned
    int* d_tmp_int = nullptr;
    void* d_hgram = nullptr; size_t hgram_cap = 0;   // scratch of the histogram passes (grow-only: an output writes dozens of histograms)
    unsigned* d_gridbar = nullptr;   // k_grid_step's grid-wide barrier: arrival count, generation
    double* d_qed_tables = nullptr;
    int64_t last_gone[3][2] = {{0, 0}, {0, 0}, {0, 0}};
    uint64_t emit_calls = 0;
    // Breit-Wheeler extension (OPAL_FLAG_PAIR_CREATION): the pairs one photon advance creates, as electron / positron records
    opal_migrant* d_pair[2] = {nullptr, nullptr}; int* d_pair_n = nullptr; int pair_cap = 0; unsigned pair_calls = 0; int64_t pairs_created = 0;
    bool ion_slot_is_user = false;   // opal_set_ion was called: the ion slot holds real ions
    // emitter queue (BucketDev::eq_slot): filled by the electrons' push + scatter of a step of opal_run_steps with emission on
    long long* d_eq_slot = nullptr; int* d_eq_cell = nullptr; int* d_eq_n = nullptr; int eq_cap = 0; bool eq_on = false;
    int64_t launches = 0;
    // lean strip kernel: cells above heavy_min particles go to the block-per-cell launch (opal_push_lean.cuh).  Switched on when a
    // push reports such a cell (StatusDev::heavy_seen), off again after heavy_quiet status reads in a row without one.
    bool widen_strips = true, pool_buckets = true;     // OPAL_STRIP_WIDEN=0 / OPAL_BUCKET_POOL=0 switch them off (diagnosis)
    bool heavy_on = false, heavy_forced = false; int heavy_quiet = 0; int heavy_min = 4096; double heavy_factor = 8.0; int64_t n_heavy_launches = 0;
    int64_t n_regrids = 0, n_halts = 0, n_bursts = 0, n_wasted = 0, n_widened = 0;   // OPAL_STATS=1: printed by opal_destroy
    double t_regrid = 0.0, t_graph = 0.0, t_replay = 0.0, t_rg[5] = {0, 0, 0, 0, 0};   // host wall-clock seconds in re-layouts / graph capture + instantiation / replays + read-backs
    ncclComm_t comm = nullptr;
    // optional per-kernel timing (opal_timing): event pairs per category
    bool timing = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev[3];
    size_t ev_used[3] = {0, 0, 0};
    int status = OPAL_OK;
    std::string err;
    double slack = 1.25; int pad = 64; bool slack_from_env = false;
    // adaptive re-layout (regrid): per charged species, steps since the last one, neighbourhood radius and slack boost
    long long adv_calls[2] = {0, 0}, last_regrid[2] = {-1000000, -1000000};
    int regrid_radius[2] = {0, 0};
    double regrid_boost[2] = {1.0, 1.0};

    bool has_left() const { return g.left_bdy == OPAL_BDY_INTERNAL; }
    bool has_right() const { return g.right_bdy == OPAL_BDY_INTERNAL; }
    bool self_nbr() const { return nranks == 1; }
    int left_rank() const { return (rank - 1 + nranks) % nranks; }    // yee.rs:309
    int right_rank() const { return (rank + 1) % nranks; }            // yee.rs:310
};

namespace {

double wall_now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
struct WallTimer { double& acc; double t0; explicit WallTimer(double& a) : acc(a), t0(wall_now()) {} ~WallTimer() { acc += wall_now() - t0; } };

bool fail(opal_ctx* c, int st, const std::string& msg) {
    if (c && c->status == OPAL_OK) { c->status = st; c->err = msg; }
    return false;
}
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fail(c, OPAL_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); return (opal_status)c->status; } } while (0)
#define CUB(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fail(c, OPAL_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); return false; } } while (0)
#define CHECK_CTX(c) do { if (!(c)) return OPAL_EINVAL; if ((c)->status != OPAL_OK) return (opal_status)(c)->status; cudaSetDevice((c)->device); } while (0)

double env_double(const char* name, double dflt) { const char* v = getenv(name); return v ? atof(v) : dflt; }
long long env_ll(const char* name, long long dflt) { const char* v = getenv(name); return v ? atoll(v) : dflt; }

BucketDev bdev(opal_ctx* c, int sp) {
    BucketDev b;
    Buckets& B = c->bk[sp];
    b.ncell = c->g.nx; b.start = B.d_start; b.count = B.d_count;
    b.c_begin = 0; b.c_end = b.ncell;
    if ((sp == OPAL_ELECTRON || sp == OPAL_ION) && c->lean[sp].range_on && c->g.halt) {       // (set for the launches of a burst only)
        b.c_begin = c->lean[sp].c_begin; b.c_end = c->lean[sp].c_end;
    }
    for (int k = 0; k < NF_FULL; k++) b.f[k] = B.d_f[k];
    b.z_star = c->z_star; b.mass_number = c->mass_number;
    if (sp == OPAL_ELECTRON && c->eq_on && c->d_eq_n) { b.eq_slot = c->d_eq_slot; b.eq_cell = c->d_eq_cell; b.eq_n = c->d_eq_n; b.eq_cap = c->eq_cap; }
    return b;
}

// the emitter queue holds at most one entry per electron of this slab
bool ensure_emit_queue(opal_ctx* c) {
    const long long want = std::max<long long>(c->bk[OPAL_ELECTRON].n_hint + c->bk[OPAL_ELECTRON].n_hint / 4 + 4096, 1 << 16);
    if (c->d_eq_n && want <= c->eq_cap) return true;
    if (want > (1ll << 30)) { c->eq_on = false; return true; }        // k_emit then scans
    if (c->d_eq_slot) cudaFree(c->d_eq_slot);
    if (c->d_eq_cell) cudaFree(c->d_eq_cell);
    c->d_eq_slot = nullptr; c->d_eq_cell = nullptr;
    if (!c->d_eq_n) { CUB(cudaMalloc(&c->d_eq_n, sizeof(int))); CUB(cudaMemset(c->d_eq_n, 0xFF, sizeof(int))); }
    CUB(cudaMalloc(&c->d_eq_slot, (size_t)want * sizeof(long long)));
    CUB(cudaMalloc(&c->d_eq_cell, (size_t)want * sizeof(int)));
    c->eq_cap = (int)want;
    return true;
}
PhotonDev pdev(opal_ctx* c, int which) {
    PhotonDev p;
    Photons& P = c->ph[which];
    p.cap = P.cap; p.n = P.d_n; p.cell = P.d_cell;
    for (int k = 0; k < NF_PHOTON; k++) p.f[k] = P.d_f[k];
    return p;
}
StepParams sparams(opal_ctx* c, int species) {
    StepParams sp;
    sp.flags = c->flags | (c->bx_zero ? OPAL_STEP_BX_ZERO : 0u); sp.species = species;
    sp.route_left = c->has_left() ? (c->self_nbr() ? 1 : 2) : 0;
    sp.route_right = c->has_right() ? (c->self_nbr() ? 1 : 2) : 0;
    sp.mover_cap = c->mover_seg_cap; sp.out_cap = c->out_cap; sp.overflow_cap = c->overflow_cap;
    sp.cell_mover_cap = (species == OPAL_ELECTRON || species == OPAL_ION) ? c->lean[species].mc : 0;
    // "heavy": far above the species' average population per cell (a uniform plasma of 1000 per cell has a million cells to keep
    // the GPU busy with; a compressed layer of a few dozen cells with thousands each has not) and above heavy_min in any case
    long long hm = c->heavy_min;
    if (species == OPAL_ELECTRON || species == OPAL_ION)
        hm = std::max<long long>(hm, (long long)(c->heavy_factor * (double)c->bk[species].n_hint / (double)std::max(c->g.nx, 1)));
    hm = std::min<long long>(hm, INT_MAX / 4);
    sp.heavy_min = c->heavy_on ? (int)hm : INT_MAX;
    sp.heavy_hint = c->heavy_min <= 0 ? INT_MAX : (int)(c->heavy_on ? hm / 2 : hm);     // (hysteresis)
    return sp;
}

void free_buckets(Buckets& B) {
    if (B.d_start) cudaFree(B.d_start);
    if (B.d_count) cudaFree(B.d_count);
    for (int k = 0; k < NF_FULL; k++) if (B.d_f[k]) { cudaFree(B.d_f[k]); B.d_f[k] = nullptr; }
    B.d_start = nullptr; B.d_count = nullptr; B.allocated = false; B.slots = 0; B.cap_slots = 0;
}

// capacity policy: demand * slack + pad, rounded up to a multiple of 32 slots
long long bucket_cap(opal_ctx* c, long long demand, bool roomy) {
    // small populations (memory is no concern): twice the demand, so the density fluctuations of e.g. a saturated
    // two-stream instability do not force a re-layout (a host round trip) every few steps
    const double slack = roomy ? std::max(c->slack, 2.0) : c->slack;
    const long long pad = roomy ? std::max(c->pad, 128) : c->pad;
    long long cap = (long long)std::ceil((double)demand * slack) + pad;
    return (cap + 31) / 32 * 32;
}

// `spare`: a set of arrays a re-layout left behind (regrid).  When it is large enough the new layout moves in without a single
// cudaMalloc / cudaFree (each a device-wide synchronisation of milliseconds: a moving density pile re-lays its buckets out every
// few hundred steps); otherwise it is freed and the new arrays get headroom for the next time.
bool alloc_buckets(opal_ctx* c, Buckets& B, const std::vector<long long>& demand, bool full, Buckets* spare = nullptr) {
    const int ncell = c->g.nx;
    B.h_start.assign(ncell + 1, 0);
    long long total = 0;
    for (int i = 0; i < ncell; i++) total += demand[i];
    const bool roomy = !c->slack_from_env && total <= 4000000 && (long long)ncell <= 4000000;
    for (int i = 0; i < ncell; i++) B.h_start[i + 1] = B.h_start[i] + bucket_cap(c, demand[i], roomy);
    B.slots = B.h_start[ncell];
    B.full = full;
    const int nf = full ? NF_FULL : NF_LEAN;
    if (spare && !c->pool_buckets) { free_buckets(*spare); spare = nullptr; }
    if (spare && spare->d_start && spare->full == full && spare->cap_slots >= B.slots) {
        // (what the arrays hold is finite -- particles of an earlier layout, or the zeroes they were allocated with -- which is all
        //  the lean push asks of a slot beyond count[])
        B.d_start = spare->d_start; B.d_count = spare->d_count; B.cap_slots = spare->cap_slots;
        for (int k = 0; k < NF_FULL; k++) B.d_f[k] = spare->d_f[k];
        *spare = Buckets();
    } else {
        const long long spare_cap = spare ? spare->cap_slots : 0;
        if (spare) free_buckets(*spare);
        // (small populations: room to spare, so that the layouts to come find the arrays they move into large enough)
        B.cap_slots = std::max<long long>(spare && roomy ? std::max(2 * B.slots, 2 * spare_cap) : B.slots, 1);
        CUB(cudaMalloc(&B.d_start, (size_t)(ncell + 1) * sizeof(long long)));
        CUB(cudaMalloc(&B.d_count, (size_t)ncell * sizeof(int)));
        for (int k = 0; k < nf; k++) {
            CUB(cudaMalloc(&B.d_f[k], (size_t)B.cap_slots * sizeof(double)));
            // zero-filled: the lean push runs its arithmetic on every lane of a chunk, also on slots no particle has ever occupied
            CUB(cudaMemsetAsync(B.d_f[k], 0, (size_t)B.cap_slots * sizeof(double), c->stream));
        }
    }
    CUB(cudaMemcpyAsync(B.d_start, B.h_start.data(), (size_t)(ncell + 1) * sizeof(long long), cudaMemcpyHostToDevice, c->stream));
    CUB(cudaMemsetAsync(B.d_count, 0, (size_t)ncell * sizeof(int), c->stream));
    CUB(cudaStreamSynchronize(c->stream));
    B.allocated = true;
    return true;
}

bool ensure_movers(opal_ctx* c, long long n_particles) {
    // small slabs: room for every particle crossing a cell each step (relativistic plasma at 0.95 CFL does), twice over,
    // because the hashed segments fill unevenly; large slabs: the thermal crossing fraction
    const double frac = env_double("OPAL_MOVER_FRACTION", n_particles <= 4000000 ? 2.0 : 0.05);
    long long seg_cap = (long long)std::ceil(2.0 * frac * (double)n_particles / OPAL_MOVER_SEGMENTS);
    if (seg_cap < 1024) seg_cap = 1024;
    // parked movers (destination bucket full, e.g. a compression front): never more than the species holds; bounded at
    // 2M records (256 MB) for the large slabs, whose per-step pile-up is a small fraction
    long long ov_cap = std::max<long long>(seg_cap, std::min<long long>(n_particles, 1ll << 21));
    if (seg_cap <= c->mover_seg_cap && ov_cap <= c->overflow_cap) return true;
    seg_cap = std::max<long long>(seg_cap, c->mover_seg_cap);
    ov_cap = std::max<long long>(ov_cap, c->overflow_cap);
    if (c->d_movers) cudaFree(c->d_movers);
    if (c->d_overflow) cudaFree(c->d_overflow);
    c->d_movers = nullptr; c->d_overflow = nullptr;
    CUB(cudaMalloc(&c->d_movers, (size_t)seg_cap * OPAL_MOVER_SEGMENTS * sizeof(opal_migrant)));
    CUB(cudaMalloc(&c->d_overflow, (size_t)ov_cap * sizeof(opal_migrant)));
    c->mover_seg_cap = (int)seg_cap;
    c->overflow_cap = (int)ov_cap;
    return true;
}

// Grow the segmented mover buffer to seg_cap records per segment.  Only between advances: its contents are consumed
// by the scatter of the advance that wrote them.
bool reserve_mover_segments(opal_ctx* c, long long seg_cap) {
    if (seg_cap <= c->mover_seg_cap) return true;
    if (seg_cap > (1ll << 30)) return fail(c, OPAL_ECAPACITY, "mover buffer: more than 2^30 records per segment");
    if (c->d_movers) cudaFree(c->d_movers);
    c->d_movers = nullptr; c->mover_seg_cap = 0;
    CUB(cudaMalloc(&c->d_movers, (size_t)seg_cap * OPAL_MOVER_SEGMENTS * sizeof(opal_migrant)));
    c->mover_seg_cap = (int)seg_cap;
    return true;
}

// The first advance after an upload / generate: count on the device how many particles the coming push will move to
// another cell (k_crossing_estimate) and make room for twice that -- a relativistic beam moves nearly every particle
// every step, which the thermal default of ensure_movers (5 % of a large population) does not hold.  Later growth of
// the crossing rate (a plasma heating up) is followed by grow_movers_after.
bool size_movers_for(opal_ctx* c, int sp) {
    Buckets& B = c->bk[sp];
    if (!B.allocated || B.movers_sized) return true;
    B.movers_sized = true;
    if (B.n_hint == 0) return true;
    c->k.crossing_estimate(c->stream, bdev(c, sp), c->g.dt / c->g.dx, c->d_energy + 5); c->launches++;
    double h[2] = {0.0, 0.0};
    CUB(cudaMemcpyAsync(h, c->d_energy + 5, 2 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CUB(cudaStreamSynchronize(c->stream));
    const double need = std::max(h[0], h[1]);
    return reserve_mover_segments(c, (long long)std::ceil(2.0 * need / OPAL_MOVER_SEGMENTS) + 1024);
}

// after an advance whose status has been fetched: a segment more than half full doubles the room for the next one
bool grow_movers_after(opal_ctx* c) {
    const int* seg = reinterpret_cast<const int*>(c->h_status + 1);
    int mx = 0;
    for (int i = 0; i < OPAL_MOVER_SEGMENTS; i++) mx = std::max(mx, seg[i]);
    if (2ll * mx <= c->mover_seg_cap) return true;
    return reserve_mover_segments(c, 2ll * mx + 1024);
}

// per-cell mover strips of a lean species: MC records per cell and direction, sized for the expected crossing fraction
// (thermal plasma at 0.95 CFL: a few % of a cell's particles per step); what does not fit goes to the segmented buffer
void reset_lean(opal_ctx* c, int sp) {
    opal_ctx::Lean& L = c->lean[sp];
    L.parity = 0; L.pending = false; L.holes = false;
    if (!L.d_hcount) return;
    for (int k = 0; k < 2; k++) cudaMemsetAsync(L.d_scount[k], 0, (size_t)2 * c->g.nx * sizeof(int), c->stream);
    cudaMemsetAsync(L.d_hcount, 0, (size_t)c->g.nx * sizeof(int), c->stream);
}

bool ensure_cell_movers(opal_ctx* c, int sp, long long n_particles) {
    if (c->full || (c->flags & OPAL_FLAG_STRICT_FP)) return true;
    opal_ctx::Lean& L = c->lean[sp];
    const double ppc = (double)n_particles / (double)c->g.nx;
    long long mc = (long long)std::ceil((env_double("OPAL_CELL_MOVER_FRACTION", 0.02) * ppc + 12.0) / 16.0) * 16;   // per direction
    if (mc < 16) mc = 16;
    if (mc > 256) mc = 256;
    mc = std::max<long long>(mc, L.mc_floor);
    if (!L.d_hcount || mc > L.mc) {
        for (int k = 0; k < 2; k++) {
            if (L.d_strips[k]) cudaFree(L.d_strips[k]);
            if (L.d_scount[k]) cudaFree(L.d_scount[k]);
            L.d_strips[k] = nullptr; L.d_scount[k] = nullptr;
        }
        if (L.d_hcount) cudaFree(L.d_hcount);
        L.d_hcount = nullptr;
        for (int k = 0; k < 2; k++) {
            CUB(cudaMalloc(&L.d_strips[k], (size_t)mc * 2 * c->g.nx * sizeof(StripRec)));
            CUB(cudaMalloc(&L.d_scount[k], (size_t)2 * c->g.nx * sizeof(int)));
        }
        CUB(cudaMalloc(&L.d_hcount, (size_t)c->g.nx * sizeof(int)));
        L.mc = (int)mc;
    }
    reset_lean(c, sp);
    return true;
}

// Strips too narrow for this plasma?  (A relativistically hot cell sends half its particles next door every step; what does not
// fit its strip goes through the segmented buffer and k_scatter: global atomics per record.)  Small grids can afford wide strips.
int strip_width_cap(const opal_ctx* c) {
    const long long by_memory = (long long)(env_double("OPAL_STRIP_BYTES", 2.0e9) / (2.0 * sizeof(StripRec) * (double)c->g.nx)) / 16 * 16;    // ~2 GB per strip buffer
    return (int)std::min<long long>(by_memory, 2048);      // (a large grid stays with what ensure_cell_movers gave it)
}
bool strips_too_narrow(const opal_ctx* c, int sp) {
    const opal_ctx::Lean& L = c->lean[sp];
    if (!L.d_hcount || L.spill_adv < 8 || L.mc >= strip_width_cap(c) || !c->widen_strips) return false;
    const double per_advance = (double)L.spill / (double)L.spill_adv;
    return per_advance > std::max(64.0, 0.002 * (double)c->bk[sp].n_hint);
}
bool settle(opal_ctx* c, int sp);
bool maybe_widen_strips(opal_ctx* c) {
    for (int sp = 0; sp < 2; sp++) {
        opal_ctx::Lean& L = c->lean[sp];
        if (!strips_too_narrow(c, sp)) { if (L.spill_adv >= 64) { L.spill = 0; L.spill_adv = 0; } continue; }
        if (!settle(c, sp)) return false;                   // the last push's movers into the buckets, free slots closed
        L.mc_floor = std::max(L.mc, std::min(strip_width_cap(c), std::max(4 * L.mc, 64)));
        L.spill = 0; L.spill_adv = 0;
        c->n_widened++;
        if (!ensure_cell_movers(c, sp, c->bk[sp].n_hint)) return false;
    }
    return true;
}

bool ensure_photons(opal_ctx* c, long long need) {
    if (need <= c->ph[0].cap) return true;
    long long cap = std::max<long long>(need * 2, 1 << 16);
    for (int w = 0; w < 2; w++) {
        Photons old = c->ph[w];
        Photons& P = c->ph[w];
        P = Photons();
        P.cap = (int)cap;
        CUB(cudaMalloc(&P.d_n, sizeof(int)));
        CUB(cudaMemsetAsync(P.d_n, 0, sizeof(int), c->stream));
        CUB(cudaMalloc(&P.d_cell, (size_t)cap * sizeof(int)));
        for (int k = 0; k < NF_PHOTON; k++) CUB(cudaMalloc(&P.d_f[k], (size_t)cap * sizeof(double)));
        if (old.cap > 0) {
            if (w == c->ph_cur) {
                CUB(cudaMemcpyAsync(P.d_n, old.d_n, sizeof(int), cudaMemcpyDeviceToDevice, c->stream));
                CUB(cudaMemcpyAsync(P.d_cell, old.d_cell, (size_t)old.cap * sizeof(int), cudaMemcpyDeviceToDevice, c->stream));
                for (int k = 0; k < NF_PHOTON; k++)
                    CUB(cudaMemcpyAsync(P.d_f[k], old.d_f[k], (size_t)old.cap * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
            }
            CUB(cudaStreamSynchronize(c->stream));
            cudaFree(old.d_n); cudaFree(old.d_cell);
            for (int k = 0; k < NF_PHOTON; k++) cudaFree(old.d_f[k]);
        }
    }
    return true;
}

bool photon_count(opal_ctx* c, int* n) {
    *n = 0;
    if (c->ph[0].cap == 0) return true;
    CUB(cudaMemcpyAsync(n, c->ph[c->ph_cur].d_n, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CUB(cudaStreamSynchronize(c->stream));
    return true;
}

bool reset_status(opal_ctx* c, bool keep_err) {
    // zero the counters (and the per-segment mover counters); the sticky error word survives
    if (keep_err) {
        CUB(cudaMemsetAsync(c->d_status, 0, offsetof(StatusDev, err), c->stream));
        CUB(cudaMemsetAsync((char*)c->d_status + offsetof(StatusDev, n_new_photons), 0,
                            sizeof(StatusDev) - offsetof(StatusDev, n_new_photons) + OPAL_MOVER_SEGMENTS * sizeof(int), c->stream));
    } else {
        CUB(cudaMemsetAsync(c->d_status, 0, sizeof(StatusDev) + OPAL_MOVER_SEGMENTS * sizeof(int), c->stream));
    }
    return true;
}

// copy the status block to the host, translate sticky device-side assertion bits
// what the status block just read says about heavy cells; returns true when the block-per-cell launch was just switched on
bool note_heavy(opal_ctx* c) {
    if (c->h_status->heavy_seen) {
        c->heavy_quiet = 0;
        if (!c->heavy_on && c->heavy_min > 0) { c->heavy_on = true; return true; }
    } else if (c->heavy_on && !c->heavy_forced && ++c->heavy_quiet >= 64) c->heavy_on = false;
    return false;
}

bool fetch_status(opal_ctx* c) {
    // the per-segment mover counts travel with it (4 KB): advance_charged sizes the mover buffer from them
    CUB(cudaMemcpyAsync(c->h_status, c->d_status, sizeof(StatusDev) + OPAL_MOVER_SEGMENTS * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CUB(cudaStreamSynchronize(c->stream));
    note_heavy(c);
    const unsigned e = c->h_status->err;
    if (e & ERR_CFL) return fail(c, OPAL_EASSERT_CFL, "push: assert!(dxi < 1.0) failed (electron.rs:313 / ion.rs:199 / photon.rs:164)");
    if (e & ERR_INDEX) return fail(c, OPAL_EASSERT_INDEX, "cell index out of range (yee.rs:501-502)");
    if (e & ERR_CHI) return fail(c, OPAL_ECHI_RANGE, "sample: chi beyond the tabulated range (photon_emission.rs:144)");
    if (e & ERR_PROB) return fail(c, OPAL_EASSERT_PROB, "absorb: assert!(partial_prob >= 0.0) failed (photon.rs:325 / photon.rs:336)");
    if (e & ERR_CAPACITY) {
        std::string which;
        if (e & ERR_CAP_MOVERS) which += " mover segments (raise OPAL_MOVER_FRACTION)";
        if (e & ERR_CAP_PARKED) which += " parked-mover list";
        if (e & ERR_CAP_OUTBOX) which += " neighbour outbox (raise OPAL_EXCHANGE_CAP)";
        if (e & ERR_CAP_PHOTONS) which += " photon store";
        if (which.empty()) which = " mover strip / photon append (raise OPAL_MOVER_FRACTION / OPAL_EXCHANGE_CAP)";
        return fail(c, OPAL_ECAPACITY, "fixed-capacity buffer overflow:" + which);
    }
    return true;
}

bool handle_overflow(opal_ctx* c, int sp, bool deposit_parked);

// Between lean steps the last push's movers wait in the strips and the buckets carry free slots (opal_push_lean.cuh).
// Everything but the step kernels reads buckets whose first count[c] slots are live and complete: append the waiting
// movers (k_strip_flush; a bucket may overflow -> re-layout), then close the free slots (k_compact_cells).
bool flush_strips(opal_ctx* c, int sp, bool inside_overflow_handling = false) {
    opal_ctx::Lean& L = c->lean[sp];
    if (!L.pending) return true;
    L.pending = false;
    if (!c->bk[sp].allocated) return true;
    // (inside the overflow handling the parked-mover list is in use: what this flush parks is appended to it)
    if (!inside_overflow_handling) CUB(cudaMemsetAsync(&c->d_status->n_overflow, 0, sizeof(int), c->stream));
    c->k.flush(c->stream, bdev(c, sp), sparams(c, sp), L.d_strips[L.parity], L.d_scount[L.parity], c->d_overflow, c->d_status); c->launches++;
    CUB(cudaMemsetAsync(L.d_scount[L.parity], 0, (size_t)2 * c->g.nx * sizeof(int), c->stream));
    if (!fetch_status(c)) return false;
    return inside_overflow_handling ? true : handle_overflow(c, sp, true);
}

bool settle(opal_ctx* c, int sp) {
    if (sp != OPAL_ELECTRON && sp != OPAL_ION) return true;
    opal_ctx::Lean& L = c->lean[sp];
    if (!flush_strips(c, sp)) return false;
    if (L.holes) {
        L.holes = false;
        if (c->bk[sp].allocated && L.d_hcount) { c->k.compact_cells(c->stream, bdev(c, sp), L.d_hcount); c->launches++; }
    }
    return true;
}

// Re-lay the buckets of species `sp` out for the demand recorded in count[] (which includes the
// movers parked in d_overflow), then scatter the parked movers.
bool regrid(opal_ctx* c, int sp, bool deposit_parked) {
    c->n_regrids++;
    WallTimer wt(c->t_regrid);
    double tq = wall_now();
    auto lap = [&](int k) { const double t = wall_now(); c->t_rg[k] += t - tq; tq = t; };
    Buckets& B = c->bk[sp];
    const int ncell = c->g.nx;
    std::vector<int> cnt(ncell);
    CUB(cudaMemcpyAsync(cnt.data(), B.d_count, (size_t)ncell * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CUB(cudaStreamSynchronize(c->stream));
    lap(0);
    std::vector<long long> demand(ncell);
    long long total = 0;
    for (int i = 0; i < ncell; i++) { demand[i] = cnt[i]; total += cnt[i]; }
    // A density pile that MOVES (a compression front, an expanding edge) overflows a fresh set of cells every few steps.
    // When re-layouts come in quick succession, size every bucket for the densest cell within a growing radius and add
    // slack, so the next re-layout is only due after the pile has travelled that far.
    if (c->adv_calls[sp] - c->last_regrid[sp] < 64) {
        // (small populations can afford a wide neighbourhood: hole_boring halts 69 times instead of 125 with 128 cells)
        const int radius_max = (int)env_ll("OPAL_REGRID_RADIUS_MAX", total <= 4000000 && !c->slack_from_env ? 128 : 32);
        c->regrid_radius[sp] = std::min(std::max(2 * c->regrid_radius[sp], 2), radius_max);
        c->regrid_boost[sp] = std::min(c->regrid_boost[sp] * 1.25, 3.0);
    } else {
        c->regrid_radius[sp] /= 2;
        c->regrid_boost[sp] = std::max(1.0, c->regrid_boost[sp] / 1.25);
    }
    c->last_regrid[sp] = c->adv_calls[sp];
    if (c->regrid_radius[sp] > 0 || c->regrid_boost[sp] > 1.0) {
        const int R = c->regrid_radius[sp];
        std::vector<long long> wide(ncell);
        for (int i = 0; i < ncell; i++) {
            long long m = 0;
            for (int j = std::max(0, i - R); j <= std::min(ncell - 1, i + R); j++) m = std::max<long long>(m, cnt[j]);
            wide[i] = (long long)std::ceil((double)m * c->regrid_boost[sp]);
        }
        // the old and the new buckets coexist during the copy: widen only when that fits comfortably
        long long wide_slots = 0;
        for (int i = 0; i < ncell; i++) wide_slots += bucket_cap(c, wide[i], false);
        const double need = (double)wide_slots * (B.full ? NF_FULL : NF_LEAN) * sizeof(double);
        size_t free_b = 0, total_b = 0;
        if (need > 1.0e9) cudaMemGetInfo(&free_b, &total_b);           // (a millisecond: only asked when it can matter)
        if (need <= 1.0e9 || need < 0.8 * (double)free_b) demand.swap(wide);
        else { c->regrid_radius[sp] = 0; c->regrid_boost[sp] = 1.0; }
    }
    lap(1);
    Buckets old = B;
    B = Buckets();
    if (!alloc_buckets(c, B, demand, old.full, &c->spare[sp])) return false;
    lap(2);
    B.n_hint = total;
    B.movers_sized = old.movers_sized;
    BucketDev src; src.ncell = ncell; src.start = old.d_start; src.count = old.d_count;
    for (int k = 0; k < NF_FULL; k++) src.f[k] = old.d_f[k];
    src.z_star = c->z_star; src.mass_number = c->mass_number;
    c->k.copy_buckets(c->stream, src, bdev(c, sp), old.full); c->launches++;
    // parked movers: their cells are final
    const int n_parked = c->h_status->n_overflow;
    CUB(cudaMemcpyAsync(c->d_tmp_int, &c->d_status->n_overflow, sizeof(int), cudaMemcpyDeviceToDevice, c->stream));
    CUB(cudaMemsetAsync(&c->d_status->n_overflow, 0, sizeof(int), c->stream));
    StepParams sp_ = sparams(c, sp);
    sp_.route_left = sp_.route_right = 0;
    BucketDev dst_b = bdev(c, sp);
    if (dst_b.eq_n) {            // the slots just moved: the emitter list of this step is void (k_emit scans instead)
        CUB(cudaMemsetAsync(dst_b.eq_n, 0xFF, sizeof(int), c->stream));
        dst_b.eq_slot = nullptr; dst_b.eq_cell = nullptr; dst_b.eq_n = nullptr; dst_b.eq_cap = 0;
    }
    c->k.scatter(c->stream, c->g, dst_b, sp_, old.full, deposit_parked, c->d_overflow, c->d_tmp_int, c->overflow_cap, 0, nullptr,
                 nullptr, nullptr, reinterpret_cast<int*>(c->d_overflow), c->d_status); c->launches++;
    (void)n_parked;
    CUB(cudaStreamSynchronize(c->stream));
    lap(3);
    // small populations keep the old arrays for the next re-layout to move into; large ones give the memory back
    if ((double)old.cap_slots * (old.full ? NF_FULL : NF_LEAN) * sizeof(double) <= 2.0e9) {
        c->spare[sp] = old;                   // (its arrays only: alloc_buckets looks at nothing else)
        c->spare[sp].allocated = false;
    } else free_buckets(old);
    lap(4);
    return fetch_status(c);
}

bool handle_overflow(opal_ctx* c, int sp, bool deposit_parked) {
    int guard = 0;
    // lean layout: the buckets are re-laid out for the demand in count[] -- bring the movers waiting in the strips in first
    if (c->h_status->n_overflow > 0 && !flush_strips(c, sp, true)) return false;
    while (c->h_status->n_overflow > 0) {
        if (c->h_status->n_overflow > c->overflow_cap) return fail(c, OPAL_ECAPACITY, "too many parked movers");
        if (!regrid(c, sp, deposit_parked)) return false;
        if (++guard > 4) return fail(c, OPAL_ECAPACITY, "re-layout did not converge");
    }
    return true;
}

// write the record counts into the header records of both outboxes
// (charged species: also the populations of the slab's two edge cells, the bound on the next step's leavers)
__global__ void k_outbox_headers(opal_migrant* out_l, opal_migrant* out_r, StatusDev* st, const int* count, int ncell) {
    if (threadIdx.x == 0) {
        out_l[0].cell = st->n_out[0]; out_r[0].cell = st->n_out[1];
        if (count) { st->edge[0] = count[0]; st->edge[1] = count[ncell - 1]; }
    }
}
__global__ void k_header_to_int(const opal_migrant* in, int* n) { if (threadIdx.x == 0) *n = (int)in[0].cell; }
// burst mode: empty the destination of the photon ping-pong -- unless the burst has halted (a memset node would wipe the live store)
__global__ void k_burst_zero_int(const StatusDev* st, int* p) {
    if (threadIdx.x == 0 && !*(volatile const int*)&st->halt) *p = 0;
}
// live particles of a species: sum of count[] (out must be zeroed)
// first and last non-empty cell of a species (out[0] = min index, out[1] = max index; preset to INT_MAX / -1)
__global__ void __launch_bounds__(256) k_occupied_range(const int* __restrict__ count, int ncell, int* out) {
    int lo = INT_MAX, hi = -1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ncell; i += gridDim.x * blockDim.x)
        if (count[i] > 0) { lo = min(lo, i); hi = max(hi, i); }
    for (int o = 16; o > 0; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
    if ((threadIdx.x & 31) == 0 && hi >= 0) { atomicMin(out, lo); atomicMax(out + 1, hi); }
}

__global__ void __launch_bounds__(256) k_total_count(const int* __restrict__ count, int ncell, unsigned long long* out) {
    unsigned long long acc = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ncell; i += gridDim.x * blockDim.x) acc += (unsigned long long)count[i];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}

// NCCL ring exchange of both outboxes (fixed size: header + out_cap records) -- replaces the
// even/odd blocking p2p of mod.rs:308-354 with one grouped, deadlock-free exchange.
bool nccl_exchange(opal_ctx* c, void* send_l, void* send_r, void* recv_l, void* recv_r, size_t bytes) {
    if (!c->comm) return fail(c, OPAL_ENCCL, "remote neighbour but no communicator (call opal_comm_init) -- or use the host-staged take/give calls");
    int rc = 0;
    rc |= g_nccl.GroupStart();
    // order matters when both neighbours are the same rank (nranks == 2): my send-to-right pairs
    // with the peer's recv-from-left
    if (c->has_right()) rc |= g_nccl.Send(send_r, bytes, NCCL_CHAR, c->right_rank(), c->comm, c->stream);
    if (c->has_left()) rc |= g_nccl.Send(send_l, bytes, NCCL_CHAR, c->left_rank(), c->comm, c->stream);
    if (c->has_left()) rc |= g_nccl.Recv(recv_l, bytes, NCCL_CHAR, c->left_rank(), c->comm, c->stream);
    if (c->has_right()) rc |= g_nccl.Recv(recv_r, bytes, NCCL_CHAR, c->right_rank(), c->comm, c->stream);
    rc |= g_nccl.GroupEnd();
    if (rc) return fail(c, OPAL_ENCCL, "ncclSend/ncclRecv failed");
    return true;
}

struct Timed {   // RAII bracket of one launch with a cudaEvent pair when timing is on
    opal_ctx* c; int cat; cudaEvent_t stop = nullptr;
    Timed(opal_ctx* c_, int cat_) : c(c_), cat(cat_) {
        if (!c->timing) return;
        if (c->ev_used[cat] == c->ev[cat].size()) {
            cudaEvent_t a, b;
            cudaEventCreate(&a); cudaEventCreate(&b);
            c->ev[cat].push_back({a, b});
        }
        auto& p = c->ev[cat][c->ev_used[cat]++];
        cudaEventRecord(p.first, c->stream);
        stop = p.second;
    }
    ~Timed() { if (stop) cudaEventRecord(stop, c->stream); }
};

bool remote(opal_ctx* c) { return !c->self_nbr() && (c->has_left() || c->has_right()); }

// The outboxes must hold every leaver of the coming push: a particle moves less than a cell per step (assert dxi < 1,
// electron.rs:313), so the leavers through an edge are at most the particles sitting in that edge cell now.
bool ensure_outbox(opal_ctx* c, int sp) {
    if (!remote(c)) return true;
    long long need = 0;
    for (int side = 0; side < 2; side++) {
        if (c->edge_bound[sp][side] < 0) {             // unknown (fresh upload / first step): read it
            if (sp == OPAL_PHOTON) {
                int n = 0;
                if (!photon_count(c, &n)) return false;
                c->edge_bound[sp][side] = n;
            } else if (c->bk[sp].allocated) {
                int v = 0;
                CUB(cudaMemcpyAsync(&v, c->bk[sp].d_count + (side == 0 ? 0 : c->g.nx - 1), sizeof(int), cudaMemcpyDeviceToHost, c->stream));
                CUB(cudaStreamSynchronize(c->stream));
                c->edge_bound[sp][side] = v;
            } else c->edge_bound[sp][side] = 0;
        }
        need = std::max(need, c->edge_bound[sp][side]);
    }
    need += 64;
    if (need <= c->out_cap) return true;
    if (need > (1ll << 24)) return fail(c, OPAL_ECAPACITY, "neighbour outbox: more than 2^24 particles in an edge cell");
    const long long cap = std::max<long long>(need + need / 4, 2ll * c->out_cap);
    for (int s_ = 0; s_ < 2; s_++) {
        if (c->d_out[s_]) cudaFree(c->d_out[s_]);
        c->d_out[s_] = nullptr;
        CUB(cudaMalloc(&c->d_out[s_], (size_t)(cap + 1) * sizeof(opal_migrant)));
    }
    c->out_cap = (int)cap;
    return true;
}

bool insert_inbox(opal_ctx* c, int sp, int side, bool dep);

bool ensure_inbox(opal_ctx* c, long long need) {
    if (need <= c->in_cap) return true;
    if (need > (1ll << 24)) return fail(c, OPAL_ECAPACITY, "neighbour inbox: more than 2^24 particles announced");
    const long long cap = std::max<long long>(need + need / 4, 2ll * c->in_cap);
    for (int side = 0; side < 2; side++) {
        opal_migrant* fresh = nullptr;
        CUB(cudaMalloc(&fresh, (size_t)(cap + 1) * sizeof(opal_migrant)));
        CUB(cudaMemcpyAsync(fresh, c->d_in[side], sizeof(opal_migrant), cudaMemcpyDeviceToDevice, c->stream));    // keep the header
        CUB(cudaStreamSynchronize(c->stream));
        cudaFree(c->d_in[side]);
        c->d_in[side] = fresh;
    }
    c->in_cap = (int)cap;
    return true;
}

// Neighbour exchange of the leavers of species `sp` (mod.rs:290-373) over NCCL, in two phases: the counts first (one
// 128-byte header each way), then exactly that many records -- the reference sends unbounded Vecs; nothing here is
// capped: the sender sized its outbox before the push (ensure_outbox), the receiver grows its inbox to what is announced.
bool exchange_migrants(opal_ctx* c, int sp, bool dep) {
    k_outbox_headers<<<1, 32, 0, c->stream>>>(c->d_out[0], c->d_out[1], c->d_status, sp == OPAL_PHOTON ? nullptr : c->bk[sp].d_count, c->g.nx);
    c->launches++;
    if (!c->comm) {                                   // host-staged exchange: opal_migrants_take / give
        c->edge_bound[sp][0] = c->edge_bound[sp][1] = -1;          // (re-read before the next push: the gives come later)
        return true;
    }
    if (!nccl_exchange(c, c->d_out[0], c->d_out[1], c->d_in[0], c->d_in[1], sizeof(opal_migrant))) return false;
    opal_migrant hdr[2];
    memset(hdr, 0, sizeof(hdr));
    for (int side = 0; side < 2; side++)
        CUB(cudaMemcpyAsync(&hdr[side], c->d_in[side], sizeof(opal_migrant), cudaMemcpyDeviceToHost, c->stream));
    if (!fetch_status(c)) return false;               // synchronises: the headers have arrived too
    const long long n_out[2] = {c->has_left() ? c->h_status->n_out[0] : 0, c->has_right() ? c->h_status->n_out[1] : 0};
    const long long n_in[2] = {c->has_left() ? hdr[0].cell : 0, c->has_right() ? hdr[1].cell : 0};
    if (!ensure_inbox(c, std::max(n_in[0], n_in[1]))) return false;
    if (n_out[0] + n_out[1] + n_in[0] + n_in[1] > 0) {
        int rc = g_nccl.GroupStart();
        // (both neighbours may be the same rank: my send-to-right pairs with its recv-from-left -- same order on both ends)
        if (n_out[1]) rc |= g_nccl.Send(c->d_out[1] + 1, (size_t)n_out[1] * sizeof(opal_migrant), NCCL_CHAR, c->right_rank(), c->comm, c->stream);
        if (n_out[0]) rc |= g_nccl.Send(c->d_out[0] + 1, (size_t)n_out[0] * sizeof(opal_migrant), NCCL_CHAR, c->left_rank(), c->comm, c->stream);
        if (n_in[0]) rc |= g_nccl.Recv(c->d_in[0] + 1, (size_t)n_in[0] * sizeof(opal_migrant), NCCL_CHAR, c->left_rank(), c->comm, c->stream);
        if (n_in[1]) rc |= g_nccl.Recv(c->d_in[1] + 1, (size_t)n_in[1] * sizeof(opal_migrant), NCCL_CHAR, c->right_rank(), c->comm, c->stream);
        rc |= g_nccl.GroupEnd();
        if (rc) return fail(c, OPAL_ENCCL, "ncclSend/ncclRecv failed");
    }
    if (n_in[0]) insert_inbox(c, sp, OPAL_LEFT, dep);
    if (n_in[1]) insert_inbox(c, sp, OPAL_RIGHT, dep);
    // what can leave in the next step: who sits in the edge cells now, plus who just arrived there
    for (int side = 0; side < 2; side++) c->edge_bound[sp][side] = (long long)c->h_status->edge[side] + n_in[side];
    if (sp == OPAL_PHOTON)       // ... plus one photon per electron of the edge cell (emit_radiation comes after the photons' advance)
        for (int side = 0; side < 2; side++) {
            c->edge_bound[sp][side] += std::max<long long>(c->edge_bound[OPAL_ELECTRON][side], 0);
            if (c->absorb_on) c->edge_bound[sp][side] *= 2;       // ... and a stimulated copy of each photon there (photon.rs:312-362)
        }
    return true;
}


// insert what the neighbours sent (d_in[side], header + records) into species `sp`
bool insert_inbox(opal_ctx* c, int sp, int side, bool dep) {
    k_header_to_int<<<1, 32, 0, c->stream>>>(c->d_in[side], c->d_tmp_int + 1 + side); c->launches++;
    const long long shift = side == OPAL_RIGHT ? c->g.nx : 0;   // mod.rs:366-368
    if (sp == OPAL_PHOTON) {
        c->k.photon_insert(c->stream, pdev(c, c->ph_cur), c->d_in[side] + 1, c->d_tmp_int + 1 + side, c->in_cap, shift, c->d_status);
    } else {
        StepParams sp_ = sparams(c, sp);
        sp_.route_left = sp_.route_right = 0;
        c->k.scatter(c->stream, c->g, bdev(c, sp), sp_, c->bk[sp].full, dep, c->d_in[side] + 1, c->d_tmp_int + 1 + side, c->in_cap, shift,
                     nullptr, nullptr, nullptr, reinterpret_cast<int*>(c->d_overflow), c->d_status);
    }
    c->launches++;
    return true;
}

// gather + push + re-bucket + migrate one charged species (mod.rs:247-376); dep: fused deposit
bool advance_charged(opal_ctx* c, int sp, bool dep, bool check, bool want_ke = false) {
    Buckets& B = c->bk[sp];
    if (!B.allocated) { c->last_gone[sp][0] = c->last_gone[sp][1] = 0; return true; }
    if (check && !size_movers_for(c, sp)) return false;      // (the unchecked callers -- graph capture -- size up front)
    if (!ensure_outbox(c, sp)) return false;
    if (!reset_status(c, true)) return false;
    c->adv_calls[sp]++;
    StepParams p = sparams(c, sp);
    bool strips = false;
    double* ke_out = nullptr;
    if (want_ke && sp == OPAL_ELECTRON) {
        ke_out = c->d_energy + 4;
        CUB(cudaMemsetAsync(ke_out, 0, sizeof(double), c->stream));
    }
    opal_ctx::Lean& L = c->lean[sp];
    if (sp == OPAL_ELECTRON && c->eq_on && c->d_eq_n) CUB(cudaMemsetAsync(c->d_eq_n, 0, sizeof(int), c->stream));   // this step's emitter list starts empty
    if (!dep && !flush_strips(c, sp)) return false;          // the call-by-call push does not read the strips
    LeanStrips ls = {L.d_strips[L.parity], L.d_scount[L.parity], L.d_strips[1 - L.parity], L.d_scount[1 - L.parity], L.d_hcount};
    { Timed tm(c, 0); strips = c->k.push(c->stream, c->g, bdev(c, sp), p, B.full, dep, c->d_movers, L.d_hcount ? &ls : nullptr, c->d_out[0] + 1, c->d_out[1] + 1,
                                         c->d_overflow, c->d_status, ke_out); c->launches++; }
    if (strips && p.heavy_min < INT_MAX) { c->launches++; c->n_heavy_launches++; }
    if (ke_out && strips) c->ke_fused = true;
    if (strips) { L.parity ^= 1; L.pending = true; L.holes = true; }      // this push's movers wait for the next one
    { Timed tm(c, 1); c->k.scatter(c->stream, c->g, bdev(c, sp), p, B.full, dep, c->d_movers, nullptr, 0, 0, nullptr, c->d_out[0] + 1, c->d_out[1] + 1,
                 reinterpret_cast<int*>(c->d_overflow), c->d_status); c->launches++; }
    if (remote(c) && !exchange_migrants(c, sp, dep)) return false;
    if (check) {
        if (!fetch_status(c)) return false;
        c->last_gone[sp][0] = c->h_status->gone[0]; c->last_gone[sp][1] = c->h_status->gone[1];
        if (strips) { L.spill += c->h_status->strip_spill[sp]; L.spill_adv++; }
        if (!handle_overflow(c, sp, dep)) return false;
        if (!grow_movers_after(c)) return false;
    }
    return true;
}

// ---- Breit-Wheeler extension: room for the pairs of one photon advance, and their insertion into the charged species ----
bool prepare_pairs(opal_ctx* c) {
    if (c->ion_slot_is_user) return fail(c, OPAL_EINVAL, "OPAL_FLAG_PAIR_CREATION keeps the positrons in the ion slot: a run with ions cannot use it");
    if (!c->full) return fail(c, OPAL_EINVAL, "pair creation needs the full particle state");
    c->z_star = 1.0; c->mass_number = 9.109383e-31 / 1.672622e-27;          // a positron: charge +e, mass m_e (constants.rs)
    for (int sp = 0; sp < 2; sp++) {
        if (c->bk[sp].allocated) continue;
        std::vector<long long> demand(c->g.nx, 0);
        if (!alloc_buckets(c, c->bk[sp], demand, true)) return false;
    }
    const int want = std::max(1 << 16, c->ph[0].cap / 8);                     // a photon that finds no room converts a step later
    if (want > c->pair_cap) {
        for (int k = 0; k < 2; k++) { if (c->d_pair[k]) cudaFree(c->d_pair[k]); c->d_pair[k] = nullptr; CUB(cudaMalloc(&c->d_pair[k], (size_t)want * sizeof(opal_migrant))); }
        c->pair_cap = want;
    }
    if (!c->d_pair_n) CUB(cudaMalloc(&c->d_pair_n, sizeof(int)));
    return true;
}

bool insert_pairs(opal_ctx* c) {
    int n = 0;
    CUB(cudaMemcpyAsync(&n, c->d_pair_n, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CUB(cudaStreamSynchronize(c->stream));
    n = std::min(n, c->pair_cap);
    if (n == 0) return true;
    c->pairs_created += n;
    CUB(cudaMemcpyAsync(c->d_pair_n, &n, sizeof(int), cudaMemcpyHostToDevice, c->stream));      // (clamped: what the scatter reads)
    for (int sp = 0; sp < 2; sp++) {
        CUB(cudaMemsetAsync(&c->d_status->n_overflow, 0, sizeof(int), c->stream));
        StepParams sp_ = sparams(c, sp);
        sp_.route_left = sp_.route_right = 0;
        c->k.scatter(c->stream, c->g, bdev(c, sp), sp_, true, false, c->d_pair[sp], c->d_pair_n, c->pair_cap, 0, nullptr, nullptr, nullptr,
                     reinterpret_cast<int*>(c->d_overflow), c->d_status);
        c->launches++;
        if (!fetch_status(c)) return false;
        if (!handle_overflow(c, sp, false)) return false;
        c->bk[sp].n_hint += n;
        c->edge_bound[sp][0] = c->edge_bound[sp][1] = -1;
    }
    return true;
}

bool advance_photons(opal_ctx* c, bool check) {
    if (c->flags & OPAL_FLAG_IMMOBILE_PHOTONS) return true;
    if (c->ph[0].cap == 0) { c->last_gone[2][0] = c->last_gone[2][1] = 0; return true; }
    if (!ensure_outbox(c, OPAL_PHOTON)) return false;
    if (!reset_status(c, true)) return false;
    const int src = c->ph_cur, dst = 1 - src;
    CUB(cudaMemsetAsync(c->ph[dst].d_n, 0, sizeof(int), c->stream));
    PairDev pairs;
    if (c->flags & OPAL_FLAG_PAIR_CREATION) {
        if (!prepare_pairs(c)) return false;
        pairs.e_recs = c->d_pair[0]; pairs.p_recs = c->d_pair[1]; pairs.n = c->d_pair_n; pairs.cap = c->pair_cap;
        pairs.seed = c->seed; pairs.call = c->pair_calls++;
        CUB(cudaMemsetAsync(c->d_pair_n, 0, sizeof(int), c->stream));
    }
    c->k.photon_push(c->stream, c->g, pdev(c, src), pdev(c, dst), sparams(c, OPAL_PHOTON), c->d_out[0] + 1, c->d_out[1] + 1, c->d_status, pairs);
    c->launches++;
    c->ph_cur = dst;
    if (pairs.n && !insert_pairs(c)) return false;
    if (remote(c) && !exchange_migrants(c, OPAL_PHOTON, false)) return false;
    if (check) {
        if (!fetch_status(c)) return false;
        c->last_gone[2][0] = c->h_status->gone[0]; c->last_gone[2][1] = c->h_status->gone[1];
    }
    return true;
}

bool do_emit(opal_ctx* c, double t, double min_e, double max_a, double max_fl) {
    Buckets& E = c->bk[OPAL_ELECTRON];
    // multi-slab: a slab that emits nothing itself still has to take part in the photon exchange of the next advance
    if (remote(c) && !ensure_photons(c, 1)) return false;
    if (!E.allocated) return true;
    if (!E.full) return fail(c, OPAL_EINVAL, "emit_radiation needs the full particle state (context created with OPAL_FLAG_LEAN_STATE)");
    // at most one photon per electron per call (electron.rs:208-251): make room up front for the electrons this slab holds
    // NOW (migrants included -- a slab that started empty may have filled up), counted on the device
    unsigned long long* d_live = reinterpret_cast<unsigned long long*>(c->d_energy + 7);
    CUB(cudaMemsetAsync(d_live, 0, sizeof(unsigned long long), c->stream));
    { int gr = (c->g.nx + 255) / 256; if (gr > 148 * 4) gr = 148 * 4; k_total_count<<<gr, 256, 0, c->stream>>>(E.d_count, c->g.nx, d_live); c->launches++; }
    unsigned long long live = 0;
    CUB(cudaMemcpyAsync(&live, d_live, sizeof(live), cudaMemcpyDeviceToHost, c->stream));
    int nph = 0;
    if (!photon_count(c, &nph)) return false;      // synchronises the stream: `live` has arrived too
    E.n_hint = (long long)live;
    if (!ensure_photons(c, (long long)nph + (long long)live + c->out_cap)) return false;
    c->k.emit(c->stream, c->g, bdev(c, OPAL_ELECTRON), pdev(c, c->ph_cur), sparams(c, OPAL_ELECTRON), c->d_qed_tables, t, min_e, max_a, max_fl,
              c->seed, c->emit_calls++, c->d_status, 0, 0);
    c->launches++;
    CUB(cudaGetLastError());
    return fetch_status(c);
}

// absorb (interactions.rs:145): see opal_absorb.cu
bool do_absorb(opal_ctx* c, double t, double max_displacement, double stop_time) {
    Buckets& E = c->bk[OPAL_ELECTRON];
    if (!E.allocated || c->ph[0].cap == 0) return true;
    if (!E.full) return fail(c, OPAL_EINVAL, "absorb needs the full particle state (context created with OPAL_FLAG_LEAN_STATE)");
    int nph = 0;
    if (!photon_count(c, &nph)) return false;
    if (nph == 0) return true;
    // every photon can stimulate at most one copy of itself per call
    if (!ensure_photons(c, 2ll * nph)) return false;
    if (c->absorb_cap < nph) {
        if (c->d_absorb_kind) { cudaFree(c->d_absorb_kind); cudaFree(c->d_absorb_slot); }
        c->absorb_cap = std::max(2 * nph, 1 << 12);
        CUB(cudaMalloc(&c->d_absorb_kind, (size_t)c->absorb_cap * sizeof(int)));
        CUB(cudaMalloc(&c->d_absorb_slot, (size_t)c->absorb_cap * sizeof(long long)));
    }
    if (!c->d_absorb_last) CUB(cudaMalloc(&c->d_absorb_last, sizeof(int)));
    const int src = c->ph_cur, dst = 1 - src;
    c->launches += absorb_launch(c->stream, c->g, bdev(c, OPAL_ELECTRON), pdev(c, src), pdev(c, dst), nph, t, max_displacement, stop_time, c->seed,
                                 (unsigned)c->absorb_calls++, (c->flags & OPAL_FLAG_NO_STIMULATED_EMISSION) ? 1 : 0, c->d_absorb_last,
                                 c->d_absorb_kind, c->d_absorb_slot, c->d_status);
    c->ph_cur = dst;
    return fetch_status(c);
}

bool do_sync(opal_ctx* c, double ly, double lz) {
    const double* halo = nullptr;
    if (remote(c)) {
        if (c->comm) {
            c->k.halo_pack(c->stream, c->g, c->d_halo_send); c->launches++;
            const size_t bytes = OPAL_HALO_CELLS * OPAL_CELL_DOUBLES * sizeof(double);
            if (!nccl_exchange(c, c->d_halo_send, c->d_halo_send + OPAL_HALO_CELLS * OPAL_CELL_DOUBLES, c->d_halo_recv,
                               c->d_halo_recv + OPAL_HALO_CELLS * OPAL_CELL_DOUBLES, bytes)) return false;
        } else {
            if ((c->has_left() && !c->halo_given[0]) || (c->has_right() && !c->halo_given[1]))
                return fail(c, OPAL_EINVAL, "synchronize: neighbour halo missing (halo_give) (yee.rs:413-419)");
        }
        halo = c->d_halo_recv;
        c->halo_given[0] = c->halo_given[1] = false;
    }
    { Timed tm(c, 2); c->k.sync_bc(c->stream, c->g, halo, ly, lz, nullptr); c->launches++; }
    return true;
}

void do_field(opal_ctx* c) {
    Timed tm(c, 2);
    GridDev nxt = c->g;
    nxt.ex = c->eb_alt[0]; nxt.ey = c->eb_alt[1]; nxt.ez = c->eb_alt[2]; nxt.bx = c->eb_alt[3]; nxt.by = c->eb_alt[4]; nxt.bz = c->eb_alt[5];
    c->k.field(c->stream, c->g, nxt, false); c->launches++;
    c->eb_alt[0] = c->g.ex; c->eb_alt[1] = c->g.ey; c->eb_alt[2] = c->g.ez; c->eb_alt[3] = c->g.bx; c->eb_alt[4] = c->g.by; c->eb_alt[5] = c->g.bz;
    c->g = nxt;
}


// ---- burst mode: the fused step as a replayed CUDA graph ---------------------------------------------------------
// Small decks (two_stream.yaml: 1e5 particles) spend their time in launch latency and in the per-step status
// read-back, not in the kernels.  Two consecutive steps (the E/B ping-pong has period two) are captured once into a
// CUDA graph and replayed; the status block is read back once per batch instead of once per species and step.  What
// the host would have reacted to -- a parked mover (bucket full) or an assert -- raises a halt flag on the device
// instead, and every later kernel of the batch returns at once, so the host finds the state exactly as it was when
// the condition arose, completes that step the ordinary way (re-layout, remaining species, field update) and resumes.
void swap_fields(opal_ctx* c) {
    double** cur[6] = {&c->g.ex, &c->g.ey, &c->g.ez, &c->g.bx, &c->g.by, &c->g.bz};
    for (int k = 0; k < 6; k++) std::swap(*cur[k], c->eb_alt[k]);
}

void swap_sources(opal_ctx* c) {
    double** cur[4] = {&c->g.rho, &c->g.jx, &c->g.jy, &c->g.jz};
    for (int k = 0; k < 4; k++) std::swap(*cur[k], c->j_alt[k]);
}

// what one iteration of main.rs:238-267 consists of; the phases run in this order
struct StepCfg {
    bool dep = true, emit = false;
    bool fused() const { return dep && !emit; }        // deposit inside the push (absorption, which also forbids it, never runs in a burst)
    double f3[3] = {0.0, 0.0, 0.0};                    // emit filters (NaN == None)
};
enum { PH_ELECTRONS = 0, PH_IONS = 1, PH_PHOTONS = 2, PH_ABSORB = 3, PH_EMIT_ROOM = 4, PH_EMIT = 5, PH_DEPOSIT = 6, PH_GRID = 7 };

// One iteration of main.rs:238-267 without the host in the loop (one slab), as few launches as the phases allow: the check
// of an advance (parked movers, asserts, mover segments filling up) rides on the next kernel of the step (GridDev::prev_phase)
// or shares a launch with the next species' begin-mark (burst_mark 3); the begin-mark of the first species rides on k_clear.
// Bursts whose steps end in the one-launch grid step: that launch also clears the source arrays of the NEXT step (a second set:
// opal_ctx::j_alt) and zeroes its per-advance counters, so no step of the burst begins with a clear launch of its own
// (run_burst clears for the first one).
bool clear_rides_on_grid_step(opal_ctx* c, const StepCfg& cfg) {
    const int last_sp = c->bk[1].allocated ? 1 : (c->bk[0].allocated ? 0 : -1);
    const bool photons_move = c->ph[0].cap > 0 && !(c->flags & OPAL_FLAG_IMMOBILE_PHOTONS);
    const bool emit = cfg.emit && c->bk[OPAL_ELECTRON].allocated;
    return cfg.fused() && !photons_move && !emit && last_sp >= 0 && env_ll("OPAL_GRID_FUSED", 1) != 0 && env_ll("OPAL_CLEAR_FUSED", 1) != 0;
}

// Returns the launch count.
int enqueue_step(opal_ctx* c, const StepCfg& cfg, uint64_t emit_call0) {
    int n = 0;
    int pend = -1;                                   // the phase whose check has not been done yet
    const int last_sp = c->bk[1].allocated ? 1 : (c->bk[0].allocated ? 0 : -1);
    const bool photons_move = c->ph[0].cap > 0 && !(c->flags & OPAL_FLAG_IMMOBILE_PHOTONS);
    const bool emit = cfg.emit && c->bk[OPAL_ELECTRON].allocated;
    // nothing between the last particle advance and the grid's share of the step: one launch does its check, halo + BCs, the
    // field update and the step counter (k_grid_step)
    const bool one_grid_launch = cfg.fused() && !photons_move && !emit && last_sp >= 0 && env_ll("OPAL_GRID_FUSED", 1) != 0;
    int* const eq_n = (c->eq_on && c->d_eq_n) ? c->d_eq_n : nullptr;
    int* ph_dst_n = photons_move ? c->ph[1 - c->ph_cur].d_n : nullptr;      // emptied by the first begin-mark of the step
    auto gdev = [&](int prev_phase, int count_step) { GridDev g = c->g; g.prev_phase = prev_phase; g.count_step = count_step; g.seg_cap = c->mover_seg_cap; return g; };
    bool cleared_with_mark = false;
    const bool clear_rides = clear_rides_on_grid_step(c, cfg);
    if (clear_rides) cleared_with_mark = true;       // by the grid launch of the step before (the first step: by run_burst)
    else if (cfg.dep) { c->k.clear(c->stream, c->g, last_sp >= 0 ? c->d_status : nullptr); n++; cleared_with_mark = last_sp >= 0; }
    for (int sp = 0; sp < 2; sp++) {
        Buckets& B = c->bk[sp];
        if (!B.allocated) continue;
        c->adv_calls[sp]++;
        StepParams p = sparams(c, sp);
        int* za = sp == OPAL_ELECTRON ? eq_n : nullptr;
        if (pend >= 0) { c->k.burst_mark(c->stream, c->d_status, 3, pend, c->mover_seg_cap, za, ph_dst_n); ph_dst_n = nullptr; n++; }
        else if (cleared_with_mark) {
            cleared_with_mark = false;
            if (za) { k_burst_zero_int<<<1, 32, 0, c->stream>>>(c->d_status, za); n++; }
        } else { c->k.burst_mark(c->stream, c->d_status, 0, sp, c->mover_seg_cap, za, ph_dst_n); ph_dst_n = nullptr; n++; }
        opal_ctx::Lean& L = c->lean[sp];
        LeanStrips ls = {L.d_strips[L.parity], L.d_scount[L.parity], L.d_strips[1 - L.parity], L.d_scount[1 - L.parity], L.d_hcount};
        const bool strips = c->k.push(c->stream, c->g, bdev(c, sp), p, B.full, cfg.fused(), c->d_movers, L.d_hcount ? &ls : nullptr, c->d_out[0] + 1, c->d_out[1] + 1,
                                      c->d_overflow, c->d_status, nullptr);
        if (strips) { L.parity ^= 1; L.pending = true; L.holes = true; }
        c->k.scatter(c->stream, c->g, bdev(c, sp), p, B.full, cfg.fused(), c->d_movers, nullptr, 0, 0, nullptr, c->d_out[0] + 1, c->d_out[1] + 1,
                     reinterpret_cast<int*>(c->d_overflow), c->d_status);
        n += 2;
        if (strips && p.heavy_min < INT_MAX) { n++; c->n_heavy_launches++; }      // the block-per-cell launch behind the push
        pend = sp;
    }
    if (photons_move) {
        const int src = c->ph_cur, dst = 1 - src;
        if (ph_dst_n) { k_burst_zero_int<<<1, 32, 0, c->stream>>>(c->d_status, c->ph[dst].d_n); n++; }      // (no begin-mark took it along)
        c->k.photon_push(c->stream, gdev(pend, 0), pdev(c, src), pdev(c, dst), sparams(c, OPAL_PHOTON), c->d_out[0] + 1, c->d_out[1] + 1, c->d_status, PairDev());
        c->ph_cur = dst;
        n++;
        pend = PH_PHOTONS;
    }
    if (emit) {
        // t and the call index of step k come from the device-side step counter (GridDev::step / laser_arr[3k + 2])
        c->k.emit(c->stream, gdev(pend, 0), bdev(c, OPAL_ELECTRON), pdev(c, c->ph_cur), sparams(c, OPAL_ELECTRON), c->d_qed_tables, 0.0, cfg.f3[0], cfg.f3[1], cfg.f3[2],
                  c->seed, emit_call0, c->d_status, std::max<long long>(c->bk[OPAL_ELECTRON].n_hint, 1), PH_EMIT_ROOM);
        n++;
        pend = PH_EMIT;
    }
    if (cfg.dep && !cfg.fused()) {
        for (int sp = 0; sp < 2; sp++) if (c->bk[sp].allocated) { c->k.deposit(c->stream, gdev(pend, 0), bdev(c, sp), sparams(c, sp), c->d_status); pend = -1; n++; }
    }
    GridDev nxt = c->g;
    nxt.ex = c->eb_alt[0]; nxt.ey = c->eb_alt[1]; nxt.ez = c->eb_alt[2]; nxt.bx = c->eb_alt[3]; nxt.by = c->eb_alt[4]; nxt.bz = c->eb_alt[5];
    if (one_grid_launch) {
        if (clear_rides) { nxt.rho = c->j_alt[0]; nxt.jx = c->j_alt[1]; nxt.jy = c->j_alt[2]; nxt.jz = c->j_alt[3]; }
        c->k.grid_step(c->stream, c->g, nxt, c->d_status, last_sp, c->d_gridbar, c->mover_seg_cap, clear_rides);
        n++;
    } else {
        c->k.sync_bc(c->stream, gdev(pend, 0), nullptr, 0.0, 0.0, nullptr);
        c->k.field(c->stream, gdev(-1, 1), nxt, false);
        n += 2;
    }
    c->eb_alt[0] = c->g.ex; c->eb_alt[1] = c->g.ey; c->eb_alt[2] = c->g.ez; c->eb_alt[3] = c->g.bx; c->eb_alt[4] = c->g.by; c->eb_alt[5] = c->g.bz;
    if (clear_rides) { c->j_alt[0] = c->g.rho; c->j_alt[1] = c->g.jx; c->j_alt[2] = c->g.jy; c->j_alt[3] = c->g.jz; }
    c->g = nxt;
    return n;
}

// Runs up to n (even) steps from laser_y/z[0..n) as graph replays.  *done = steps completed; *halted tells whether the
// step after them stopped half-way (c->h_status then holds the counters and halt_species of that moment).
bool run_burst(opal_ctx* c, const StepCfg& cfg, int64_t n, double t0, const double* ly, const double* lz, int64_t* done, bool* halted) {
    *done = 0; *halted = false;
    c->n_bursts++;
    if (c->laser_cap < 3 * n) {
        if (c->d_laser) cudaFree(c->d_laser);
        c->d_laser = nullptr; c->laser_cap = 0;
        CUB(cudaMalloc(&c->d_laser, (size_t)(3 * n) * sizeof(double)));
        c->laser_cap = 3 * n;
    }
    // per step: the two laser amplitudes and the time, accumulated as main.rs:266 does (t += dt)
    std::vector<double> lh((size_t)(3 * n), 0.0);
    { double t = t0; for (int64_t k = 0; k < n; k++) { lh[3 * k] = ly ? ly[k] : 0.0; lh[3 * k + 1] = lz ? lz[k] : 0.0; lh[3 * k + 2] = t; t += c->g.dt; } }
    CUB(cudaMemcpyAsync(c->d_laser, lh.data(), lh.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    CUB(cudaMemsetAsync(c->d_status, 0, sizeof(StatusDev) + OPAL_MOVER_SEGMENTS * sizeof(int), c->stream));
    // Where are the particles?  The lean push launches one block per cell; a target that fills a quarter of the grid need not
    // pay for the vacuum around it.  A particle moves less than a cell per step, so nothing leaves [first - n, last + n] within
    // the n steps of this burst (+ 4: the movers waiting in the strips, and the cells that emptied in the last two steps and
    // still have to write their empty strip counts).  A range that touches a periodic edge is not restricted.
    int h_range[4] = {INT_MAX, -1, INT_MAX, -1};
    const bool restrict_ok = cfg.fused() && env_ll("OPAL_PUSH_RANGE", 1) != 0;
    if (restrict_ok) {
        CUB(cudaMemcpyAsync(c->d_tmp_int + 4, h_range, sizeof(h_range), cudaMemcpyHostToDevice, c->stream));
        for (int sp = 0; sp < 2; sp++) if (c->bk[sp].allocated && c->lean[sp].d_hcount) {
            int gr = (c->g.nx + 255) / 256; if (gr > 148 * 4) gr = 148 * 4;
            k_occupied_range<<<gr, 256, 0, c->stream>>>(c->bk[sp].d_count, c->g.nx, c->d_tmp_int + 4 + 2 * sp); c->launches++;
        }
        CUB(cudaMemcpyAsync(h_range, c->d_tmp_int + 4, sizeof(h_range), cudaMemcpyDeviceToHost, c->stream));
    }
    CUB(cudaStreamSynchronize(c->stream));
    for (int sp = 0; sp < 2; sp++) {
        opal_ctx::Lean& L = c->lean[sp];
        L.range_on = false;
        if (!restrict_ok || !c->bk[sp].allocated || !L.d_hcount || h_range[2 * sp + 1] < 0) continue;
        const long long lo = (long long)h_range[2 * sp] - n - 4, hi = (long long)h_range[2 * sp + 1] + n + 4;
        const bool wraps = c->self_nbr() && c->has_left();                       // periodic single slab: the two edges are neighbours
        if (wraps && (lo <= 0 || hi >= c->g.nx - 1)) continue;
        L.c_begin = (int)std::max<long long>(lo, 0); L.c_end = (int)std::min<long long>(hi + 1, c->g.nx);
        L.range_on = L.c_end - L.c_begin < c->g.nx;
    }
    c->g.halt = &c->d_status->halt; c->g.step = &c->d_status->step; c->g.laser_arr = c->d_laser; c->g.bst = c->d_status;
    auto leave = [&]() { c->g.halt = nullptr; c->g.step = nullptr; c->g.laser_arr = nullptr; c->g.bst = nullptr; c->g.prev_phase = -1; c->g.count_step = 0;
                         c->lean[0].range_on = false; c->lean[1].range_on = false; };
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    int per_step = 0;
    const int parity0[2] = {c->lean[0].parity, c->lean[1].parity};     // two captured steps leave the strip parity where it was
    const int ph_cur0 = c->ph_cur;                                     // ... and the photon ping-pong
    const uint64_t emit_call0 = c->emit_calls;
    const bool clear_rides = clear_rides_on_grid_step(c, cfg);
    if (clear_rides) { c->k.clear(c->stream, c->g, c->d_status); c->launches++; }      // the first step's; every later one's rides on a grid launch
    const double tg0 = wall_now();
    cudaError_t e = cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal);
    if (e == cudaSuccess) {
        per_step = enqueue_step(c, cfg, emit_call0);
        enqueue_step(c, cfg, emit_call0);        // the second half of the ping-pong: c->g is back where it started
        e = cudaStreamEndCapture(c->stream, &graph);
    }
    if (e == cudaSuccess) e = cudaGraphInstantiate(&exec, graph, 0);
    if (e != cudaSuccess) { if (graph) cudaGraphDestroy(graph); leave(); return fail(c, OPAL_ECUDA, std::string("burst graph: ") + cudaGetErrorString(e)); }
    c->t_graph += wall_now() - tg0;
    const double tr0 = wall_now();
    bool ok = true;
    int64_t s = 0;
    int spill_seen[2] = {0, 0};
    while (s < n && ok) {
        const int64_t batch = std::min<int64_t>(n - s, 128);
        for (int64_t i = 0; i < batch; i += 2) {
            if (cudaGraphLaunch(exec, c->stream) != cudaSuccess) { ok = fail(c, OPAL_ECUDA, "cudaGraphLaunch failed"); break; }
        }
        if (!ok) break;
        c->launches += batch * per_step;
        if (cudaMemcpyAsync(c->h_status, c->d_status, sizeof(StatusDev) + OPAL_MOVER_SEGMENTS * sizeof(int), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
            cudaStreamSynchronize(c->stream) != cudaSuccess) { ok = fail(c, OPAL_ECUDA, "burst status read-back failed"); break; }
        for (int sp = 0; sp < 2; sp++) {            // (the block was zeroed when the burst began)
            c->lean[sp].spill += c->h_status->strip_spill[sp] - spill_seen[sp]; spill_seen[sp] = c->h_status->strip_spill[sp];
            c->lean[sp].spill_adv += c->h_status->step - *done;
        }
        *done = c->h_status->step;
        if (c->h_status->halt) { *halted = true; c->n_halts++; c->n_wasted += (s + batch) - *done; note_heavy(c); break; }
        s += batch;
        if (note_heavy(c)) break;                    // heavy cells have appeared: the caller starts a new burst with the block-per-cell launch in it
        if (strips_too_narrow(c, 0) || strips_too_narrow(c, 1)) break;      // ... or one with wider mover strips
    }
    c->t_replay += wall_now() - tr0;
    { const double td0 = wall_now(); cudaGraphExecDestroy(exec); cudaGraphDestroy(graph); c->t_graph += wall_now() - td0; }
    leave();
    c->burst_steps += *done;
    if (ok && (*done & 1)) { swap_fields(c); if (clear_rides) swap_sources(c); }      // an odd number of steps ran: the fields (and sources) sit in the other buffer set
    const int hphase = *halted ? c->h_status->halt_species : -1;   // the phase of the half-finished step that raised the halt
    for (int sp = 0; sp < 2; sp++) {
        // pushes of this species that ran: one per completed step, plus the half-finished step up to the halting phase
        const int64_t pushes = *done + ((*halted && c->bk[sp].allocated && sp <= hphase) ? 1 : 0);
        if (c->lean[sp].d_hcount && c->bk[sp].allocated) c->lean[sp].parity = parity0[sp] ^ (int)(pushes & 1);
    }
    if (c->ph[0].cap > 0 && !(c->flags & OPAL_FLAG_IMMOBILE_PHOTONS)) c->ph_cur = ph_cur0 ^ (int)((*done + (hphase >= PH_PHOTONS ? 1 : 0)) & 1);
    if (cfg.emit) c->emit_calls = emit_call0 + (uint64_t)*done + (hphase >= PH_EMIT ? 1 : 0);
    return ok;
}

}  // namespace

// =============================== C ABI ======================================================
extern "C" {

opal_status opal_create(opal_ctx** out, int device, int64_t nx, double xmin, double dx, double dt, int left, int right,
                        int rank, int nranks, uint64_t seed, uint32_t flags) {
    if (!out || nx < 1 || nx > (1ll << 30) || nranks < 1 || rank < 0 || rank >= nranks) return OPAL_EINVAL;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return OPAL_ECUDA;   // no CPU fallback
    if (device < 0 || device >= ndev) return OPAL_EINVAL;
    opal_ctx* c = new opal_ctx();
    *out = c;
    c->device = device; c->rank = rank; c->nranks = nranks; c->flags = flags; c->seed = seed;
    c->full = !(flags & OPAL_FLAG_LEAN_STATE);
    c->k = (flags & OPAL_FLAG_STRICT_FP) ? strict::table() : fast::table();
    c->slack = env_double("OPAL_BUCKET_SLACK", 1.25);
    c->slack_from_env = getenv("OPAL_BUCKET_SLACK") != nullptr || getenv("OPAL_BUCKET_PAD") != nullptr;
    c->pad = (int)env_ll("OPAL_BUCKET_PAD", 64);
    // OPAL_HEAVY_MIN: cell population above which the lean push may hand a cell to a whole block (0: never); OPAL_HEAVY_FORCE=1
    // keeps that second launch on from the first step (tests)
    c->heavy_min = (int)std::max<long long>(0, std::min<long long>(env_ll("OPAL_HEAVY_MIN", 4096), INT_MAX / 4));
    c->heavy_factor = std::max(0.0, env_double("OPAL_HEAVY_FACTOR", 8.0));       // ... and above this many times the average population
    c->heavy_forced = c->heavy_on = env_ll("OPAL_HEAVY_FORCE", 0) != 0 && c->heavy_min > 0;
    c->widen_strips = env_ll("OPAL_STRIP_WIDEN", 1) != 0;
    c->pool_buckets = env_ll("OPAL_BUCKET_POOL", 1) != 0;
    CU(cudaSetDevice(device));
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    GridDev& g = c->g;
    g.nx = (int)nx; g.xmin = xmin; g.dx = dx; g.dt = dt; g.self_nbr = nranks == 1;
    g.halt = nullptr; g.step = nullptr; g.laser_arr = nullptr; g.bst = nullptr; g.prev_phase = -1; g.count_step = 0; g.seg_cap = 0;
    // yee.rs:278-292
    if (rank == 0 && left == OPAL_BDY_LASER) { g.left_bdy = OPAL_BDY_LASER; g.left = OPAL_LASER_BDY_SIZE; }
    else { g.left_bdy = OPAL_BDY_INTERNAL; g.left = OPAL_GHOST_SIZE; }
    if (rank == nranks - 1 && right == OPAL_BDY_ABSORBING) { g.right_bdy = OPAL_BDY_ABSORBING; g.right = OPAL_ABSORBING_BDY_SIZE; }
    else if (rank == nranks - 1 && right == OPAL_BDY_CONDUCTING) { g.right_bdy = OPAL_BDY_CONDUCTING; g.right = OPAL_CONDUCTING_BDY_SIZE; }
    else { g.right_bdy = OPAL_BDY_INTERNAL; g.right = OPAL_GHOST_SIZE; }
    g.len = g.nx + g.left + g.right;
    const size_t n = (size_t)g.len;
    CU(cudaMalloc(&c->grid_block, 20 * n * sizeof(double)));
    CU(cudaMemsetAsync(c->grid_block, 0, 20 * n * sizeof(double), c->stream));
    double* p = c->grid_block;
    g.rho = p; g.jx = p + n; g.jy = p + 2 * n; g.jz = p + 3 * n;
    g.ex = p + 4 * n; g.ey = p + 5 * n; g.ez = p + 6 * n; g.bx = p + 7 * n; g.by = p + 8 * n; g.bz = p + 9 * n;
    for (int k = 0; k < 6; k++) c->eb_alt[k] = p + (10 + k) * n;
    for (int k = 0; k < 4; k++) c->j_alt[k] = p + (16 + k) * n;
    CU(cudaMalloc(&c->d_status, sizeof(StatusDev) + OPAL_MOVER_SEGMENTS * sizeof(int)));
    CU(cudaMemsetAsync(c->d_status, 0, sizeof(StatusDev) + OPAL_MOVER_SEGMENTS * sizeof(int), c->stream));
    CU(cudaMallocHost(&c->h_status, sizeof(StatusDev) + OPAL_MOVER_SEGMENTS * sizeof(int)));
    memset(c->h_status, 0, sizeof(StatusDev));
    CU(cudaMalloc(&c->d_energy, 8 * sizeof(double)));
    CU(cudaMallocHost(&c->h_energy, 8 * sizeof(double)));
    CU(cudaMalloc(&c->d_tmp_int, 8 * sizeof(int)));
    CU(cudaMalloc(&c->d_gridbar, 2 * sizeof(unsigned)));
    CU(cudaMemsetAsync(c->d_gridbar, 0, 2 * sizeof(unsigned), c->stream));
    c->out_cap = c->in_cap = (int)env_ll("OPAL_EXCHANGE_CAP", 4096);      // initial sizes: both grow on demand
    for (int s = 0; s < 2; s++) {
        CU(cudaMalloc(&c->d_out[s], (size_t)(c->out_cap + 1) * sizeof(opal_migrant)));
        CU(cudaMalloc(&c->d_in[s], (size_t)(c->in_cap + 1) * sizeof(opal_migrant)));
        CU(cudaMemsetAsync(c->d_in[s], 0, sizeof(opal_migrant), c->stream));
    }
    CU(cudaMalloc(&c->d_halo_send, 2 * OPAL_HALO_CELLS * OPAL_CELL_DOUBLES * sizeof(double)));
    CU(cudaMalloc(&c->d_halo_recv, 2 * OPAL_HALO_CELLS * OPAL_CELL_DOUBLES * sizeof(double)));
    CU(cudaMalloc(&c->d_qed_tables, sizeof(H_QED_TABLES)));
    CU(cudaMemcpyAsync(c->d_qed_tables, H_QED_TABLES, sizeof(H_QED_TABLES), cudaMemcpyHostToDevice, c->stream));
    if (!ensure_movers(c, 1024)) return (opal_status)c->status;
    CU(cudaStreamSynchronize(c->stream));
    return OPAL_OK;
}

opal_status opal_destroy(opal_ctx* c) {
    if (!c) return OPAL_EINVAL;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (const char* e = getenv("OPAL_STATS")) if (e[0] == '1')
        fprintf(stderr, "opal_b200 stats: launches %lld, burst steps %lld in %lld bursts, %lld halts (%lld replayed steps skipped behind them), %lld re-layouts; %lld block-per-cell launches enqueued, strips widened %lld times (now %d / %d records); host seconds: re-layouts %.3f (counts %.3f, sizing %.3f, alloc %.3f, copy %.3f, free %.3f), graph build %.3f, replays %.3f\n",
                (long long)c->launches, (long long)c->burst_steps, (long long)c->n_bursts, (long long)c->n_halts, (long long)c->n_wasted, (long long)c->n_regrids, (long long)c->n_heavy_launches, (long long)c->n_widened, c->lean[0].mc, c->lean[1].mc, c->t_regrid, c->t_rg[0], c->t_rg[1], c->t_rg[2], c->t_rg[3], c->t_rg[4], c->t_graph, c->t_replay);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    for (int k = 0; k < 3; k++) for (auto& p : c->ev[k]) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
    for (int s = 0; s < 2; s++) { free_buckets(c->bk[s]); free_buckets(c->spare[s]); }
    if (c->d_absorb_kind) cudaFree(c->d_absorb_kind);
    if (c->d_absorb_slot) cudaFree(c->d_absorb_slot);
    if (c->d_absorb_last) cudaFree(c->d_absorb_last);
    for (int w = 0; w < 2; w++) {
        if (c->ph[w].d_n) cudaFree(c->ph[w].d_n);
        if (c->ph[w].d_cell) cudaFree(c->ph[w].d_cell);
        for (int k = 0; k < NF_PHOTON; k++) if (c->ph[w].d_f[k]) cudaFree(c->ph[w].d_f[k]);
    }
    void* frees[] = {c->grid_block, c->d_movers, c->d_overflow, c->d_out[0], c->d_out[1], c->d_in[0], c->d_in[1], c->d_halo_send,
                     c->d_halo_recv, c->d_status, c->d_energy, c->d_tmp_int, c->d_qed_tables, c->d_laser, c->d_hgram, c->d_eq_slot, c->d_eq_cell, c->d_eq_n, c->d_pair[0], c->d_pair[1], c->d_pair_n, c->d_gridbar};
    for (void* p : frees) if (p) cudaFree(p);
    for (int sp = 0; sp < 2; sp++) {
        for (int k = 0; k < 2; k++) { if (c->lean[sp].d_strips[k]) cudaFree(c->lean[sp].d_strips[k]); if (c->lean[sp].d_scount[k]) cudaFree(c->lean[sp].d_scount[k]); }
        if (c->lean[sp].d_hcount) cudaFree(c->lean[sp].d_hcount);
    }
    if (c->h_status) cudaFreeHost(c->h_status);
    if (c->h_energy) cudaFreeHost(c->h_energy);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return OPAL_OK;
}

const char* opal_last_error(opal_ctx* c) { return c ? c->err.c_str() : "null context"; }

opal_status opal_grid_shape(opal_ctx* c, int64_t* total, int64_t* l, int64_t* r) {
    if (!c) return OPAL_EINVAL;
    if (total) *total = c->g.len;
    if (l) *l = c->g.left;
    if (r) *r = c->g.right;
    return OPAL_OK;
}

opal_status opal_set_ion(opal_ctx* c, double z_star, double a) {
    CHECK_CTX(c);
    if (c->flags & OPAL_FLAG_PAIR_CREATION) return (opal_status)(fail(c, OPAL_EINVAL, "OPAL_FLAG_PAIR_CREATION keeps the positrons in the ion slot: a run with ions cannot use it"), c->status);
    c->ion_slot_is_user = true;
    c->z_star = z_star; c->mass_number = a;
    return OPAL_OK;
}

opal_status opal_particles_upload(opal_ctx* c, int species, int64_t n, const int64_t* cell, const double* x, const double* ux,
                                  const double* uy, const double* uz, const double* w, const double* tau) {
    CHECK_CTX(c);
    c->ke_fused = false;
    if (species >= 0 && species <= 2) c->edge_bound[species][0] = c->edge_bound[species][1] = -1;
    if (n < 0 || n > (1ll << 31) - 64 || (n > 0 && (!cell || !x || !ux || !uy || !uz || !w))) return OPAL_EINVAL;
    const double dx = c->g.dx, dt = c->g.dt;
    const double inf = std::numeric_limits<double>::infinity();
    if (species == OPAL_PHOTON) {
        if (!ensure_photons(c, std::max<int64_t>(n, 1))) return (opal_status)c->status;
        Photons& P = c->ph[c->ph_cur];
        std::vector<int> hc((size_t)n);
        std::vector<double> hf[NF_PHOTON];
        for (auto& v : hf) v.assign((size_t)n, 0.0);
        for (int64_t i = 0; i < n; i++) {
            if (cell[i] < 0 || cell[i] >= c->g.nx) return (opal_status)(fail(c, OPAL_EINVAL, "upload: cell out of range"), c->status);
            hc[i] = (int)cell[i];
            const double k0 = std::sqrt(ux[i] * ux[i] + uy[i] * uy[i] + uz[i] * uz[i]);   // photon.rs:95-116
            hf[P_X][i] = x[i]; hf[P_PREVX][i] = x[i] - C_LIGHT * ux[i] * dt / (k0 * dx);
            hf[P_KX][i] = ux[i]; hf[P_KY][i] = uy[i]; hf[P_KZ][i] = uz[i]; hf[P_W][i] = w[i];
            hf[P_TAU][i] = tau ? tau[i] : inf; hf[P_BIRTH][i] = -inf;
            hf[P_TAU_ABS][i] = hf[P_TAU_ST][i] = hf[P_TAU][i];            // with_optical_depth (photon.rs:118-124)
        }
        const int nn = (int)n;
        CU(cudaMemcpyAsync(P.d_n, &nn, sizeof(int), cudaMemcpyHostToDevice, c->stream));
        if (n > 0) {
            CU(cudaMemcpyAsync(P.d_cell, hc.data(), (size_t)n * sizeof(int), cudaMemcpyHostToDevice, c->stream));
            for (int k = 0; k < NF_PHOTON; k++) CU(cudaMemcpyAsync(P.d_f[k], hf[k].data(), (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
        }
        CU(cudaStreamSynchronize(c->stream));
        return OPAL_OK;
    }
    if (species != OPAL_ELECTRON && species != OPAL_ION) return OPAL_EINVAL;
    Buckets& B = c->bk[species];
    free_buckets(B);
    B = Buckets();
    // n == 0 on a lone slab: the species does not exist and every later call skips it.  On a slab of a multi-slab run the
    // species may live on a neighbour: allocate empty buckets so this slab takes part in the per-step neighbour exchange
    // (every rank of the reference exchanges every step, mod.rs:308-354) and can receive migrants.
    if (n == 0 && c->nranks == 1) return OPAL_OK;
    const int ncell = c->g.nx;
    std::vector<long long> demand(ncell, 0);
    for (int64_t i = 0; i < n; i++) {
        if (cell[i] < 0 || cell[i] >= ncell) return (opal_status)(fail(c, OPAL_EINVAL, "upload: cell out of range"), c->status);
        demand[cell[i]]++;
    }
    if (!alloc_buckets(c, B, demand, c->full)) return (opal_status)c->status;
    B.n_hint = n;
    if (!ensure_movers(c, n)) return (opal_status)c->status;
    if (!ensure_cell_movers(c, species, n)) return (opal_status)c->status;
    // host-side bucketing into the slot layout (stable: keeps the caller's order inside a cell)
    const int nf = c->full ? NF_FULL : NF_LEAN;
    std::vector<std::vector<double>> hf(nf, std::vector<double>((size_t)B.slots, 0.0));
    std::vector<int> fill(ncell, 0);
    for (int64_t i = 0; i < n; i++) {
        const long long s = B.h_start[cell[i]] + fill[cell[i]]++;
        hf[F_X][s] = x[i]; hf[F_UX][s] = ux[i]; hf[F_UY][s] = uy[i]; hf[F_UZ][s] = uz[i]; hf[F_W][s] = w[i];
        if (c->full) {
            const double u2 = ux[i] * ux[i] + uy[i] * uy[i] + uz[i] * uz[i];
            if (species == OPAL_ELECTRON) {           // electron.rs:85-103
                const double gamma = std::sqrt(1.0 + u2);
                hf[F_GAMMA][s] = gamma;
                hf[F_PREVX][s] = x[i] - C_LIGHT * ux[i] * dt / (gamma * dx);
                hf[F_TAU][s] = tau ? tau[i] : inf;
            } else {                                  // ion.rs:81-101
                const double gm1 = u2 / (1.0 + std::sqrt(1.0 + u2));
                const double vx = C_LIGHT * ux[i] / (1.0 + gm1);
                hf[F_GAMMA][s] = gm1;
                hf[F_PREVX][s] = x[i] - vx * dt / dx;
                hf[F_TAU][s] = inf;
            }
        }
    }
    CU(cudaMemcpyAsync(B.d_count, fill.data(), (size_t)ncell * sizeof(int), cudaMemcpyHostToDevice, c->stream));
    for (int k = 0; k < nf; k++) CU(cudaMemcpyAsync(B.d_f[k], hf[k].data(), (size_t)B.slots * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return OPAL_OK;
}

opal_status opal_particles_generate(opal_ctx* c, int species, int64_t npc, double weight, const double u_th[3], const double drift[3],
                                    int two_stream, int64_t global_cell_offset, uint64_t seed) {
    CHECK_CTX(c);
    c->ke_fused = false;
    if (species == OPAL_ELECTRON || species == OPAL_ION) c->edge_bound[species][0] = c->edge_bound[species][1] = -1;
    if ((species != OPAL_ELECTRON && species != OPAL_ION) || npc < 1 || !u_th || !drift) return OPAL_EINVAL;
    Buckets& B = c->bk[species];
    free_buckets(B);
    B = Buckets();
    std::vector<long long> demand(c->g.nx, npc);
    if (!alloc_buckets(c, B, demand, c->full)) return (opal_status)c->status;
    B.n_hint = npc * (long long)c->g.nx;
    if (!ensure_movers(c, B.n_hint)) return (opal_status)c->status;
    if (!ensure_cell_movers(c, species, B.n_hint)) return (opal_status)c->status;
    c->k.generate(c->stream, bdev(c, species), species, npc, weight, u_th, drift, two_stream, global_cell_offset, seed, c->full, c->g.dx, c->g.dt);
    c->launches++;
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    return OPAL_OK;
}

opal_status opal_particles_count(opal_ctx* c, int species, int64_t* n) {
    CHECK_CTX(c);
    if (!settle(c, species)) return (opal_status)c->status;
    if (!n) return OPAL_EINVAL;
    *n = 0;
    if (species == OPAL_PHOTON) {
        int v = 0;
        if (!photon_count(c, &v)) return (opal_status)c->status;
        *n = v;
        return OPAL_OK;
    }
    if (species != OPAL_ELECTRON && species != OPAL_ION) return OPAL_EINVAL;
    Buckets& B = c->bk[species];
    if (!B.allocated) return OPAL_OK;
    std::vector<int> cnt(c->g.nx);
    CU(cudaMemcpyAsync(cnt.data(), B.d_count, cnt.size() * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    long long t = 0;
    for (int v : cnt) t += v;
    *n = t;
    return OPAL_OK;
}

opal_status opal_particles_download(opal_ctx* c, int species, int64_t cap, int64_t* cell, double* x, double* prev_x, double* ux,
                                    double* uy, double* uz, double* w, double* gamma, double* chi, double* tau, double* work,
                                    double* y, double* z, double* birth) {
    CHECK_CTX(c);
    if (!settle(c, species)) return (opal_status)c->status;
    const double inf = std::numeric_limits<double>::infinity();
    if (species == OPAL_PHOTON) {
        int n = 0;
        if (!photon_count(c, &n)) return (opal_status)c->status;
        if (n > cap) return OPAL_ECAPACITY;
        if (n == 0) return OPAL_OK;
        Photons& P = c->ph[c->ph_cur];
        std::vector<int> hc(n);
        std::vector<double> hf[NF_PHOTON];
        CU(cudaMemcpyAsync(hc.data(), P.d_cell, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        for (int k = 0; k < NF_PHOTON; k++) { hf[k].resize(n); CU(cudaMemcpyAsync(hf[k].data(), P.d_f[k], (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, c->stream)); }
        CU(cudaStreamSynchronize(c->stream));
        // Population::advance leaves the store sorted by cell (mod.rs:288)
        std::vector<int> order(n);
        for (int i = 0; i < n; i++) order[i] = i;
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return hc[a] < hc[b]; });
        for (int o = 0; o < n; o++) {
            const int i = order[o];
            if (cell) cell[o] = hc[i];
            if (x) x[o] = hf[P_X][i];
            if (prev_x) prev_x[o] = hf[P_PREVX][i];
            if (ux) ux[o] = hf[P_KX][i];
            if (uy) uy[o] = hf[P_KY][i];
            if (uz) uz[o] = hf[P_KZ][i];
            if (w) w[o] = hf[P_W][i];
            if (gamma) gamma[o] = std::sqrt(hf[P_KX][i] * hf[P_KX][i] + hf[P_KY][i] * hf[P_KY][i] + hf[P_KZ][i] * hf[P_KZ][i]);
            if (chi) chi[o] = hf[P_CHI][i];
            if (tau) tau[o] = hf[P_TAU][i];
            if (work) work[o] = 0.0;
            if (y) y[o] = hf[P_Y][i];
            if (z) z[o] = hf[P_Z][i];
            if (birth) birth[o] = hf[P_BIRTH][i];
        }
        return OPAL_OK;
    }
    if (species != OPAL_ELECTRON && species != OPAL_ION) return OPAL_EINVAL;
    Buckets& B = c->bk[species];
    if (!B.allocated) return OPAL_OK;
    const int ncell = c->g.nx;
    std::vector<int> cnt(ncell);
    CU(cudaMemcpyAsync(cnt.data(), B.d_count, (size_t)ncell * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    std::vector<long long> off(ncell);
    long long n = 0;
    for (int i = 0; i < ncell; i++) { off[i] = n; n += cnt[i]; }
    if (n > cap) return OPAL_ECAPACITY;
    if (n == 0) return OPAL_OK;
    long long* d_off = nullptr; long long* d_cell = nullptr;
    CU(cudaMalloc(&d_off, (size_t)ncell * sizeof(long long)));
    CU(cudaMemcpyAsync(d_off, off.data(), (size_t)ncell * sizeof(long long), cudaMemcpyHostToDevice, c->stream));
    if (cell) CU(cudaMalloc(&d_cell, (size_t)n * sizeof(long long)));
    double* host_dst[NF_FULL] = {x, ux, uy, uz, w, prev_x, gamma, chi, tau, work, y, z};
    double* d_out[NF_FULL] = {nullptr};
    const int nf = B.full ? NF_FULL : NF_LEAN;
    for (int k = 0; k < nf; k++) if (host_dst[k]) CU(cudaMalloc(&d_out[k], (size_t)n * sizeof(double)));
    c->k.gather_out(c->stream, bdev(c, species), B.full, species, d_off, d_cell, d_out); c->launches++;
    if (cell) CU(cudaMemcpyAsync(cell, d_cell, (size_t)n * sizeof(long long), cudaMemcpyDeviceToHost, c->stream));
    for (int k = 0; k < nf; k++) if (host_dst[k]) CU(cudaMemcpyAsync(host_dst[k], d_out[k], (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaFree(d_off); if (d_cell) cudaFree(d_cell);
    for (int k = 0; k < nf; k++) if (d_out[k]) cudaFree(d_out[k]);
    if (B.full) {
        if (species == OPAL_ION) {
            if (gamma) for (long long i = 0; i < n; i++) gamma[i] = 1.0 + gamma[i];
            if (chi) for (long long i = 0; i < n; i++) chi[i] = 0.0;
            if (tau) for (long long i = 0; i < n; i++) tau[i] = inf;
        }
    } else {
        // lean layout: derived quantities are recomputed, unobservable ones reported as zero
        for (long long i = 0; i < n; i++) {
            if (gamma) {
                if (!ux || !uy || !uz) return (opal_status)(fail(c, OPAL_EINVAL, "lean download of gamma needs ux, uy, uz"), c->status);
                gamma[i] = std::sqrt(1.0 + (ux[i] * ux[i] + uy[i] * uy[i] + uz[i] * uz[i]));
            }
            if (prev_x) prev_x[i] = std::numeric_limits<double>::quiet_NaN();
            if (chi) chi[i] = 0.0;
            if (tau) tau[i] = inf;
            if (work) work[i] = 0.0;
            if (y) y[i] = 0.0;
            if (z) z[i] = 0.0;
        }
    }
    if (birth) for (long long i = 0; i < n; i++) birth[i] = 0.0;
    return OPAL_OK;
}

opal_status opal_fields_upload(opal_ctx* c, const double* cells11) {
    CHECK_CTX(c);
    const int len = c->g.len;
    std::vector<double> soa((size_t)10 * len);
    for (int i = 0; i < len; i++)
        for (int k = 0; k < 10; k++) soa[(size_t)k * len + i] = cells11[(size_t)i * 11 + 1 + k];
    c->bx_zero = true;
    for (int i = 0; i < len; i++) if (soa[(size_t)7 * len + i] != 0.0) { c->bx_zero = false; break; }
    double* dst[10] = {c->g.rho, c->g.jx, c->g.jy, c->g.jz, c->g.ex, c->g.ey, c->g.ez, c->g.bx, c->g.by, c->g.bz};
    for (int k = 0; k < 10; k++) CU(cudaMemcpyAsync(dst[k], soa.data() + (size_t)k * len, (size_t)len * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return OPAL_OK;
}

opal_status opal_fields_download(opal_ctx* c, double* cells11) {
    CHECK_CTX(c);
    const int len = c->g.len;
    std::vector<double> soa((size_t)10 * len);
    double* src[10] = {c->g.rho, c->g.jx, c->g.jy, c->g.jz, c->g.ex, c->g.ey, c->g.ez, c->g.bx, c->g.by, c->g.bz};
    for (int k = 0; k < 10; k++) CU(cudaMemcpyAsync(soa.data() + (size_t)k * len, src[k], (size_t)len * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    for (int i = 0; i < len; i++) {
        cells11[(size_t)i * 11] = c->g.xmin + ((double)i - (double)c->g.left) * c->g.dx;   // yee.rs:296
        for (int k = 0; k < 10; k++) cells11[(size_t)i * 11 + 1 + k] = soa[(size_t)k * len + i];
    }
    return OPAL_OK;
}

opal_status opal_advance(opal_ctx* c, int species) {
    CHECK_CTX(c);
    c->ke_fused = false;
    if (species == OPAL_PHOTON) { advance_photons(c, true); return (opal_status)c->status; }
    if (species != OPAL_ELECTRON && species != OPAL_ION) return OPAL_EINVAL;
    if (!c->full) return (opal_status)(fail(c, OPAL_EINVAL, "lean-state context: only opal_run_steps is available"), c->status);
    advance_charged(c, species, false, true);
    return (opal_status)c->status;
}

opal_status opal_last_leavers(opal_ctx* c, int species, int64_t* l, int64_t* r) {
    if (!c || species < 0 || species > 2) return OPAL_EINVAL;
    if (l) *l = c->last_gone[species][0];
    if (r) *r = c->last_gone[species][1];
    return OPAL_OK;
}

opal_status opal_migrants_take(opal_ctx* c, int species, int side, int64_t cap, int64_t* n, opal_migrant* out) {
    CHECK_CTX(c);
    if (side != OPAL_LEFT && side != OPAL_RIGHT) return OPAL_EINVAL;
    (void)species;   // the outbox holds the leavers of the most recent advance()
    *n = 0;
    if (!remote(c)) return OPAL_OK;
    if ((side == OPAL_LEFT && !c->has_left()) || (side == OPAL_RIGHT && !c->has_right())) return OPAL_OK;
    opal_migrant hdr;
    CU(cudaMemcpyAsync(&hdr, c->d_out[side], sizeof(hdr), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *n = hdr.cell;
    if (*n > cap) return OPAL_ECAPACITY;
    if (*n > 0) {
        CU(cudaMemcpyAsync(out, c->d_out[side] + 1, (size_t)*n * sizeof(opal_migrant), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
    }
    return OPAL_OK;
}

opal_status opal_migrants_give(opal_ctx* c, int species, int side, int64_t n, const opal_migrant* in) {
    CHECK_CTX(c);
    if (!settle(c, species)) return (opal_status)c->status;
    c->ke_fused = false;
    if (side != OPAL_LEFT && side != OPAL_RIGHT || n < 0 || species < 0 || species > 2) return OPAL_EINVAL;
    if (n == 0) return OPAL_OK;
    if (!ensure_inbox(c, n)) return (opal_status)c->status;
    opal_migrant hdr;
    memset(&hdr, 0, sizeof(hdr));
    hdr.cell = n;
    CU(cudaMemcpyAsync(c->d_in[side], &hdr, sizeof(hdr), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->d_in[side] + 1, in, (size_t)n * sizeof(opal_migrant), cudaMemcpyHostToDevice, c->stream));
    if (species == OPAL_PHOTON) { if (!ensure_photons(c, c->ph[0].cap / 2 + n)) return (opal_status)c->status; }
    else if (!c->bk[species].allocated) {
        std::vector<long long> demand(c->g.nx, 0);
        if (!alloc_buckets(c, c->bk[species], demand, c->full)) return (opal_status)c->status;
    }
    CU(cudaMemsetAsync(&c->d_status->n_overflow, 0, sizeof(int), c->stream));
    insert_inbox(c, species, side, false);
    if (!fetch_status(c)) return (opal_status)c->status;
    if (species != OPAL_PHOTON) handle_overflow(c, species, false);
    return (opal_status)c->status;
}

opal_status opal_emit_radiation(opal_ctx* c, double t, const double* min_energy_mev, const double* max_angle, const double* max_fl) {
    CHECK_CTX(c);
    c->ke_fused = false;
    const double nan = std::numeric_limits<double>::quiet_NaN();
    do_emit(c, t, min_energy_mev ? *min_energy_mev : nan, max_angle ? *max_angle : nan, max_fl ? *max_fl : nan);
    return (opal_status)c->status;
}

opal_status opal_absorb(opal_ctx* c, double t, const double* max_displacement, const double* stop_time) {
    CHECK_CTX(c);
    c->ke_fused = false;
    const double nan = std::numeric_limits<double>::quiet_NaN();
    do_absorb(c, t, max_displacement ? *max_displacement : nan, stop_time ? *stop_time : nan);
    return (opal_status)c->status;
}

opal_status opal_set_absorption(opal_ctx* c, int enabled, const double* max_displacement, const double* stop_time) {
    CHECK_CTX(c);
    if (enabled && !c->full) return (opal_status)(fail(c, OPAL_EINVAL, "absorption needs the full particle state (context created with OPAL_FLAG_LEAN_STATE)"), c->status);
    const double nan = std::numeric_limits<double>::quiet_NaN();
    c->absorb_on = enabled != 0;
    c->absorb_max_displacement = max_displacement ? *max_displacement : nan;
    c->absorb_stop_time = stop_time ? *stop_time : nan;
    return OPAL_OK;
}

opal_status opal_photon_pair_depths(opal_ctx* c, int set, int64_t n, double* tau_abs, double* tau_st) {
    CHECK_CTX(c);
    int nph = 0;
    if (!photon_count(c, &nph)) return (opal_status)c->status;
    if (n != nph || (n > 0 && (!tau_abs || !tau_st))) return OPAL_EINVAL;
    if (n == 0) return OPAL_OK;
    Photons& P = c->ph[c->ph_cur];
    std::vector<int> hc((size_t)n), order((size_t)n);
    std::vector<double> a((size_t)n), b((size_t)n);
    CU(cudaMemcpyAsync(hc.data(), P.d_cell, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    if (!set) {
        CU(cudaMemcpyAsync(a.data(), P.d_f[P_TAU_ABS], (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaMemcpyAsync(b.data(), P.d_f[P_TAU_ST], (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    }
    CU(cudaStreamSynchronize(c->stream));
    for (int64_t i = 0; i < n; i++) order[(size_t)i] = (int)i;
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return hc[(size_t)x] < hc[(size_t)y]; });      // the download order
    for (int64_t o = 0; o < n; o++) {
        const size_t i = (size_t)order[(size_t)o];
        if (set) { a[i] = tau_abs[o]; b[i] = tau_st[o]; }
        else { tau_abs[o] = a[i]; tau_st[o] = b[i]; }
    }
    if (set) {
        CU(cudaMemcpyAsync(P.d_f[P_TAU_ABS], a.data(), (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemcpyAsync(P.d_f[P_TAU_ST], b.data(), (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
        CU(cudaStreamSynchronize(c->stream));
    }
    return OPAL_OK;
}

opal_status opal_clear(opal_ctx* c) {
    CHECK_CTX(c);
    c->k.clear(c->stream, c->g, nullptr); c->launches++;
    return OPAL_OK;
}

opal_status opal_deposit(opal_ctx* c, int species) {
    CHECK_CTX(c);
    if (species == OPAL_PHOTON) return OPAL_OK;   // charge 0 (photon.rs:197-199)
    if (species != OPAL_ELECTRON && species != OPAL_ION) return OPAL_EINVAL;
    if (!c->full) return (opal_status)(fail(c, OPAL_EINVAL, "lean-state context: only opal_run_steps is available"), c->status);
    if (!c->bk[species].allocated) return OPAL_OK;
    c->k.deposit(c->stream, c->g, bdev(c, species), sparams(c, species), c->d_status); c->launches++;
    return OPAL_OK;
}

opal_status opal_halo_take(opal_ctx* c, int side, double* out) {
    CHECK_CTX(c);
    if (side != OPAL_LEFT && side != OPAL_RIGHT) return OPAL_EINVAL;
    c->k.halo_pack(c->stream, c->g, c->d_halo_send); c->launches++;
    CU(cudaMemcpyAsync(out, c->d_halo_send + side * OPAL_HALO_CELLS * OPAL_CELL_DOUBLES, OPAL_HALO_CELLS * OPAL_CELL_DOUBLES * sizeof(double),
                       cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return OPAL_OK;
}

opal_status opal_halo_give(opal_ctx* c, int side, const double* in) {
    CHECK_CTX(c);
    if (side != OPAL_LEFT && side != OPAL_RIGHT) return OPAL_EINVAL;
    CU(cudaMemcpyAsync(c->d_halo_recv + side * OPAL_HALO_CELLS * OPAL_CELL_DOUBLES, in, OPAL_HALO_CELLS * OPAL_CELL_DOUBLES * sizeof(double),
                       cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->halo_given[side] = true;
    return OPAL_OK;
}

opal_status opal_synchronize(opal_ctx* c, double t, double laser_y, double laser_z) {
    CHECK_CTX(c);
    (void)t;
    do_sync(c, laser_y, laser_z);
    return (opal_status)c->status;
}

opal_status opal_laser_x(opal_ctx* c, double* x) {
    if (!c || !x) return OPAL_EINVAL;
    *x = c->g.xmin + ((double)(c->g.left - 2) - (double)c->g.left) * c->g.dx;
    return OPAL_OK;
}

opal_status opal_field_advance(opal_ctx* c) {
    CHECK_CTX(c);
    do_field(c);
    return OPAL_OK;
}

// One iteration of main.rs:238-267 the ordinary way (status read back after every advance), from phase `from` on: the
// whole step for from == PH_ELECTRONS; the rest of a step a graph-replayed burst left half-finished otherwise.
static bool run_phases(opal_ctx* c, const StepCfg& cfg, int from, double t, double ly, double lz, bool want_ke) {
    const bool fused = cfg.fused() && !c->absorb_on && !(c->flags & OPAL_FLAG_PAIR_CREATION);   // (leptons born mid-step are deposited by the separate pass)
    if (from <= PH_ELECTRONS) {
        if (cfg.dep) { c->k.clear(c->stream, c->g, nullptr); c->launches++; }
        if (!advance_charged(c, OPAL_ELECTRON, fused, true, want_ke)) return false;
    }
    if (from <= PH_IONS && !advance_charged(c, OPAL_ION, fused, true)) return false;
    if (from <= PH_PHOTONS && !advance_photons(c, c->ph[0].cap > 0)) return false;
    if (from <= PH_ABSORB && c->absorb_on && !do_absorb(c, t, c->absorb_max_displacement, c->absorb_stop_time)) return false;      // main.rs:246-248
    if (from <= PH_EMIT && cfg.emit && !do_emit(c, t, cfg.f3[0], cfg.f3[1], cfg.f3[2])) return false;
    if (from <= PH_DEPOSIT && cfg.dep && !fused) {
        for (int sp = 0; sp < 2; sp++) if (c->bk[sp].allocated) { c->k.deposit(c->stream, c->g, bdev(c, sp), sparams(c, sp), c->d_status); c->launches++; }
    }
    if (!do_sync(c, ly, lz)) return false;
    do_field(c);
    return true;
}

opal_status opal_run_steps(opal_ctx* c, int64_t nsteps, double t0, const double* laser_y, const double* laser_z,
                           int current_deposition, int photon_emission, const double* f3) {
    CHECK_CTX(c);
    if (remote(c) && !c->comm) return (opal_status)(fail(c, OPAL_ENCCL, "run_steps with remote neighbours needs opal_comm_init"), c->status);
    if (!c->full && (photon_emission || c->absorb_on || !current_deposition))
        return (opal_status)(fail(c, OPAL_EINVAL, "lean-state context supports only the fused push+deposit step (deposition on, emission and absorption off)"), c->status);
    const double nan = std::numeric_limits<double>::quiet_NaN();
    StepCfg cfg;
    cfg.dep = current_deposition != 0; cfg.emit = photon_emission != 0;
    for (int k = 0; k < 3; k++) cfg.f3[k] = f3 ? f3[k] : nan;
    if (cfg.emit && c->bk[OPAL_ELECTRON].allocated && !c->bk[OPAL_ELECTRON].full)
        return (opal_status)(fail(c, OPAL_EINVAL, "emit_radiation needs the full particle state (context created with OPAL_FLAG_LEAN_STATE)"), c->status);
    double t = t0;
    c->ke_fused = false;
    int64_t s0 = 0;
    c->eq_on = cfg.emit && !c->absorb_on && c->bk[OPAL_ELECTRON].allocated && !(c->flags & OPAL_FLAG_STRICT_FP);
    if (c->eq_on && !ensure_emit_queue(c)) return (opal_status)c->status;
    // room for the cell-crossing particles of a freshly loaded population (before any graph capture: it reads back a count)
    if (!size_movers_for(c, OPAL_ELECTRON) || !size_movers_for(c, OPAL_ION)) return (opal_status)c->status;
    // Launch- and sync-bound decks: replay the step as a CUDA graph (all but the last one or two steps, which also sum the
    // kinetic energy and report the leaver counts).  Everything a single slab does in a step is covered -- fused or separate
    // deposit, photons, emission; absorption, remote neighbours and per-kernel timing take the ordinary path.
    const bool burst_ok = !remote(c) && !c->timing && !c->absorb_on && !(c->flags & OPAL_FLAG_PAIR_CREATION) && env_ll("OPAL_BURST", 1) != 0 &&
                          (long long)c->bk[0].n_hint + c->bk[1].n_hint <= env_ll("OPAL_BURST_MAX_PARTICLES", 20000000);
    while (burst_ok && nsteps - s0 >= 6 && c->status == OPAL_OK) {
        if (cfg.emit && c->bk[OPAL_ELECTRON].allocated) {
            // room for a few steps' worth of photons up front (every electron can emit one per step; a few per cent do)
            int nph = 0;
            if (!photon_count(c, &nph)) break;
            if (!ensure_photons(c, (long long)nph + 2 * c->bk[OPAL_ELECTRON].n_hint)) break;
        }
        if (!maybe_widen_strips(c)) break;
        const int64_t n = ((nsteps - s0 - 1) / 2) * 2;
        int64_t done = 0;
        bool halted = false;
        if (!run_burst(c, cfg, n, t, laser_y ? laser_y + s0 : nullptr, laser_z ? laser_z + s0 : nullptr, &done, &halted)) break;
        s0 += done;
        for (int64_t k = 0; k < done; k++) t += c->g.dt;             // main.rs:266
        if (!halted) { if (done >= n) break; continue; }     // (a burst that ended early at a step boundary: start the next one)
        // step s0 stopped half-way: finish it the ordinary way, then resume
        const unsigned err = c->h_status->err;
        const int hphase = c->h_status->halt_species;
        if (err) { fetch_status(c); break; }                      // an assert: sticky status, as in the ordinary path
        int resume = hphase + 1;
        if (hphase == PH_ELECTRONS || hphase == PH_IONS) {        // a bucket overflowed: re-lay out, re-insert the parked movers
            c->last_gone[hphase][0] = c->h_status->gone[0]; c->last_gone[hphase][1] = c->h_status->gone[1];
            if (!handle_overflow(c, hphase, cfg.fused())) break;
            if (!grow_movers_after(c)) break;                     // (a mover segment more than half full halts a burst too)
        } else if (hphase == PH_EMIT_ROOM) {                      // the photon store is full: do_emit (the next phase) grows it
            resume = PH_EMIT;
        }
        if (!run_phases(c, cfg, resume, t, laser_y ? laser_y[s0] : 0.0, laser_z ? laser_z[s0] : 0.0, false)) break;
        s0++;
        t += c->g.dt;
    }
    for (int64_t s = s0; s < nsteps && c->status == OPAL_OK; s++) {   // main.rs:238-267
        if (!maybe_widen_strips(c)) break;
        // the last step of the call also sums the electrons' kinetic energy (lean strip kernel): opal_energy then needs no particle pass
        if (!run_phases(c, cfg, PH_ELECTRONS, t, laser_y ? laser_y[s] : 0.0, laser_z ? laser_z[s] : 0.0, s == nsteps - 1 && !cfg.emit && !c->absorb_on)) break;
        t += c->g.dt;
    }
    if (c->eq_on) { c->eq_on = false; cudaMemsetAsync(c->d_eq_n, 0xFF, sizeof(int), c->stream); }     // the list belongs to a step of this call
    if (c->status == OPAL_OK) {
        cudaError_t e = cudaStreamSynchronize(c->stream);
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) fail(c, OPAL_ECUDA, std::string("run_steps: ") + cudaGetErrorString(e));
    }
    return (opal_status)c->status;
}

static void* hgram_scratch(opal_ctx* c, size_t bytes) {
    if (bytes > c->hgram_cap) {
        if (c->d_hgram) cudaFree(c->d_hgram);
        c->d_hgram = nullptr; c->hgram_cap = 0;
        const size_t want = std::max<size_t>(bytes, 1 << 20);
        if (cudaMalloc(&c->d_hgram, want) != cudaSuccess) { fail(c, OPAL_ENOMEM, "histogram scratch"); return nullptr; }
        c->hgram_cap = want;
    }
    return c->d_hgram;
}

static bool hgram_args_ok(int species, int ndim, const int* q) {
    if (species < 0 || species > 2 || (ndim != 1 && ndim != 2) || !q) return false;
    for (int d = 0; d < ndim; d++) if (q[d] < 0 || q[d] >= OPAL_Q_COUNT_) return false;
    return true;
}

opal_status opal_hgram_minmax(opal_ctx* c, int species, int ndim, const int* q, const int* lg, double* mn, double* mx, int64_t* n) {
    CHECK_CTX(c);
    if (!settle(c, species)) return (opal_status)c->status;
    if (!hgram_args_ok(species, ndim, q) || !mn || !mx || !n) return OPAL_EINVAL;
    for (int d = 0; d < ndim; d++) { mn[d] = std::numeric_limits<double>::max(); mx[d] = -std::numeric_limits<double>::max(); }   // hgram.rs:189
    opal_status st = opal_particles_count(c, species, n);
    if (st != OPAL_OK || *n == 0) return st;
    unsigned long long* d_mm = static_cast<unsigned long long*>(hgram_scratch(c, 4 * sizeof(unsigned long long) + 4 * sizeof(double)));
    if (!d_mm) return (opal_status)c->status;
    double* d_out = reinterpret_cast<double*>(d_mm + 4);
    if (species == OPAL_PHOTON) hgram_minmax_photon(c->stream, c->g, pdev(c, c->ph_cur), ndim, q, lg, d_mm, d_out);
    else hgram_minmax_charged(c->stream, c->g, bdev(c, species), species, c->bk[species].full, ndim, q, lg, d_mm, d_out);
    c->launches += 3;
    double h[4];
    CU(cudaMemcpyAsync(h, d_out, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    for (int d = 0; d < ndim; d++) { mn[d] = h[2 * d]; mx[d] = h[2 * d + 1]; }
    return OPAL_OK;
}

opal_status opal_hgram_bin(opal_ctx* c, int species, int ndim, const int* q, const int* lg, const double* gmin, const double* bin_sz,
                           const int64_t* nbins, int weight_kind, int log_density_adjust, double* counts, double* total) {
    CHECK_CTX(c);
    if (!settle(c, species)) return (opal_status)c->status;
    if (!hgram_args_ok(species, ndim, q) || !gmin || !bin_sz || !nbins || !counts || !total) return OPAL_EINVAL;
    if (weight_kind != OPAL_W_WEIGHT && weight_kind != OPAL_W_ENERGY) return OPAL_EINVAL;
    long long nb[2] = {1, 1};
    size_t len = 1;
    for (int d = 0; d < ndim; d++) {
        if (nbins[d] < 1 || nbins[d] > (1ll << 28)) return OPAL_EINVAL;
        nb[d] = nbins[d];
        len *= (size_t)nbins[d];
    }
    if (len > (1ull << 30)) return OPAL_EINVAL;
    for (size_t i = 0; i < len; i++) counts[i] = 0.0;
    *total = 0.0;
    int64_t n = 0;
    opal_status st = opal_particles_count(c, species, &n);
    if (st != OPAL_OK || n == 0) return st;
    double* d_counts = static_cast<double*>(hgram_scratch(c, (len + 1) * sizeof(double)));
    if (!d_counts) return (opal_status)c->status;
    CU(cudaMemsetAsync(d_counts, 0, (len + 1) * sizeof(double), c->stream));
    if (species == OPAL_PHOTON) hgram_bin_photon(c->stream, c->g, pdev(c, c->ph_cur), ndim, q, lg, gmin, bin_sz, nb, weight_kind, log_density_adjust, d_counts, d_counts + len);
    else hgram_bin_charged(c->stream, c->g, bdev(c, species), species, c->bk[species].full, ndim, q, lg, gmin, bin_sz, nb, weight_kind, log_density_adjust,
                           d_counts, d_counts + len);
    c->launches++;
    CU(cudaMemcpyAsync(counts, d_counts, len * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(total, d_counts + len, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return OPAL_OK;
}

// ---- opal_hgram_bin_static: opal_hgram_bin for callers that may not allocate (no counterpart in the reference).
// The bin counters -- which opal_hgram_bin keeps in scratch memory it cudaMallocs (and grows) on demand -- live in a file-scope device
// array that every GPU loading the library holds until exit.  Its NAME is a host-side stand-in in this host code, not a device
// address: the device address is looked up once per device with cudaGetSymbolAddress (that device current) and passed to the
// kernels like any other pointer.  All calls share the array (it keeps the last call's counts: cleared on the stream before
// each use); calls that overlap on one GPU -- two contexts, two host threads -- race on it.
static __device__ __align__(256) double g_hgram_static_counts[OPAL_HGRAM_STATIC_MAX_BINS + 1];   // [bins][total]

namespace {
constexpr int HGRAM_STATIC_MAX_DEVICES = 64;
double* g_hgram_static_ptr[HGRAM_STATIC_MAX_DEVICES] = {nullptr};

// the device address of g_hgram_static_counts on the context's device (current: CHECK_CTX), nullptr + error state on failure
double* hgram_static_scratch(opal_ctx* c) {
    if (c->device < 0 || c->device >= HGRAM_STATIC_MAX_DEVICES) { fail(c, OPAL_EINVAL, "opal_hgram_bin_static: device ordinal out of range"); return nullptr; }
    if (!g_hgram_static_ptr[c->device]) {
        double* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_hgram_static_counts);
        if (e != cudaSuccess || !p) {
            fail(c, OPAL_ECUDA, std::string("opal_hgram_bin_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
            return nullptr;
        }
        g_hgram_static_ptr[c->device] = p;
    }
    return g_hgram_static_ptr[c->device];
}
}  // namespace

opal_status opal_hgram_bin_static(opal_ctx* c, int species, int ndim, const int* q, const int* lg, const double* gmin, const double* bin_sz,
                                  const int64_t* nbins, int weight_kind, int log_density_adjust, double* counts, double* total) {
    CHECK_CTX(c);
    if (!hgram_args_ok(species, ndim, q) || !gmin || !bin_sz || !nbins || !counts || !total) return OPAL_EINVAL;
    if (weight_kind != OPAL_W_WEIGHT && weight_kind != OPAL_W_ENERGY) return OPAL_EINVAL;
    long long nb[2] = {1, 1};
    unsigned long long len = 1;
    for (int d = 0; d < ndim; d++) {
        if (nbins[d] < 1 || nbins[d] > (1ll << 28)) return OPAL_EINVAL;
        nb[d] = nbins[d];
        len *= (unsigned long long)nbins[d];
    }
    if (len > (unsigned long long)OPAL_HGRAM_STATIC_MAX_BINS) {
        // not a sticky error: the context stays usable, the message is left for opal_last_error
        c->err = "opal_hgram_bin_static: " + std::to_string(len) + " bins exceed the static scratch of " + std::to_string((long long)OPAL_HGRAM_STATIC_MAX_BINS) +
                 " bins (use opal_hgram_bin)";
        return OPAL_ECAPACITY;
    }
    if (species != OPAL_PHOTON && (c->lean[species].pending || c->lean[species].holes)) {
        // settling the lean layout's mover strips may have to re-lay the buckets out, which allocates
        c->err = "opal_hgram_bin_static: the cell hand-over of the last step is still pending; call any reader (opal_particles_count) first";
        return OPAL_EINVAL;
    }
    for (unsigned long long i = 0; i < len; i++) counts[i] = 0.0;
    *total = 0.0;
    const bool present = species == OPAL_PHOTON ? c->ph[0].cap > 0 : c->bk[species].allocated;
    if (!present) return OPAL_OK;
    double* d_counts = hgram_static_scratch(c);
    if (!d_counts) return (opal_status)c->status;
    CU(cudaMemsetAsync(d_counts, 0, (len + 1) * sizeof(double), c->stream));
    if (species == OPAL_PHOTON) hgram_bin_photon(c->stream, c->g, pdev(c, c->ph_cur), ndim, q, lg, gmin, bin_sz, nb, weight_kind, log_density_adjust, d_counts, d_counts + len);
    else hgram_bin_charged(c->stream, c->g, bdev(c, species), species, c->bk[species].full, ndim, q, lg, gmin, bin_sz, nb, weight_kind, log_density_adjust,
                           d_counts, d_counts + len);
    c->launches++;
    CU(cudaMemcpyAsync(counts, d_counts, len * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(total, d_counts + len, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return OPAL_OK;
}

opal_status opal_energy(opal_ctx* c, double* em, double* ke_e, double* ke_i, double* ke_ph) {
    CHECK_CTX(c);
    BucketDev be, bi;
    PhotonDev pp;
    const bool fused_ke = c->ke_fused && c->bk[0].allocated;
    if (!fused_ke && !settle(c, OPAL_ELECTRON)) return (opal_status)c->status;
    if (!settle(c, OPAL_ION)) return (opal_status)c->status;
    const bool he = c->bk[0].allocated && !fused_ke, hi = c->bk[1].allocated, hp = c->ph[0].cap > 0;
    if (he) be = bdev(c, 0);
    if (hi) bi = bdev(c, 1);
    if (hp) pp = pdev(c, c->ph_cur);
    c->k.energy(c->stream, c->g, he ? &be : nullptr, hi ? &bi : nullptr, hp ? &pp : nullptr, c->full, c->d_energy);
    c->launches += 1 + (he ? 1 : 0) + (hi ? 1 : 0) + (hp ? 1 : 0);
    CU(cudaMemcpyAsync(c->h_energy, c->d_energy, 5 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (em) *em = c->h_energy[0];
    if (ke_e) *ke_e = fused_ke ? c->h_energy[4] : c->h_energy[1];
    if (ke_i) *ke_i = c->h_energy[2];
    if (ke_ph) *ke_ph = c->h_energy[3];
    return OPAL_OK;
}

opal_status opal_comm_unique_id(void* id128) {
    if (!id128) return OPAL_EINVAL;
    if (!g_nccl.load()) return OPAL_ENCCL;
    ncclUniqueId id;
    if (g_nccl.GetUniqueId(&id)) return OPAL_ENCCL;
    memcpy(id128, &id, 128);
    return OPAL_OK;
}

opal_status opal_comm_init(opal_ctx* c, const void* id128) {
    CHECK_CTX(c);
    if (!id128) return OPAL_EINVAL;
    if (!g_nccl.load()) return (opal_status)(fail(c, OPAL_ENCCL, "libnccl.so.2 not found"), c->status);
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    if (g_nccl.CommInitRank(&c->comm, c->nranks, id, c->rank)) return (opal_status)(fail(c, OPAL_ENCCL, "ncclCommInitRank failed"), c->status);
    return OPAL_OK;
}

opal_status opal_stream(opal_ctx* c, void** s) {
    if (!c || !s) return OPAL_EINVAL;
    *s = (void*)c->stream;
    return OPAL_OK;
}

opal_status opal_timing(opal_ctx* c, int enable) {
    CHECK_CTX(c);
    CU(cudaStreamSynchronize(c->stream));
    c->timing = enable != 0;
    for (int k = 0; k < 3; k++) c->ev_used[k] = 0;
    return OPAL_OK;
}

opal_status opal_kernel_times(opal_ctx* c, double ms3[3], int64_t launches3[3]) {
    CHECK_CTX(c);
    if (!ms3 || !launches3) return OPAL_EINVAL;
    CU(cudaStreamSynchronize(c->stream));
    for (int k = 0; k < 3; k++) {
        double total = 0.0;
        for (size_t i = 0; i < c->ev_used[k]; i++) {
            float ms = 0.f;
            CU(cudaEventElapsedTime(&ms, c->ev[k][i].first, c->ev[k][i].second));
            total += ms;
        }
        ms3[k] = total; launches3[k] = (int64_t)c->ev_used[k];
    }
    return OPAL_OK;
}

opal_status opal_burst_steps(opal_ctx* c, int64_t* n) {
    if (!c || !n) return OPAL_EINVAL;
    *n = c->burst_steps;
    return OPAL_OK;
}

opal_status opal_launch_count(opal_ctx* c, int64_t* n) {
    if (!c || !n) return OPAL_EINVAL;
    *n = c->launches;
    return OPAL_OK;
}

// ---- pure functions (no context): run on the current device's default stream -----------------
opal_status opal_qed_rate(int64_t n, const double* chi, const double* gamma, double* out) {
    if (n < 0 || (n > 0 && (!chi || !gamma || !out))) return OPAL_EINVAL;
    if (n == 0) return OPAL_OK;
    double *d_chi = nullptr, *d_g = nullptr, *d_o = nullptr;
    const size_t b = (size_t)n * sizeof(double);
    if (cudaMalloc(&d_chi, b) != cudaSuccess || cudaMalloc(&d_g, b) != cudaSuccess || cudaMalloc(&d_o, b) != cudaSuccess) return OPAL_ECUDA;
    cudaMemcpy(d_chi, chi, b, cudaMemcpyHostToDevice);
    cudaMemcpy(d_g, gamma, b, cudaMemcpyHostToDevice);
    fast::table().qed_rate(0, (int)n, d_chi, d_g, d_o);
    cudaError_t e = cudaMemcpy(out, d_o, b, cudaMemcpyDeviceToHost);
    cudaFree(d_chi); cudaFree(d_g); cudaFree(d_o);
    return e == cudaSuccess ? OPAL_OK : OPAL_ECUDA;
}

opal_status opal_qed_sample(int64_t n, const double* chi, const double* gamma, const double* r1, const double* r2, const double* r3,
                            int classical, double* out3) {
    if (n < 0 || (n > 0 && (!chi || !gamma || !r1 || !r2 || !r3 || !out3))) return OPAL_EINVAL;
    if (n == 0) return OPAL_OK;
    const size_t b = (size_t)n * sizeof(double);
    double* d[5]; double* d_o = nullptr; double* d_t = nullptr; StatusDev* d_st = nullptr;
    const double* h[5] = {chi, gamma, r1, r2, r3};
    for (int k = 0; k < 5; k++) { if (cudaMalloc(&d[k], b) != cudaSuccess) return OPAL_ECUDA; cudaMemcpy(d[k], h[k], b, cudaMemcpyHostToDevice); }
    if (cudaMalloc(&d_o, 3 * b) != cudaSuccess || cudaMalloc(&d_t, sizeof(H_QED_TABLES)) != cudaSuccess || cudaMalloc(&d_st, sizeof(StatusDev)) != cudaSuccess) return OPAL_ECUDA;
    cudaMemcpy(d_t, H_QED_TABLES, sizeof(H_QED_TABLES), cudaMemcpyHostToDevice);
    cudaMemset(d_st, 0, sizeof(StatusDev));
    fast::table().qed_sample(0, (int)n, d_t, d[0], d[1], d[2], d[3], d[4], classical, d_o, d_st);
    cudaError_t e = cudaMemcpy(out3, d_o, 3 * b, cudaMemcpyDeviceToHost);
    StatusDev st;
    cudaMemcpy(&st, d_st, sizeof(st), cudaMemcpyDeviceToHost);
    for (int k = 0; k < 5; k++) cudaFree(d[k]);
    cudaFree(d_o); cudaFree(d_t); cudaFree(d_st);
    if (e != cudaSuccess) return OPAL_ECUDA;
    return (st.err & ERR_CHI) ? OPAL_ECHI_RANGE : OPAL_OK;
}

opal_status opal_bw_rate(int64_t n, const double* chi, double* out) {
    if (n < 0 || (n > 0 && (!chi || !out))) return OPAL_EINVAL;
    if (n == 0) return OPAL_OK;
    double *d_c = nullptr, *d_o = nullptr;
    const size_t b = (size_t)n * sizeof(double);
    if (cudaMalloc(&d_c, b) != cudaSuccess || cudaMalloc(&d_o, b) != cudaSuccess) return OPAL_ECUDA;
    cudaMemcpy(d_c, chi, b, cudaMemcpyHostToDevice);
    fast::table().bw_rate(0, (int)n, d_c, d_o);
    const cudaError_t e = cudaMemcpy(out, d_o, b, cudaMemcpyDeviceToHost);
    cudaFree(d_c); cudaFree(d_o);
    return e == cudaSuccess ? OPAL_OK : OPAL_ECUDA;
}

opal_status opal_bw_sample(int64_t n, const double* chi, const double* r1, const double* r2, double* out) {
    if (n < 0 || (n > 0 && (!chi || !r1 || !r2 || !out))) return OPAL_EINVAL;
    if (n == 0) return OPAL_OK;
    const size_t b = (size_t)n * sizeof(double);
    double* d[4] = {nullptr, nullptr, nullptr, nullptr};
    const double* h[3] = {chi, r1, r2};
    for (int k = 0; k < 4; k++) if (cudaMalloc(&d[k], b) != cudaSuccess) return OPAL_ECUDA;
    for (int k = 0; k < 3; k++) cudaMemcpy(d[k], h[k], b, cudaMemcpyHostToDevice);
    fast::table().bw_sample(0, (int)n, d[0], d[1], d[2], d[3]);
    const cudaError_t e = cudaMemcpy(out, d[3], b, cudaMemcpyDeviceToHost);
    for (int k = 0; k < 4; k++) cudaFree(d[k]);
    return e == cudaSuccess ? OPAL_OK : OPAL_ECUDA;
}

opal_status opal_pairs_created(opal_ctx* c, int64_t* n) {
    if (!c || !n) return OPAL_EINVAL;
    *n = c->pairs_created;
    return OPAL_OK;
}

opal_status opal_fields_at(opal_ctx* c, int64_t n, const int64_t* cell, const double* xi, double* out6) {
    CHECK_CTX(c);
    if (n <= 0) return OPAL_OK;
    long long* d_c = nullptr; double *d_x = nullptr, *d_o = nullptr;
    CU(cudaMalloc(&d_c, (size_t)n * sizeof(long long)));
    CU(cudaMalloc(&d_x, (size_t)n * sizeof(double)));
    CU(cudaMalloc(&d_o, (size_t)n * 6 * sizeof(double)));
    CU(cudaMemcpyAsync(d_c, cell, (size_t)n * sizeof(long long), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_x, xi, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    c->k.fields_at(c->stream, c->g, (int)n, d_c, d_x, d_o, c->d_status); c->launches++;
    CU(cudaMemcpyAsync(out6, d_o, (size_t)n * 6 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    fetch_status(c);
    cudaFree(d_c); cudaFree(d_x); cudaFree(d_o);
    return (opal_status)c->status;
}

}  // extern "C"
