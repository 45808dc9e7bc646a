// main.cpp -- the opal driver over the C ABI: src/main.rs:47-286 with every hot-path call replaced by its
// libopal_b200.so entry point (include/opal_b200.h).  The reference's driver is Rust; no Rust toolchain exists in
// this image, so the host side is C++ (INTEGRATION.md shows the Rust `extern "C"` block for the same calls).
//
//   opal_b200_driver input.yaml [--ranks N] [--device D] [--lean|--full] [--outputs K] [--steps K] [--seed S] [--dump] [--quiet]
//
// --ranks N    one process per GPU (the reference: mpirun -n N), x-slabs from GridDesign, NCCL ring inside the library
// --lean/--full  keep only (x, u, w) per particle -- the 80 B/particle-step layout; automatic when QED is off, deposition on and
//              no r/work/chi output is requested -- or insist on the reference's whole per-particle state
// --outputs K  stop after output K (for tests; default: run to `end`)
// --steps K    at most K steps between outputs (for tests: the unmodified example decks, shortened; default: the deck's)
// --seed S     seed the loader and the library with S + rank instead of the rank (main.rs:53 hard-wires the rank: one
//              realisation per deck; the statistical QED comparisons need several)
// --dump       also write {index}_{species}.bin at every output: the raw per-particle arrays (int64 n, then n doubles each of
//              cell, x, ux, uy, uz, weight, gamma, chi) -- for sample-level tests (KS on spectra), not a reference output
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "config.h"
#include "constants.h"
#include "grid_design.h"
#include "loader.h"
#include "output.h"
#include "ranks.h"

using namespace opal_driver;

namespace {

void ck(opal_status st, opal_ctx* ctx, const char* where) {
    if (st != OPAL_OK) {
        static const char* const names[] = {"OPAL_OK", "OPAL_EINVAL", "OPAL_ECUDA", "OPAL_ENCCL", "OPAL_ENOMEM", "OPAL_EASSERT_CFL",
                                            "OPAL_EASSERT_INDEX", "OPAL_ECHI_RANGE", "OPAL_ECAPACITY", "OPAL_EASSERT_PROB"};
        const char* name = (int)st >= 0 && (int)st < 10 ? names[(int)st] : "?";
        const char* msg = ctx ? OPAL_FN(last_error)(ctx) : (st == OPAL_ECUDA ? "no usable CUDA device (the library has no CPU fallback)" : "");
        throw std::runtime_error(std::string(where) + " failed (status " + std::to_string((int)st) + " " + name + "): " + (msg ? msg : ""));
    }
}

template <class F> bool optional_read(F f) {      // `.ok()` / `.unwrap_or()` of main.rs:72-82
    try { f(); return true; } catch (const ConfigError&) { return false; }
}

struct Options {
    std::string input;
    int ranks = 1, device = -1;
    bool lean = false, full = false, quiet = false, dump = false;
    long outputs = -1, steps = -1;
    unsigned long long seed = 0;
};

void dump_species(opal_ctx* ctx, int species, const char* name, const std::string& dir, size_t index, int rank) {
    int64_t n = 0;
    ck(OPAL_FN(particles_count)(ctx, species, &n), ctx, "particles_count");
    std::vector<int64_t> cell((size_t)n);
    std::vector<double> f[7];
    for (auto& v : f) v.resize((size_t)n);
    if (n > 0)
        ck(OPAL_FN(particles_download)(ctx, species, n, cell.data(), f[0].data(), nullptr, f[1].data(), f[2].data(), f[3].data(), f[4].data(),
                                       f[5].data(), f[6].data(), nullptr, nullptr, nullptr, nullptr, nullptr), ctx, "particles_download");
    const std::string path = dir + "/" + std::to_string(index) + "_" + name + (rank > 0 ? ".r" + std::to_string(rank) : std::string()) + ".bin";
    FILE* fp = std::fopen(path.c_str(), "wb");
    if (!fp) throw std::runtime_error("cannot write " + path);
    std::fwrite(&n, sizeof(n), 1, fp);
    std::vector<double> c((size_t)n);
    for (int64_t i = 0; i < n; i++) c[(size_t)i] = (double)cell[(size_t)i];
    std::fwrite(c.data(), sizeof(double), (size_t)n, fp);
    for (auto& v : f) std::fwrite(v.data(), sizeof(double), (size_t)n, fp);
    std::fclose(fp);
}

int run(const Options& opt, Ranks& ranks) {
    const int id = ranks.rank(), np = ranks.nranks();
    Xoshiro256StarStar rng((uint64_t)id + opt.seed);                                // main.rs:53 (--seed 0: the reference's choice)

    std::string output_dir = ".";                                                   // main.rs:61 (an empty parent means the cwd here)
    {
        const size_t slash = opt.input.find_last_of('/');
        if (slash != std::string::npos) output_dir = slash == 0 ? "/" : opt.input.substr(0, slash);
    }

    Config input = Config::from_file(opt.input);                                    // main.rs:65-66
    input.with_context("constants");

    const size_t nx = input.read_usize("control", "nx");                           // main.rs:68-76
    const double xmin = input.read_f64("control", "xmin");
    const double dx = input.read_f64("control", "dx");
    const double dt = 0.95 * dx / constants::SPEED_OF_LIGHT;
    const double tstart = input.read_f64("control", "start");
    const double tend = input.read_f64("control", "end");
    const bool current_deposition = input.read_bool("control", "current_deposition");
    const size_t output_frequency = input.read_usize("control", "n_outputs");
    bool balance = true;
    optional_read([&] { balance = input.read_bool("control", "balance"); });

    const bool photon_emission = input.read_bool("qed", "photon_emission");         // main.rs:78-79
    const bool photon_absorption = input.read_bool("qed", "photon_absorption");
    // EXTENSION (not in the reference's schema; absent == false): qed.pair_creation -- nonlinear Breit-Wheeler pair creation,
    // OPAL_FLAG_PAIR_CREATION of include/opal_b200.h.  Positrons take the ion slot and are written out as species "positron"
    // with the electrons' output list; a deck with ions cannot switch it on.
    bool pair_creation = false;
    optional_read([&] { pair_creation = input.read_bool("qed", "pair_creation"); });
    double photon_energy_min = 0.0, photon_angle_max = 0.0, max_formation_length = 0.0;
    const bool has_emin = optional_read([&] { photon_energy_min = 1.0e-6 * input.read_f64("qed", "photon_energy_min") / constants::ELEMENTARY_CHARGE; });
    const bool has_amax = optional_read([&] { photon_angle_max = input.read_f64("qed", "photon_angle_max"); });
    const bool has_fl = optional_read([&] { max_formation_length = input.read_f64("qed", "max_formation_length"); });
    double disable_qed_after = 0.0, disable_absorption_after = 0.0;                 // main.rs:84-85
    const bool has_dqa = optional_read([&] { disable_qed_after = input.read_f64("qed", "disable_qed_after"); });
    const bool has_daa = optional_read([&] { disable_absorption_after = input.read_f64("qed", "disable_absorption_after"); });

    // laser (main.rs:88-104): both components or a config error; no laser section => periodic
    std::function<double(double, double)> laser_y, laser_z;
    int left_bdy = OPAL_BDY_INTERNAL, right_bdy = OPAL_BDY_INTERNAL;
    if (input.contains("laser")) {
        laser_y = input.func2("laser", "Ey", "t", "x");
        laser_z = input.func2("laser", "Ez", "t", "x");
        left_bdy = OPAL_BDY_LASER;
        right_bdy = OPAL_BDY_ABSORBING;
    }

    const size_t min_size = 2 * OPAL_GHOST_SIZE;                                    // YeeGrid::min_size, yee.rs:782-784
    GridDesign design = balance ? GridDesign::balanced(np, nx, xmin, dx, min_size, input.func("electrons", "ne", "x"))   // main.rs:108-113
                                : GridDesign::unbalanced(np, nx, xmin, dx, min_size);
    const size_t my_nx = design.nx[(size_t)id];
    const double my_xmin = xmin + design.offset[(size_t)id];

    // ---- the slab context (YeeGrid::build, yee.rs:272-319) ----
    // The lean (x, u, w) layout -- the fused fast kernel -- is valid when nothing observable needs the rest of the
    // particle state: QED off, deposition on, and no r / work / chi distribution requested.  It is chosen automatically
    // then (--full keeps the reference's whole per-particle state, --lean insists).
    bool lean_ok = !photon_emission && !photon_absorption && current_deposition;
    for (const char* sec : {"electrons", "ions"}) {
        std::vector<std::string> outs;
        optional_read([&] { outs = input.read_strings(sec, "output"); });
        for (const std::string& o : outs) {
            const OutputSpec spec = OutputSpec::parse(o);
            for (int q : spec.quantity) if (q == OPAL_Q_R || q == OPAL_Q_WORK || q == OPAL_Q_CHI) lean_ok = false;
        }
    }
    if (opt.lean && !lean_ok) throw std::runtime_error("--lean: this deck needs the full particle state (QED on, deposition off, or r/work/chi outputs)");
    const bool lean = opt.lean || (lean_ok && !opt.full);
    if (pair_creation && !(photon_emission || photon_absorption)) throw std::runtime_error("qed.pair_creation needs photons (qed.photon_emission or qed.photon_absorption)");
    const uint32_t flags = (lean ? OPAL_FLAG_LEAN_STATE : 0u) | (pair_creation ? OPAL_FLAG_PAIR_CREATION : 0u);
    opal_ctx* ctx = nullptr;
    const int device = opt.device >= 0 ? opt.device + id : id;
    ck(OPAL_FN(create)(&ctx, device, (int64_t)my_nx, my_xmin, dx, dt, left_bdy, right_bdy, id, np, (uint64_t)id + opt.seed, flags), nullptr, "opal_create");
    ranks.connect(ctx);                                                             // NCCL ring (np > 1)
    if (pair_creation && np > 1 && ranks.host_staged()) throw std::runtime_error("qed.pair_creation with several ranks needs the library's NCCL exchange");
    const bool staged = ranks.host_staged();                                        // no communicator in the library (ranks.h)
    if (staged && lean) {                                                           // a lean context takes opal_run_steps only
        OPAL_FN(destroy)(ctx);
        ctx = nullptr;
        ck(OPAL_FN(create)(&ctx, device, (int64_t)my_nx, my_xmin, dx, dt, left_bdy, right_bdy, id, np, (uint64_t)id + opt.seed, 0u), nullptr, "opal_create");
    }

    // ---- particle initialisation (main.rs:117-168) ----
    auto upload = [&](int species, const ParticleArrays& p) {
        ck(OPAL_FN(particles_upload)(ctx, species, (int64_t)p.size(), p.cell.data(), p.x.data(), p.ux.data(), p.uy.data(), p.uz.data(),
                                     p.weight.data(), p.tau.empty() ? nullptr : p.tau.data()), ctx, "particles_upload");
    };
    const size_t epc = input.read_usize("electrons", "npc");
    const std::vector<std::string> eospec = input.read_strings("electrons", "output");
    if (epc > 0) {
        upload(OPAL_ELECTRON, load_population(epc, input.func("electrons", "ne", "x"), input.func3("electrons", "ux", "x", "urand", "nrand"),
                                              input.func3("electrons", "uy", "x", "urand", "nrand"), input.func3("electrons", "uz", "x", "urand", "nrand"),
                                              (int64_t)my_nx, my_xmin, dx, rng, true));
    }
    const size_t ipc = input.read_usize("ions", "npc");
    std::vector<std::string> iospec;
    std::string ion_name;
    if (ipc > 0 && pair_creation) throw std::runtime_error("qed.pair_creation keeps the positrons in the ion slot: set ions.npc to 0");
    if (ipc > 0) {
        iospec = input.read_strings("ions", "output");
        ion_name = input.read_string("ions", "name");
        const double ion_charge = input.read_f64("ions", "Z");
        const double ion_mass = input.read_f64("ions", "A");
        ck(OPAL_FN(set_ion)(ctx, ion_charge, ion_mass), ctx, "set_ion");            // with_charge_state, main.rs:145-147
        upload(OPAL_ION, load_population(ipc, input.func("ions", "ni", "x"), input.func3("ions", "ux", "x", "urand", "nrand"),
                                         input.func3("ions", "uy", "x", "urand", "nrand"), input.func3("ions", "uz", "x", "urand", "nrand"),
                                         (int64_t)my_nx, my_xmin, dx, rng, false));
    }
    std::vector<std::string> pospec;
    bool photons_on = false;
    if (photon_emission || photon_absorption) {                                     // main.rs:154-170
        const size_t ppc = input.read_usize("photons", "npc");
        pospec = input.read_strings("photons", "output");
        photons_on = true;
        ParticleArrays p;
        if (ppc > 0) {
            p = load_population(ppc, input.func("photons", "nph", "x"), input.func3("photons", "ux", "x", "urand", "nrand"),
                                input.func3("photons", "uy", "x", "urand", "nrand"), input.func3("photons", "uz", "x", "urand", "nrand"),
                                (int64_t)my_nx, my_xmin, dx, rng, true);
        }
        if (ppc > 0 || np > 1) upload(OPAL_PHOTON, p);
    }

    // main.rs:246-248: absorb(..., disable_qed_after, disable_absorption_after) -- the first lands in absorb's
    // max_displacement parameter, the second in stop_time (interactions.rs:145); kept as the reference has it
    if (photon_absorption)
        ck(OPAL_FN(set_absorption)(ctx, 1, has_dqa ? &disable_qed_after : nullptr, has_daa ? &disable_absorption_after : nullptr), ctx, "set_absorption");

    // ---- time loop (main.rs:184-267) ----
    double t = tstart;
    const size_t total_steps = (size_t)((tend - tstart) / dt);
    size_t steps_bt_output = total_steps / output_frequency;
    if (opt.steps >= 0 && (size_t)opt.steps < steps_bt_output) steps_bt_output = (size_t)opt.steps;
    double x2 = 0.0;
    ck(OPAL_FN(laser_x)(ctx, &x2), ctx, "laser_x");
    if (id == 0 && !opt.quiet)
#ifdef OPAL_ORACLE_BUILD          // oracle/opal_ref_driver: the same driver linked over the CPU restatement (test infrastructure)
        std::printf("Running %d task%s on the CPU oracle (libopal_oracle: one x-slab per task)...\n", np, np > 1 ? "s" : "");
#else
        std::printf("Running %d task%s on %d GPU%s (libopal_b200: one x-slab per GPU)...\n", np, np > 1 ? "s" : "", np, np > 1 ? "s" : "");
#endif

    double output_seconds = 0.0, grid_seconds = 0.0;
    auto write_all = [&](size_t index) {                                             // main.rs:218-222
        const auto w0 = std::chrono::steady_clock::now();
        write_grid(ctx, output_dir, index, ranks);
        grid_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count();
        write_population(ctx, OPAL_ELECTRON, "electron", eospec, output_dir, index, ranks);
        if (ipc > 0) write_population(ctx, OPAL_ION, ion_name, iospec, output_dir, index, ranks);
        if (pair_creation) write_population(ctx, OPAL_ION, "positron", eospec, output_dir, index, ranks);
        if (photons_on) write_population(ctx, OPAL_PHOTON, "photon", pospec, output_dir, index, ranks);
        write_energies(ctx, output_dir, index, ranks);
        if (opt.dump) {
            dump_species(ctx, OPAL_ELECTRON, "electron", output_dir, index, id);
            if (ipc > 0) dump_species(ctx, OPAL_ION, ion_name.c_str(), output_dir, index, id);
            if (pair_creation) dump_species(ctx, OPAL_ION, "positron", output_dir, index, id);
            if (photons_on) dump_species(ctx, OPAL_PHOTON, "photon", output_dir, index, id);
        }
        output_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count();
    };

    double filters[3];
    const double nan = std::nan("");
    filters[0] = has_emin ? photon_energy_min : nan;
    filters[1] = has_amax ? photon_angle_max : nan;
    filters[2] = has_fl ? max_formation_length : nan;

    // One iteration of main.rs:238-267 call by call, the driver moving leavers (mod.rs:296-373) and halo strips
    // (yee.rs:353-409) between the slabs itself: the path of a host that keeps its own p2p (host_staged(), ranks.h).
    const bool has_side[2] = {!(id == 0 && left_bdy != OPAL_BDY_INTERNAL), !(id == np - 1 && right_bdy != OPAL_BDY_INTERNAL)};   // yee.rs:329-343
    auto staged_advance = [&](int species) {
        ck(OPAL_FN(advance)(ctx, species), ctx, "advance");
        std::vector<char> out[2], in[2];
        for (int side = 0; side < 2; side++) {
            if (!has_side[side]) continue;
            int64_t cap = 4096, n = 0;
            for (;;) {
                out[side].resize((size_t)cap * sizeof(opal_migrant));
                const opal_status st = OPAL_FN(migrants_take)(ctx, species, side, cap, &n, reinterpret_cast<opal_migrant*>(out[side].data()));
                if (st == OPAL_ECAPACITY && n > cap) { cap = n; continue; }
                ck(st, ctx, "migrants_take");
                break;
            }
            out[side].resize((size_t)n * sizeof(opal_migrant));
        }
        ranks.ring_exchange(out[OPAL_LEFT], out[OPAL_RIGHT], in[OPAL_LEFT], in[OPAL_RIGHT]);
        for (int side = 0; side < 2; side++)
            if (has_side[side] && !in[side].empty())
                ck(OPAL_FN(migrants_give)(ctx, species, side, (int64_t)(in[side].size() / sizeof(opal_migrant)),
                                          reinterpret_cast<const opal_migrant*>(in[side].data())), ctx, "migrants_give");
    };
    auto staged_step = [&](double tj, double lyj, double lzj) {
        if (epc > 0) staged_advance(OPAL_ELECTRON);
        if (ipc > 0) staged_advance(OPAL_ION);
        if (photons_on) staged_advance(OPAL_PHOTON);
        if (photon_absorption)
            ck(OPAL_FN(absorb)(ctx, tj, has_dqa ? &disable_qed_after : nullptr, has_daa ? &disable_absorption_after : nullptr), ctx, "absorb");
        if (photon_emission)
            ck(OPAL_FN(emit_radiation)(ctx, tj, has_emin ? &photon_energy_min : nullptr, has_amax ? &photon_angle_max : nullptr,
                                       has_fl ? &max_formation_length : nullptr), ctx, "emit_radiation");
        if (current_deposition) {
            ck(OPAL_FN(clear)(ctx), ctx, "clear");
            if (epc > 0) ck(OPAL_FN(deposit)(ctx, OPAL_ELECTRON), ctx, "deposit");
            if (ipc > 0) ck(OPAL_FN(deposit)(ctx, OPAL_ION), ctx, "deposit");
        }
        std::vector<char> out[2], in[2];
        for (int side = 0; side < 2; side++) {
            if (!has_side[side]) continue;
            out[side].resize(8 * 11 * sizeof(double));
            ck(OPAL_FN(halo_take)(ctx, side, reinterpret_cast<double*>(out[side].data())), ctx, "halo_take");
        }
        ranks.ring_exchange(out[OPAL_LEFT], out[OPAL_RIGHT], in[OPAL_LEFT], in[OPAL_RIGHT]);
        for (int side = 0; side < 2; side++)
            if (has_side[side] && in[side].size() == 8 * 11 * sizeof(double))
                ck(OPAL_FN(halo_give)(ctx, side, reinterpret_cast<const double*>(in[side].data())), ctx, "halo_give");
        ck(OPAL_FN(synchronize)(ctx, tj, lyj, lzj), ctx, "synchronize");
        ck(OPAL_FN(field_advance)(ctx), ctx, "field_advance");
    };

    const auto wall0 = std::chrono::steady_clock::now();
    double loop_seconds = 0.0;
    int64_t particle_steps = 0;
    std::vector<double> ly(steps_bt_output, 0.0), lz(steps_bt_output, 0.0);
    size_t i = 0;
    for (; i < output_frequency; i++) {
        write_all(i);
        if (id == 0 && !opt.quiet) {
            // the reference's progress lines (main.rs:224-233): SimulationTime, PrettyDuration and ettc of setup.rs:374-437
            const double rt = std::chrono::duration<double>(std::chrono::steady_clock::now() - wall0).count();
            if (i > 0) {
                const double ettc = rt * (double)(output_frequency - i) / (double)i;   // (total - current) / current in steps
                std::printf("Output %4zu at t = %s, RT = %s, ETTC = %s...\n", i, fmt_simulation_time(t).c_str(), fmt_pretty_duration(rt).c_str(),
                            fmt_pretty_duration(ettc).c_str());
            } else {
                std::printf("Output %4zu at t = %s...\n", i, fmt_simulation_time(t).c_str());
            }
            std::fflush(stdout);
        }
        if (opt.outputs >= 0 && (long)i >= opt.outputs) break;
        // the laser amplitudes of the next steps_bt_output steps, at the times main.rs:266 accumulates
        const double t0 = t;
        for (size_t j = 0; j < steps_bt_output; j++) {
            if (left_bdy == OPAL_BDY_LASER && id == 0) { ly[j] = laser_y(t, x2); lz[j] = laser_z(t, x2); }   // yee.rs:456-462
            t += dt;
        }
        int64_t n_e = 0, n_i = 0;
        OPAL_FN(particles_count)(ctx, OPAL_ELECTRON, &n_e);
        OPAL_FN(particles_count)(ctx, OPAL_ION, &n_i);
        const auto c0 = std::chrono::steady_clock::now();
        if (staged) {
            double tj = t0;
            for (size_t j = 0; j < steps_bt_output; j++, tj += dt) staged_step(tj, ly[j], lz[j]);
        } else
        ck(OPAL_FN(run_steps)(ctx, (int64_t)steps_bt_output, t0, ly.data(), lz.data(), current_deposition ? 1 : 0, photon_emission ? 1 : 0, filters),
           ctx, "run_steps");                                                        // main.rs:238-267, steps_bt_output times
        const double interval_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - c0).count();
        loop_seconds += interval_s;
        if (id == 0 && getenv("OPAL_STATS")) {                                        // instrumentation: where a run spends its time
            std::fprintf(stderr, "interval %zu: %zu steps in %.1f ms = %.1f us/step", i, steps_bt_output, 1e3 * interval_s, 1e6 * interval_s / (double)steps_bt_output);
            if (atoi(getenv("OPAL_STATS")) >= 2) {                                    // per-kernel device time (the library then runs step by step)
                double ms[3]; int64_t nl[3];
                if (OPAL_FN(kernel_times)(ctx, ms, nl) == OPAL_OK && nl[0] > 0)
                    std::fprintf(stderr, "; kernels per step: push %.1f us, scatter %.1f us, grid %.1f us", 1e3 * ms[0] / (double)steps_bt_output,
                                 1e3 * ms[1] / (double)steps_bt_output, 1e3 * ms[2] / (double)steps_bt_output);
                OPAL_FN(timing)(ctx, 1);                                              // (restarts the sums: the next interval's)
            }
            std::fprintf(stderr, "\n");
        }
        particle_steps += (n_e + n_i) * (int64_t)steps_bt_output;
    }
    if (i == output_frequency) {
        write_all(output_frequency);                                                 // main.rs:271-276
        if (id == 0 && !opt.quiet) {
            const double rt = std::chrono::duration<double>(std::chrono::steady_clock::now() - wall0).count();
            std::printf("Output %4zu at t = %s, RT = %s\n", output_frequency, fmt_simulation_time(t).c_str(), fmt_pretty_duration(rt).c_str());   // main.rs:278-282
        }
    }
    double ps[2] = {(double)particle_steps, 0.0};
    ranks.sum(ps, 1);
    if (id == 0 && !opt.quiet && loop_seconds > 0.0)
        std::printf("time loop: %.3f s, %.4e charged-particle-steps/s\n", loop_seconds, ps[0] / loop_seconds);
    if (id == 0 && getenv("OPAL_STATS"))
        std::fprintf(stderr, "outputs: %.3f s (grid files %.3f s), whole run since the first output: %.3f s\n", output_seconds, grid_seconds, std::chrono::duration<double>(std::chrono::steady_clock::now() - wall0).count());
    OPAL_FN(destroy)(ctx);
    return 0;
}

}  // namespace

int main(int argc, char** argv) {
    Options opt;
    for (int a = 1; a < argc; a++) {
        const std::string s = argv[a];
        if (s == "--ranks" && a + 1 < argc) opt.ranks = std::atoi(argv[++a]);
        else if (s == "--device" && a + 1 < argc) opt.device = std::atoi(argv[++a]);
        else if (s == "--outputs" && a + 1 < argc) opt.outputs = std::atol(argv[++a]);
        else if (s == "--steps" && a + 1 < argc) opt.steps = std::atol(argv[++a]);
        else if (s == "--seed" && a + 1 < argc) opt.seed = std::strtoull(argv[++a], nullptr, 10);
        else if (s == "--dump") opt.dump = true;
        else if (s == "--lean") opt.lean = true;
        else if (s == "--full") opt.full = true;
        else if (s == "--quiet") opt.quiet = true;
        else if (!s.empty() && s[0] != '-' && opt.input.empty()) opt.input = s;
        else { std::fprintf(stderr, "unknown argument %s\n", s.c_str()); return 2; }
    }
    if (opt.input.empty()) {                                                         // main.rs:57-59
        std::fprintf(stderr, "%s\n", ConfigError{ConfigErrorKind::MissingFile, "", ""}.message().c_str());
        return 1;
    }
    if (opt.ranks < 1) opt.ranks = 1;
    return launch_ranks(opt.ranks, [&](Ranks& ranks) -> int {
        try {
            return run(opt, ranks);
        } catch (const ConfigError& e) {
            std::fprintf(stderr, "Error: %s\n", e.message().c_str());
        } catch (const std::exception& e) {
            std::fprintf(stderr, "Error: %s\n", e.what());
        }
        return 1;
    });
}
