// output.h -- what the reference writes every `steps_bt_output` steps (main.rs:218-222, 272-276):
//   {dir}/{index}_grid.dat                       YeeGrid::write_data     yee.rs:749-780 (+ interpolate :818-837)
//   {dir}/{index}_{species}_{axes}[...].fits     Population::write_data  particle/mod.rs:383-568, hgram.rs
//   {dir}/{index}_energy.dat                     write_energies          main.rs:23-42
// The histograms are built on the device (opal_hgram_minmax / opal_hgram_bin); this file holds the host-side
// bookkeeping of Histogram::generate_{1d,2d} and the FITS primary-HDU writer (the reference links CFITSIO through
// the `fitsio 0.15.0` crate, which is not vendored: the layout follows the FITS standard and the keys of hgram.rs:404-419).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/opal_b200.h"

namespace opal_driver {

// Rust's float formatting, used by every text output of the reference
std::string fmt_lower_exp(double v);              // {:e}   shortest round-trip digits, e.g. 1.5e-7, 0e0
void append_lower_exp(std::string& out, double v);  // the same, appended
std::string fmt_lower_exp6(double v);             // {:.6e} e.g. 1.200135e-6
std::string fmt_display(double v);                // {}     shortest round-trip digits, positional, e.g. 0.00015
std::string fmt_simulation_time(double seconds);  // SimulationTime (setup.rs:417-437): "{: >8.2} unit", nearest SI prefix from as to s
std::string fmt_pretty_duration(double seconds);  // PrettyDuration (setup.rs:400-414): [d"d "]hh:mm:ss of the whole seconds

// sums / minima / maxima over the slabs of a run (the reference: MPI all_reduce); one slab => identity
struct Reducer {
    virtual ~Reducer() {}
    virtual void min(double* v, int n) {}
    virtual void max(double* v, int n) {}
    virtual void sum(double* v, size_t n) {}
    virtual void sum_i64(int64_t* v, int n) {}
    virtual int rank() const { return 0; }
    virtual int nranks() const { return 1; }
    // concatenation of per-slab arrays on rank 0, in rank order (YeeGrid::write_data's receive loop, yee.rs:757-760)
    virtual std::vector<double> gather(const std::vector<double>& mine) { return mine; }
};

enum class BinSpecKind { Automatic, LogScaled, FixedNumber, FixedSize };     // hgram.rs:9-14
struct BinSpec {
    BinSpecKind kind = BinSpecKind::Automatic;
    size_t n = 0;
    double dx = 0.0;
    static BinSpec from(const std::string& s);                               // hgram.rs:27-41
};

struct Histogram {                                                            // hgram.rs:69-83
    size_t dim = 0;
    double total = 0.0, bin_vol = 0.0;
    std::vector<double> min, max, cts, bin_sz;
    std::vector<size_t> bins;
    std::string name, bunit;
    std::vector<std::string> axis, unit;
    void write_fits(const std::string& filename) const;                       // hgram.rs:394-426; a leading '!' means overwrite
};

// one entry of a species' `output:` list, parsed as mod.rs:451-500.  valid == false: the reference silently skips it
struct OutputSpec {
    bool valid = false;
    std::vector<std::string> axes;          // 1 or 2 names
    std::vector<int> quantity;              // opal_quantity per axis
    std::vector<std::string> units;
    BinSpec bspec;
    std::string weight;                     // "weight" | "auto" | "energy"
    int weight_kind = OPAL_W_WEIGHT;
    static OutputSpec parse(const std::string& o);
    std::string filename(const std::string& dir, size_t index, const std::string& species_name) const;   // mod.rs:512-519, 531-538
};

// Histogram::generate_1d / generate_2d (hgram.rs:167-392) with the two particle passes done by the library.
// Returns false when the species has no particles anywhere (hgram.rs:206-208: None).
bool generate_histogram(opal_ctx* ctx, int species, const OutputSpec& spec, Reducer& red, Histogram* out);

// Population::write_data for one species
void write_population(opal_ctx* ctx, int species, const std::string& species_name, const std::vector<std::string>& outputs,
                      const std::string& dir, size_t index, Reducer& red);
// YeeGrid::write_data
void write_grid(opal_ctx* ctx, const std::string& dir, size_t index, Reducer& red);
// write_energies
void write_energies(opal_ctx* ctx, const std::string& dir, size_t index, Reducer& red);

}  // namespace opal_driver
