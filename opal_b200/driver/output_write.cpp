// output_write.cpp -- Histogram::generate_{1d,2d}, Population::write_data, YeeGrid::write_data, write_energies
// (see output.h).  All particle passes run inside the library (device-side in libopal_b200.so).
#include <cfloat>
#include <cmath>
#include <fstream>
#include <stdexcept>

#include "output.h"

namespace opal_driver {

namespace {

void ck(opal_status st, opal_ctx* ctx, const char* where) {
    if (st != OPAL_OK) {
        const char* msg = OPAL_FN(last_error)(ctx);
        throw std::runtime_error(std::string(where) + " failed (status " + std::to_string((int)st) + "): " + (msg ? msg : ""));
    }
}

size_t number_of_bins(double min, double max, size_t n, const BinSpec& b) {       // hgram.rs:130-143
    if (min == max) return 1;
    switch (b.kind) {
        case BinSpecKind::FixedNumber: return b.n;
        case BinSpecKind::FixedSize: {
            const double v = std::ceil((max - min) / b.dx);
            return v > 0.0 ? (size_t)v : 0;
        }
        default: {
            const double v = std::ceil(2.0 * std::cbrt((double)n));
            return v > 0.0 ? (size_t)v : 0;
        }
    }
}

}  // namespace

bool generate_histogram(opal_ctx* ctx, int species, const OutputSpec& spec, Reducer& red, Histogram* h) {
    const int dim = (int)spec.axes.size();
    const bool log_axis = spec.bspec.kind == BinSpecKind::LogScaled;
    int q[2] = {0, 0}, lg[2] = {log_axis ? 1 : 0, log_axis ? 1 : 0};
    for (int d = 0; d < dim; d++) q[d] = spec.quantity[d];

    // local, then global min / max / count (hgram.rs:186-204, 290-316)
    double mn[2] = {DBL_MAX, DBL_MAX}, mx[2] = {-DBL_MAX, -DBL_MAX};
    int64_t num = 0;
    ck(OPAL_FN(hgram_minmax)(ctx, species, dim, q, lg, mn, mx, &num), ctx, "hgram_minmax");
    red.min(mn, dim);
    red.max(mx, dim);
    red.sum_i64(&num, 1);
    if (num == 0) return false;                                                    // hgram.rs:206-208

    int64_t nbins[2] = {1, 1};
    double bin_sz[2] = {0.0, 0.0};       // what the binning pass divides by
    double volume = 1.0;
    h->bin_sz.assign((size_t)dim, 0.0);
    for (int d = 0; d < dim; d++) {
        nbins[d] = (int64_t)number_of_bins(mn[d], mx[d], (size_t)num, spec.bspec);
        if (nbins[d] < 1) throw std::runtime_error("histogram with zero bins requested");
        if (mn[d] == mx[d]) {
            bin_sz[d] = 0.0;                                                       // hgram.rs:149-151
        } else {
            bin_sz[d] = spec.bspec.kind == BinSpecKind::FixedSize ? spec.bspec.dx : (mx[d] - mn[d]) / (double)nbins[d];   // hgram.rs:153-158
            volume *= bin_sz[d];
        }
        h->bin_sz[(size_t)d] = bin_sz[d];
    }
    if (dim == 1) {
        // generate_1d divides by bin_vol, which is 1.0 for a degenerate axis (hgram.rs:209-220), and reports
        // bin_sz = 0 when there is a single bin (hgram.rs:275)
        if (mn[0] == mx[0]) bin_sz[0] = 1.0;
        h->bin_sz[0] = nbins[0] <= 1 ? 0.0 : bin_sz[0];
    }

    const size_t len = (size_t)nbins[0] * (size_t)(dim == 2 ? nbins[1] : 1);
    h->cts.assign(len, 0.0);
    double total = 0.0;
    ck(OPAL_FN(hgram_bin)(ctx, species, dim, q, lg, mn, bin_sz, nbins, spec.weight_kind, /*HeightSpec::Density*/ 1, h->cts.data(), &total), ctx,
       "hgram_bin");
    red.sum(&total, 1);                                                            // hgram.rs:261, 369
    for (double& c : h->cts) c = c / volume;                                       // HeightSpec::Density, hgram.rs:266, 374
    red.sum(h->cts.data(), len);                                                   // hgram.rs:271, 379

    h->dim = (size_t)dim;
    h->total = total;
    h->bin_vol = volume;
    h->min.assign(mn, mn + dim);
    h->max.assign(mx, mx + dim);
    h->bins.clear();
    for (int d = 0; d < dim; d++) h->bins.push_back((size_t)nbins[d]);
    h->axis = spec.axes;
    h->unit = spec.units;
    if (dim == 1) {
        h->name = "hgram/density/" + spec.axes[0];                                  // hgram.rs:276-277
        h->bunit = "1/" + spec.units[0];
    } else {
        h->name = "hgram/density/" + spec.axes[0] + "_" + spec.axes[1];             // hgram.rs:387-388
        h->bunit = "1/(" + spec.units[0] + "." + spec.units[1] + ")";
    }
    return true;
}

void write_population(opal_ctx* ctx, int species, const std::string& species_name, const std::vector<std::string>& outputs,
                      const std::string& dir, size_t index, Reducer& red) {
    for (const std::string& o : outputs) {                                          // mod.rs:451
        const OutputSpec spec = OutputSpec::parse(o);
        if (!spec.valid) continue;                                                  // mod.rs:502, 564
        Histogram h;
        const bool some = generate_histogram(ctx, species, spec, red, &h);
        if (red.rank() == 0 && some) h.write_fits(spec.filename(dir, index, species_name));   // mod.rs:557-562
    }
}

void write_grid(opal_ctx* ctx, const std::string& dir, size_t index, Reducer& red) {
    int64_t total = 0, left = 0, right = 0;
    ck(OPAL_FN(grid_shape)(ctx, &total, &left, &right), ctx, "grid_shape");
    std::vector<double> cells((size_t)total * OPAL_CELL_DOUBLES);
    ck(OPAL_FN(fields_download)(ctx, cells.data()), ctx, "fields_download");
    // YeeGrid::interpolate (yee.rs:818-837): ghost cells stripped, staggered components averaged to the cell's left edge
    const size_t first = (size_t)(left - 1), last = (size_t)(total - right);       // slice [left-1, len-right)
    std::vector<double> mine;
    mine.reserve((last - first - 1) * OPAL_CELL_DOUBLES);
    for (size_t i = first; i + 1 < last; i++) {
        const double* a = &cells[i * OPAL_CELL_DOUBLES];
        const double* b = a + OPAL_CELL_DOUBLES;                                    // {x, rho, j[3], E[3], B[3]}
        const double out[OPAL_CELL_DOUBLES] = {b[0], b[1], 0.5 * (a[2] + b[2]), b[3], b[4], 0.5 * (a[5] + b[5]), b[6], b[7],
                                               b[8], 0.5 * (a[9] + b[9]), 0.5 * (a[10] + b[10])};
        mine.insert(mine.end(), out, out + OPAL_CELL_DOUBLES);
    }
    const std::vector<double> global = red.gather(mine);                            // yee.rs:755-760
    if (red.rank() != 0) return;
    std::ofstream f(dir + "/" + std::to_string(index) + "_grid.dat", std::ios::trunc);
    if (!f) throw std::runtime_error("cannot create grid output in " + dir);
    std::string text;                                                               // the whole file, written once
    text.reserve(global.size() * 12);
    for (size_t i = 0; i + OPAL_CELL_DOUBLES <= global.size(); i += OPAL_CELL_DOUBLES) {   // yee.rs:765-775: ten {:e} and a final {}
        for (int k = 0; k < 10; k++) { append_lower_exp(text, global[i + k]); text += ' '; }
        text += fmt_display(global[i + 10]);
        text += '\n';
    }
    f.write(text.data(), (std::streamsize)text.size());
}

void write_energies(opal_ctx* ctx, const std::string& dir, size_t index, Reducer& red) {
    double e[4] = {0.0, 0.0, 0.0, 0.0};
    ck(OPAL_FN(energy)(ctx, &e[0], &e[1], &e[2], &e[3]), ctx, "energy");
    red.sum(e, 4);                                                                  // reduce_into_root, yee.rs:801-808, mod.rs:232-239
    if (red.rank() != 0) return;
    std::ofstream f(dir + "/" + std::to_string(index) + "_energy.dat", std::ios::trunc);   // main.rs:32-38
    if (!f) throw std::runtime_error("cannot create energy output in " + dir);
    f << "em_field " << fmt_lower_exp6(e[0]) << "\n";
    f << "electrons " << fmt_lower_exp6(e[1]) << "\n";
    f << "ions " << fmt_lower_exp6(e[2]) << "\n";
    f << "photons " << fmt_lower_exp6(e[3]) << "\n";
}

}  // namespace opal_driver
