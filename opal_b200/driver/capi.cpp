// capi.cpp -- C-callable access to the driver's host-side logic (config reader, expression evaluator, Rust-style
// number formatting, output-spec parsing, particle loader) so the CPU test-suite can exercise it through ctypes.
#include <cstring>
#include <string>

#include "config.h"
#include "grid_design.h"
#include "loader.h"
#include "output.h"

using namespace opal_driver;

namespace {
int copy_out(const std::string& s, char* buf, int cap) {
    if (!buf || cap <= 0) return (int)s.size();
    const int n = (int)std::min<size_t>(s.size(), (size_t)cap - 1);
    std::memcpy(buf, s.data(), (size_t)n);
    buf[n] = 0;
    return (int)s.size();
}
// 0 ok, 1 MissingFile, 2 MissingSection, 3 MissingField, 4 ConversionFailure
int code(const ConfigError& e) { return 1 + (int)e.kind; }
}  // namespace

extern "C" {

int opal_drv_expr_eval(const char* text, double* out) {
    try {
        Context ctx;
        *out = Expr::parse(text).eval(ctx);
        return 0;
    } catch (const ExprError&) {
        return 1;
    }
}

// kind: 0 f64, 1 usize, 2 bool, 3 string (into buf), 4 func(arg)(x), 5 func3(x,urand,nrand)(x,y,z), 6 func2(t,x)(x,y), 7 contains(section)
int opal_drv_config_read(const char* yaml, const char* section, const char* field, int kind, const char* arg, double x, double y, double z,
                         double* out, char* buf, int cap) {
    try {
        Config c = Config::from_string(yaml);
        c.with_context("constants");
        switch (kind) {
            case 0: *out = c.read_f64(section, field); break;
            case 1: *out = (double)c.read_usize(section, field); break;
            case 2: *out = c.read_bool(section, field) ? 1.0 : 0.0; break;
            case 3: {
                std::string joined;
                for (const std::string& s : c.read_strings(section, field)) joined += (joined.empty() ? "" : "\n") + s;
                copy_out(joined, buf, cap);
                break;
            }
            case 4: *out = c.func(section, field, arg)(x); break;
            case 5: *out = c.func3(section, field, "x", "urand", "nrand")(x, y, z); break;
            case 6: *out = c.func2(section, field, "t", "x")(x, y); break;
            case 7: *out = c.contains(section) ? 1.0 : 0.0; break;
            default: return -1;
        }
        return 0;
    } catch (const ConfigError& e) {
        copy_out(e.message(), buf, cap);
        return code(e);
    }
}

// kind: 0 {:e}, 1 {:.6e}, 2 {}, 3 SimulationTime, 4 PrettyDuration
int opal_drv_format(int kind, double v, char* buf, int cap) {
    if (kind == 3) return copy_out(fmt_simulation_time(v), buf, cap);
    if (kind == 4) return copy_out(fmt_pretty_duration(v), buf, cap);
    return copy_out(kind == 0 ? fmt_lower_exp(v) : (kind == 1 ? fmt_lower_exp6(v) : fmt_display(v)), buf, cap);
}

// parses one `output:` entry; returns validity; info = "ndim q0 q1 binspec_kind n dx weight_kind", file = its file name
int opal_drv_output_spec(const char* entry, const char* dir, long index, const char* species, char* info, int info_cap, char* file, int file_cap) {
    const OutputSpec sp = OutputSpec::parse(entry);
    if (!sp.valid) return 0;
    char tmp[256];
    std::snprintf(tmp, sizeof tmp, "%d %d %d %d %zu %.17g %d", (int)sp.axes.size(), sp.quantity[0], sp.axes.size() > 1 ? sp.quantity[1] : -1,
                  (int)sp.bspec.kind, sp.bspec.n, sp.bspec.dx, sp.weight_kind);
    copy_out(tmp, info, info_cap);
    copy_out(sp.filename(dir, (size_t)index, species), file, file_cap);
    return 1;
}

// writes a histogram with the given content as FITS (for the reader test)
int opal_drv_write_fits(const char* path, int dim, const long* bins, const double* min, const double* bin_sz, const double* cts, double total) {
    try {
        Histogram h;
        h.dim = (size_t)dim;
        size_t len = 1;
        for (int d = 0; d < dim; d++) {
            h.bins.push_back((size_t)bins[d]); h.min.push_back(min[d]); h.max.push_back(min[d] + bin_sz[d] * (double)bins[d]); h.bin_sz.push_back(bin_sz[d]);
            h.axis.push_back(d == 0 ? "x" : "px"); h.unit.push_back(d == 0 ? "m" : "MeV/c");
            len *= (size_t)bins[d];
        }
        h.cts.assign(cts, cts + len);
        h.total = total;
        h.name = dim == 1 ? "hgram/density/x" : "hgram/density/x_px";
        h.bunit = dim == 1 ? "1/m" : "1/(m.MeV/c)";
        h.write_fits(path);
        return 0;
    } catch (...) {
        return 1;
    }
}

// the first n uniforms of Xoshiro256StarStar::seed_from_u64(seed)
void opal_drv_rng_stream(unsigned long long seed, long n, double* out) {
    Xoshiro256StarStar r(seed);
    for (long i = 0; i < n; i++) out[i] = r.gen_f64();
}

// GridDesign::{unbalanced, balanced} with ne(x) given as an expression of x; writes numtasks cell counts and offsets
int opal_drv_grid_design(int balanced, int numtasks, long size, double xmin, double dx, long min_subsize, const char* ne_expr, long* nx_out,
                         double* offset_out) {
    try {
        GridDesign d;
        if (balanced) {
            Context ctx;
            const Expr e = Expr::parse(ne_expr);
            const char* names[1] = {"x"};
            d = GridDesign::balanced(numtasks, (size_t)size, xmin, dx, (size_t)min_subsize, [&](double x) { return e.eval(ctx, names, &x, 1); });
        } else {
            d = GridDesign::unbalanced(numtasks, (size_t)size, xmin, dx, (size_t)min_subsize);
        }
        for (int i = 0; i < numtasks; i++) { nx_out[i] = (long)d.nx[(size_t)i]; offset_out[i] = d.offset[(size_t)i]; }
        return 0;
    } catch (...) {
        return 1;
    }
}

// Population::new on `ncells` cells from the deck's section (ne/ni field name given); returns the particle count, fills up to cap entries
long opal_drv_load(const char* yaml, const char* section, const char* density_field, long npc, long ncells, double slab_xmin, double dx,
                   unsigned long long seed, int with_tau, long cap, long long* cell, double* x, double* ux, double* uy, double* uz, double* w, double* tau) {
    try {
        Config c = Config::from_string(yaml);
        c.with_context("constants");
        Xoshiro256StarStar rng(seed);
        const ParticleArrays p = load_population((size_t)npc, c.func(section, density_field, "x"), c.func3(section, "ux", "x", "urand", "nrand"),
                                                 c.func3(section, "uy", "x", "urand", "nrand"), c.func3(section, "uz", "x", "urand", "nrand"),
                                                 ncells, slab_xmin, dx, rng, with_tau != 0);
        const long n = (long)p.size();
        for (long i = 0; i < n && i < cap; i++) {
            cell[i] = p.cell[(size_t)i]; x[i] = p.x[(size_t)i]; ux[i] = p.ux[(size_t)i]; uy[i] = p.uy[(size_t)i]; uz[i] = p.uz[(size_t)i];
            w[i] = p.weight[(size_t)i];
            if (with_tau) tau[i] = p.tau[(size_t)i];
        }
        return n;
    } catch (...) {
        return -1;
    }
}

}  // extern "C"
