// output.cpp -- see output.h
#include "output.h"

#include <algorithm>
#include <cfloat>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <stdexcept>

namespace opal_driver {

// ---- Rust float formatting -------------------------------------------------------------------
namespace {

// shortest decimal digits that round-trip (std::to_chars, the same Ryu-style shortest representation Rust prints):
// digits (no dot) and the decimal exponent of the first digit
void shortest_digits(double v, std::string* digits, int* exp10) {
    char buf[64];
    const auto r = std::to_chars(buf, buf + sizeof buf, v, std::chars_format::scientific);
    *r.ptr = 0;
    const char* e = std::strchr(buf, 'e');
    *exp10 = std::atoi(e + 1);
    digits->clear();
    for (const char* c = buf; c < e; c++) if (*c >= '0' && *c <= '9') digits->push_back(*c);
    while (digits->size() > 1 && digits->back() == '0') digits->pop_back();
}

std::string special(double v) {
    if (std::isnan(v)) return "NaN";
    return v < 0 ? "-inf" : "inf";
}

}  // namespace

// {:e} appended to `out` without a temporary per number (a grid file is 11 numbers on each of nx lines, every output):
// std::to_chars' shortest scientific form is Rust's but for the exponent, which Rust prints without '+' and without padding
void append_lower_exp(std::string& out, double v) {
    if (!std::isfinite(v)) { out += special(v); return; }
    if (v == 0.0) { out += std::signbit(v) ? "-0e0" : "0e0"; return; }
    char buf[48];
    const auto r = std::to_chars(buf, buf + sizeof buf, v, std::chars_format::scientific);
    const char* e = r.ptr - 1;
    while (e > buf && *e != 'e') e--;
    out.append(buf, (size_t)(e - buf));
    out += 'e';
    const char* p = e + 1;
    if (*p == '-') out += '-';
    if (*p == '-' || *p == '+') p++;
    while (*p == '0' && p + 1 < r.ptr) p++;
    out.append(p, (size_t)(r.ptr - p));
}

std::string fmt_lower_exp(double v) {
    std::string out;
    append_lower_exp(out, v);
    return out;
}

std::string fmt_lower_exp6(double v) {
    if (!std::isfinite(v)) return special(v);
    char buf[40];
    std::snprintf(buf, sizeof buf, "%.6e", v);           // d.dddddde[+-]XX  ->  Rust prints the exponent without sign padding
    char* e = std::strchr(buf, 'e');
    const int ex = std::atoi(e + 1);
    *e = 0;
    return std::string(buf) + "e" + std::to_string(ex);
}

std::string fmt_display(double v) {
    if (!std::isfinite(v)) return special(v);
    std::string sign = std::signbit(v) ? "-" : "";
    v = std::fabs(v);
    if (v == 0.0) return sign + "0";
    std::string d;
    int e;
    shortest_digits(v, &d, &e);
    std::string out;
    if (e >= 0) {
        if ((int)d.size() <= e + 1) out = d + std::string((size_t)(e + 1) - d.size(), '0');
        else out = d.substr(0, (size_t)e + 1) + "." + d.substr((size_t)e + 1);
    } else {
        out = "0." + std::string((size_t)(-e - 1), '0') + d;
    }
    return sign + out;
}

// ---- BinSpec / OutputSpec ----------------------------------------------------------------------
BinSpec BinSpec::from(const std::string& s) {            // hgram.rs:27-41
    BinSpec b;
    if (!s.empty() && s.find_first_not_of("0123456789") == std::string::npos) {          // parse::<usize>
        b.kind = BinSpecKind::FixedNumber;
        b.n = (size_t)std::strtoull(s.c_str(), nullptr, 10);
        return b;
    }
    if (!s.empty() && s[0] == '+' && s.size() > 1 && s.find_first_not_of("0123456789", 1) == std::string::npos) {
        b.kind = BinSpecKind::FixedNumber;
        b.n = (size_t)std::strtoull(s.c_str() + 1, nullptr, 10);
        return b;
    }
    char* end = nullptr;
    const double dx = std::strtod(s.c_str(), &end);
    if (!s.empty() && end != s.c_str() && *end == 0) {   // parse::<f64>
        b.kind = BinSpecKind::FixedSize;
        b.dx = dx;
        return b;
    }
    if (s == "log") b.kind = BinSpecKind::LogScaled;
    return b;                                             // "auto" and anything else: Automatic
}

static int quantity_of(const std::string& s, std::string* unit) {    // mod.rs:455-486
    static const char* names[] = {"x", "r", "energy", "px", "py", "pz", "p_perp", "theta", "phi", "longitude", "latitude", "work", "chi"};
    static const char* units[] = {"m", "m", "MeV", "MeV/c", "MeV/c", "MeV/c", "MeV/c", "rad", "rad", "rad", "rad", "J", "1"};
    for (int i = 0; i < OPAL_Q_COUNT_; i++)
        if (s == names[i]) { *unit = units[i]; return i; }
    return -1;
}

OutputSpec OutputSpec::parse(const std::string& o) {
    OutputSpec sp;
    std::vector<std::string> ss;
    {
        size_t a = 0;
        while (true) {
            const size_t b = o.find(':', a);
            ss.push_back(o.substr(a, b == std::string::npos ? std::string::npos : b - a));
            if (b == std::string::npos) break;
            a = b + 1;
        }
    }
    sp.weight = "weight";
    if (ss.size() >= 2 && !ss.back().empty() && ss.back().front() == '(' && ss.back().back() == ')') {    // mod.rs:455-466
        std::string last = ss.back();
        ss.pop_back();
        while (!last.empty() && last.front() == '(') last.erase(0, 1);       // trim_start_matches / trim_end_matches
        while (!last.empty() && last.back() == ')') last.pop_back();
        std::vector<std::string> parts;
        size_t a = 0;
        while (true) {
            const size_t b = last.find(';', a);
            parts.push_back(last.substr(a, b == std::string::npos ? std::string::npos : b - a));
            if (b == std::string::npos) break;
            a = b + 1;
        }
        if (parts.size() == 1) sp.weight = parts[0];
        else if (parts.size() == 2) { sp.bspec = BinSpec::from(parts[0]); sp.weight = parts[1]; }
    }
    bool ok = true;
    for (const std::string& s : ss) {
        std::string unit;
        const int q = quantity_of(s, &unit);
        if (q < 0) ok = false;
        sp.axes.push_back(s);
        sp.quantity.push_back(q);
        sp.units.push_back(unit);
    }
    if (sp.weight == "energy") sp.weight_kind = OPAL_W_ENERGY;
    else if (sp.weight == "weight" || sp.weight == "auto") sp.weight_kind = OPAL_W_WEIGHT;
    else ok = false;
    sp.valid = ok && (ss.size() == 1 || ss.size() == 2);     // mod.rs:503, 540: other dimensionalities produce nothing
    return sp;
}

std::string OutputSpec::filename(const std::string& dir, size_t index, const std::string& species_name) const {
    std::string f = "!" + dir + "/" + std::to_string(index) + "_" + species_name + "_" + axes[0];
    if (axes.size() == 2) f += "-" + axes[1];
    if (weight != "weight") f += "_" + weight;
    if (bspec.kind == BinSpecKind::LogScaled) f += "_log";
    return f + ".fits";
}

// ---- FITS primary HDU -----------------------------------------------------------------------------
namespace {

std::string card(const std::string& key, const std::string& value_field) {
    char buf[81];
    std::snprintf(buf, sizeof buf, "%-8.8s= %s", key.c_str(), value_field.c_str());
    std::string c(buf);
    c.resize(80, ' ');
    return c;
}
std::string card_int(const std::string& key, long long v) {
    char buf[32];
    std::snprintf(buf, sizeof buf, "%20lld", v);
    return card(key, buf);
}
std::string card_logical(const std::string& key, bool v) { return card(key, std::string(19, ' ') + (v ? "T" : "F")); }
std::string card_double(const std::string& key, double v) {
    char buf[40];
    std::snprintf(buf, sizeof buf, "%20.15G", v);          // CFITSIO's ffpkyd writes doubles with 15 decimals in fixed format
    std::string s(buf);
    if (s.find_first_of(".EN") == std::string::npos) {     // keep it a floating-point literal
        std::snprintf(buf, sizeof buf, "%20s", (s.substr(s.find_first_not_of(' ')) + ".").c_str());
        s = buf;
    }
    return card(key, s);
}
std::string card_string(const std::string& key, const std::string& v) {
    std::string q;
    for (char ch : v) { q.push_back(ch); if (ch == '\'') q.push_back('\''); }
    if (q.size() < 8) q.resize(8, ' ');
    return card(key, "'" + q + "'");
}

}  // namespace

void Histogram::write_fits(const std::string& filename) const {
    std::string path = filename;
    if (!path.empty() && path[0] == '!') path.erase(0, 1);          // CFITSIO: clobber an existing file
    std::string hdr;
    hdr += card_logical("SIMPLE", true);
    hdr += card_int("BITPIX", -64);                                 // ImageType::Double
    hdr += card_int("NAXIS", (long long)dim);
    for (size_t i = 0; i < dim; i++) hdr += card_int("NAXIS" + std::to_string(i + 1), (long long)bins[dim - 1 - i]);   // fitsio reverses row-major dimensions
    hdr += card_logical("EXTEND", true);
    for (size_t i = 0; i < dim; i++) {                              // hgram.rs:404-410
        const std::string n = std::to_string(i + 1);
        hdr += card_double("CRPIX" + n, 1.0);
        hdr += card_double("CRVAL" + n, min[i] + 0.5 * bin_sz[i]);
        hdr += card_double("CDELT" + n, bin_sz[i]);
        hdr += card_string("CNAME" + n, axis[i]);
        hdr += card_string("CUNIT" + n, unit[i]);
    }
    hdr += card_string("BUNIT", bunit);                             // hgram.rs:412-414
    hdr += card_double("TOTAL", total);
    hdr += card_string("OBJECT", name);
    double mn = 0.0, mx = 0.0;                                      // hgram.rs:416-419
    if (!cts.empty()) { mn = *std::min_element(cts.begin(), cts.end()); mx = *std::max_element(cts.begin(), cts.end()); }
    hdr += card_double("DATAMIN", mn);
    hdr += card_double("DATAMAX", mx);
    std::string end = "END";
    end.resize(80, ' ');
    hdr += end;
    hdr.resize((hdr.size() + 2879) / 2880 * 2880, ' ');

    std::string data(cts.size() * 8, '\0');
    for (size_t i = 0; i < cts.size(); i++) {                       // IEEE-754 big endian
        uint64_t bits;
        std::memcpy(&bits, &cts[i], 8);
        for (int b = 0; b < 8; b++) data[i * 8 + b] = (char)((bits >> (56 - 8 * b)) & 0xff);
    }
    data.resize((data.size() + 2879) / 2880 * 2880, '\0');
    std::ofstream f(path, std::ios::binary | std::ios::trunc);
    if (!f) throw std::runtime_error("cannot create " + path);
    f.write(hdr.data(), (std::streamsize)hdr.size());
    f.write(data.data(), (std::streamsize)data.size());
}

// SimulationTime's Display (setup.rs:417-437; pinned by the reference's time_format test, setup.rs:487-492)
std::string fmt_simulation_time(double seconds) {
    double power = 3.0 * std::floor(std::log10(std::fabs(seconds)) / 3.0);
    power = std::max(std::min(power, 0.0), -18.0);          // NaN (t == 0: log10 -> -inf -> min/max keep -18) as in Rust's f64::min/max
    if (std::isnan(power)) power = 0.0;
    const char* unit = " s";
    double scale = 1.0;
    switch ((int)power) {
        case -18: unit = "as"; scale = 1.0e18; break;
        case -15: unit = "fs"; scale = 1.0e15; break;
        case -12: unit = "ps"; scale = 1.0e12; break;
        case -9: unit = "ns"; scale = 1.0e9; break;
        case -6: unit = "\u03bcs"; scale = 1.0e6; break;
        case -3: unit = "ms"; scale = 1.0e3; break;
        default: break;
    }
    char buf[64];
    std::snprintf(buf, sizeof buf, "%8.2f %s", scale * seconds, unit);
    return buf;
}

// PrettyDuration's Display (setup.rs:400-414)
std::string fmt_pretty_duration(double seconds) {
    unsigned long long t = seconds > 0.0 ? (unsigned long long)seconds : 0ull;
    const unsigned long long s = t % 60; t /= 60;
    const unsigned long long min = t % 60; t /= 60;
    const unsigned long long hr = t % 24, d = t / 24;
    char buf[64];
    if (d > 0) std::snprintf(buf, sizeof buf, "%llud %02llu:%02llu:%02llu", d, hr, min, s);
    else std::snprintf(buf, sizeof buf, "%02llu:%02llu:%02llu", hr, min, s);
    return buf;
}

}  // namespace opal_driver
