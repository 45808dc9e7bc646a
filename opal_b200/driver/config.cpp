// config.cpp -- see config.h
#include "config.h"

#include <cmath>
#include <cstdlib>
#include <fstream>
#include <sstream>

#include "constants.h"

namespace opal_driver {

std::string ConfigError::message() const {   // setup.rs:51-66
    const std::string help = "Usage: mpirun -n np ./opal input-file";
    switch (kind) {
        case ConfigErrorKind::MissingFile: return "Unable to open configuration file.\n" + help;
        case ConfigErrorKind::MissingSection: return "Could not find section \"" + section + "\".\n" + help;
        case ConfigErrorKind::MissingField: return "Could not find field \"" + field + "\" in section \"" + section + "\".\n" + help;
        default: return "Could not convert field \"" + field + "\" in section \"" + section + "\" to target type.\n" + help;
    }
}

Config Config::from_file(const std::string& path) {
    std::ifstream in(path);
    if (!in) throw ConfigError{ConfigErrorKind::MissingFile, "", ""};
    std::stringstream ss;
    ss << in.rdbuf();
    return from_string(ss.str());
}

Config Config::from_string(const std::string& text) {
    Config c;
    try {
        c.input_ = Yaml::load(text);
    } catch (const YamlError&) {
        throw ConfigError{ConfigErrorKind::MissingFile, "", ""};      // setup.rs:96 maps a scan error to MissingFile
    }
    if (c.input_.map.empty()) throw ConfigError{ConfigErrorKind::MissingFile, "", ""};
    return c;
}

static double rust_parse_f64(const std::string& s, bool* ok) {
    // text of a Yaml::Real; ".inf"/".nan" spellings fail Rust's f64::from_str and so does the reference (setup.rs:166, 288)
    if (!s.empty() && (s[0] == '.' || s.compare(0, 2, "-.") == 0 || s.compare(0, 2, "+.") == 0) && s.find_first_of("0123456789") == std::string::npos) {
        *ok = false;
        return 0.0;
    }
    char* end = nullptr;
    const double v = std::strtod(s.c_str(), &end);
    *ok = end && *end == 0;
    return v;
}

Config& Config::with_context(const std::string& section) {
    using namespace constants;
    ctx_.var("m", ELECTRON_MASS).var("me", ELECTRON_MASS).var("mp", PROTON_MASS).var("c", SPEED_OF_LIGHT)
        .var("e", ELEMENTARY_CHARGE).var("eV", ELEMENTARY_CHARGE).var("keV", 1.0e3 * ELEMENTARY_CHARGE).var("MeV", 1.0e6 * ELEMENTARY_CHARGE)
        .var("femto", 1.0e-15).var("pico", 1.0e-12).var("nano", 1.0e-9).var("micro", 1.0e-6).var("milli", 1.0e-3);
    ctx_.func("step", 3, [](const double* a, int) { return (a[0] >= a[1] && a[0] < a[2]) ? 1.0 : 0.0; });
    ctx_.func("gauss", 3, [](const double* a, int) { const double d = a[0] - a[1]; return std::exp(-(d * d) / (2.0 * (a[2] * a[2]))); });
    ctx_.func("critical", 1, [](const double* a, int) {
        return VACUUM_PERMITTIVITY * ELECTRON_MASS * (a[0] * a[0]) / (ELEMENTARY_CHARGE * ELEMENTARY_CHARGE);
    });
    ctx_.func("gauss_pulse_re", 4, [](const double* a, int) {          // setup.rs:113-122
        const double t = a[0], x = a[1], omega = a[2], sigma = a[3];
        const double phi = omega * (t - x / SPEED_OF_LIGHT);
        const double carrier = std::sin(phi) + phi * std::cos(phi) / (sigma * sigma);
        const double envelope = std::exp(-(phi * phi) / (2.0 * (sigma * sigma)));
        return carrier * envelope;
    });
    ctx_.func("gauss_pulse_im", 4, [](const double* a, int) {          // setup.rs:124-133
        const double t = a[0], x = a[1], omega = a[2], sigma = a[3];
        const double phi = omega * (t - x / SPEED_OF_LIGHT);
        const double carrier = std::cos(phi) - phi * std::sin(phi) / (sigma * sigma);
        const double envelope = std::exp(-(phi * phi) / (2.0 * (sigma * sigma)));
        return carrier * envelope;
    });

    const Yaml& sec = input_[section];
    if (sec.is_bad() || sec.kind != Yaml::Hash) return *this;
    const Context tmp = ctx_;      // "a constant cannot depend on other constants yet" (setup.rs:160)
    for (const auto& kv : sec.map) {
        const Yaml& v = kv.second;
        if (v.kind == Yaml::Real) {                    // setup.rs:166-168
            bool ok = false;
            const double num = rust_parse_f64(v.s, &ok);
            if (ok) ctx_.var(kv.first, num);
        } else if (v.kind == Yaml::String) {           // setup.rs:169-173
            try {
                ctx_.var(kv.first, Expr::parse(v.s).eval(tmp));
            } catch (const ExprError&) {
            }
        }                                              // integers, booleans, ... are ignored (setup.rs:174)
    }
    return *this;
}

bool Config::contains(const std::string& section) const { return !input_[section].is_bad(); }

const Yaml& Config::locate(const std::string& section, const std::string& field) const {
    if (input_[section].is_bad()) throw ConfigError{ConfigErrorKind::MissingSection, section, field};
    const Yaml& v = input_[section][field];
    if (v.is_bad()) throw ConfigError{ConfigErrorKind::MissingField, section, field};
    return v;
}

double Config::read_f64(const std::string& section, const std::string& field) const {
    const Yaml& v = locate(section, field);
    const ConfigError fail{ConfigErrorKind::ConversionFailure, section, field};
    if (v.kind == Yaml::Real) {
        bool ok = false;
        const double d = rust_parse_f64(v.s, &ok);
        if (!ok) throw fail;
        return d;
    }
    if (v.kind == Yaml::Integer) return (double)v.i;
    if (v.kind == Yaml::String) {
        try {
            return Expr::parse(v.s).eval(ctx_);
        } catch (const ExprError&) {
            throw fail;
        }
    }
    throw fail;
}

long long Config::read_i64(const std::string& section, const std::string& field) const {
    const Yaml& v = locate(section, field);
    if (v.kind != Yaml::Integer) throw ConfigError{ConfigErrorKind::ConversionFailure, section, field};
    return v.i;
}

size_t Config::read_usize(const std::string& section, const std::string& field) const {
    const long long i = read_i64(section, field);
    if (i < 0) throw ConfigError{ConfigErrorKind::ConversionFailure, section, field};
    return (size_t)i;
}

bool Config::read_bool(const std::string& section, const std::string& field) const {
    const Yaml& v = locate(section, field);
    if (v.kind != Yaml::Boolean) throw ConfigError{ConfigErrorKind::ConversionFailure, section, field};
    return v.b;
}

std::string Config::read_string(const std::string& section, const std::string& field) const {
    const Yaml& v = locate(section, field);
    if (v.kind != Yaml::String) throw ConfigError{ConfigErrorKind::ConversionFailure, section, field};
    return v.s;
}

std::vector<std::string> Config::read_strings(const std::string& section, const std::string& field) const {
    const Yaml& v = locate(section, field);
    if (v.kind == Yaml::String) return {v.s};
    if (v.kind == Yaml::Array) {
        std::vector<std::string> got;
        for (const Yaml& y : v.items) if (y.kind == Yaml::String) got.push_back(y.s);    // non-strings are dropped (setup.rs:346)
        if (!got.empty()) return got;
    }
    throw ConfigError{ConfigErrorKind::ConversionFailure, section, field};
}

Expr Config::bound_expr(const std::string& section, const std::string& field, const char* const* names, int n) const {
    const Yaml& v = locate(section, field);
    const ConfigError fail{ConfigErrorKind::ConversionFailure, section, field};
    if (v.kind != Yaml::String && v.kind != Yaml::Real) throw fail;
    try {
        Expr e = Expr::parse(v.s);
        // meval's bind_with_context fails at bind time when a free variable or function is unknown: probe once
        const double zeros[3] = {0.0, 0.0, 0.0};
        (void)e.eval(ctx_, names, zeros, n);
        return e;
    } catch (const ExprError&) {
        throw fail;
    }
}

std::function<double(double)> Config::func(const std::string& section, const std::string& field, const char* arg) const {
    const char* names[1] = {arg};
    Expr e = bound_expr(section, field, names, 1);
    const Context* ctx = &ctx_;
    const std::string a(arg);
    return [e, ctx, a](double x) { const char* nm[1] = {a.c_str()}; return e.eval(*ctx, nm, &x, 1); };
}

std::function<double(double, double)> Config::func2(const std::string& section, const std::string& field, const char* a0, const char* a1) const {
    const char* names[2] = {a0, a1};
    Expr e = bound_expr(section, field, names, 2);
    const Context* ctx = &ctx_;
    const std::string s0(a0), s1(a1);
    return [e, ctx, s0, s1](double x, double y) {
        const char* nm[2] = {s0.c_str(), s1.c_str()};
        const double v[2] = {x, y};
        return e.eval(*ctx, nm, v, 2);
    };
}

std::function<double(double, double, double)> Config::func3(const std::string& section, const std::string& field, const char* a0, const char* a1,
                                                            const char* a2) const {
    const char* names[3] = {a0, a1, a2};
    Expr e = bound_expr(section, field, names, 3);
    const Context* ctx = &ctx_;
    const std::string s0(a0), s1(a1), s2(a2);
    return [e, ctx, s0, s1, s2](double x, double y, double z) {
        const char* nm[3] = {s0.c_str(), s1.c_str(), s2.c_str()};
        const double v[3] = {x, y, z};
        return e.eval(*ctx, nm, v, 3);
    };
}

}  // namespace opal_driver
