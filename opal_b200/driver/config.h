// config.h -- opal's input-file reader: the interface of `Config` in src/setup.rs:80-282
// (from_file, with_context, contains, read<T>, func, func2, func3) with the same error kinds
// (setup.rs:35-76) and the same type dispatch on the YAML node kind (setup.rs:283-370).
#pragma once
#include <functional>
#include <string>
#include <vector>

#include "expr.h"
#include "yaml_subset.h"

namespace opal_driver {

enum class ConfigErrorKind { MissingFile, MissingSection, MissingField, ConversionFailure };   // setup.rs:37-42

struct ConfigError {
    ConfigErrorKind kind;
    std::string section, field;
    std::string message() const;          // the text of setup.rs:54-66
};

class Config {
public:
    static Config from_file(const std::string& path);      // setup.rs:84
    static Config from_string(const std::string& text);    // setup.rs:93
    // setup.rs:109-181: default constants/functions, then the scalar entries of `section` ("constants")
    Config& with_context(const std::string& section);
    bool contains(const std::string& section) const;       // setup.rs:184

    // read<T> (setup.rs:191-207); each throws ConfigError
    double read_f64(const std::string& section, const std::string& field) const;             // setup.rs:283-300
    long long read_i64(const std::string& section, const std::string& field) const;          // setup.rs:302-310
    size_t read_usize(const std::string& section, const std::string& field) const;           // setup.rs:312-318
    bool read_bool(const std::string& section, const std::string& field) const;              // setup.rs:320-328
    std::string read_string(const std::string& section, const std::string& field) const;     // setup.rs:362-370
    std::vector<std::string> read_strings(const std::string& section, const std::string& field) const;   // setup.rs:330-360

    // func / func2 / func3 (setup.rs:211-282): the field's expression with 1-3 bound arguments
    std::function<double(double)> func(const std::string& section, const std::string& field, const char* arg) const;
    std::function<double(double, double)> func2(const std::string& section, const std::string& field, const char* a0, const char* a1) const;
    std::function<double(double, double, double)> func3(const std::string& section, const std::string& field, const char* a0, const char* a1,
                                                        const char* a2) const;

    const Context& context() const { return ctx_; }

private:
    const Yaml& locate(const std::string& section, const std::string& field) const;   // MissingSection / MissingField
    Expr bound_expr(const std::string& section, const std::string& field, const char* const* names, int n) const;
    Yaml input_;
    Context ctx_;
};

}  // namespace opal_driver
