// yaml_subset.h -- the part of YAML that opal input files use (docs/input.md, examples/*.yaml):
// a block mapping of sections, each a block mapping of `key: value` pairs whose values are plain or
// quoted scalars, flow sequences `[a, b:c, d]` or folded/literal block scalars (`>` / `|`).
//
// Scalars are typed the way `yaml-rust 0.4.3` (the reference's loader, setup.rs:95) resolves plain scalars,
// because the reference dispatches on that type (setup.rs:164-177, 283-370):
//   true/false -> Boolean;  ~ / null -> Null;  parses as i64 (also 0x.., 0o.., +n) -> Integer;
//   parses as f64 (also .inf/.nan) -> Real (text kept);  otherwise, and every quoted scalar -> String.
#pragma once
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

namespace opal_driver {

struct YamlError {
    std::string what;
    int line;
};

struct Yaml {
    enum Kind { Bad, Null, Boolean, Integer, Real, String, Array, Hash } kind = Bad;
    bool b = false;
    int64_t i = 0;
    std::string s;                                   // Real keeps its text (yaml-rust does), String its value
    std::vector<Yaml> items;                         // Array
    std::vector<std::pair<std::string, Yaml>> map;   // Hash, in file order

    bool is_bad() const { return kind == Bad; }
    const Yaml& operator[](const std::string& key) const;   // Bad when absent or not a Hash (yaml-rust's BadValue)

    static Yaml load(const std::string& text);       // throws YamlError
    static Yaml scalar(const std::string& plain);    // type resolution of one plain scalar
};

}  // namespace opal_driver
