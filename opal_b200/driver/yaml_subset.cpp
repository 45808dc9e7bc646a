// yaml_subset.cpp -- see yaml_subset.h
#include "yaml_subset.h"

#include <cctype>
#include <cerrno>
#include <cstdlib>
#include <cstring>

namespace opal_driver {

namespace {

const Yaml kBad;

std::string rstrip(std::string s) {
    while (!s.empty() && (s.back() == ' ' || s.back() == '\t' || s.back() == '\r')) s.pop_back();
    return s;
}
std::string strip(const std::string& s) {
    size_t a = 0;
    while (a < s.size() && (s[a] == ' ' || s[a] == '\t')) a++;
    return rstrip(s.substr(a));
}
int indent_of(const std::string& s) {
    int n = 0;
    while ((size_t)n < s.size() && s[n] == ' ') n++;
    return n;
}
// cut a trailing comment: '#' at the start or preceded by white space, outside quotes
std::string cut_comment(const std::string& s) {
    char q = 0;
    for (size_t i = 0; i < s.size(); i++) {
        const char ch = s[i];
        if (q) { if (ch == q) q = 0; continue; }
        if (ch == '"' || ch == '\'') { if (i == 0 || s[i - 1] == ' ' || s[i - 1] == '[' || s[i - 1] == ',') q = ch; continue; }
        if (ch == '#' && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) return s.substr(0, i);
    }
    return s;
}

bool parse_i64(const std::string& s, int64_t* out) {
    if (s.empty()) return false;
    const char* p = s.c_str();
    char* end = nullptr;
    errno = 0;
    long long v;
    if (s.size() > 2 && s[0] == '0' && s[1] == 'x') v = std::strtoll(p + 2, &end, 16);
    else if (s.size() > 2 && s[0] == '0' && s[1] == 'o') v = std::strtoll(p + 2, &end, 8);
    else {
        size_t k = (s[0] == '+' || s[0] == '-') ? 1 : 0;
        if (k >= s.size()) return false;
        for (size_t j = k; j < s.size(); j++) if (!std::isdigit((unsigned char)s[j])) return false;
        v = std::strtoll(p, &end, 10);
    }
    if (errno || !end || *end) return false;
    *out = v;
    return true;
}

// Rust's f64::from_str grammar: [+-] (digits [. digits*] | . digits) [e [+-] digits] | inf | infinity | nan
bool parse_f64(const std::string& s) {
    if (s == ".inf" || s == "-.inf" || s == "+.inf" || s == ".nan" || s == ".Inf" || s == ".INF" || s == ".NaN" || s == ".NAN") return true;
    size_t i = 0;
    if (i < s.size() && (s[i] == '+' || s[i] == '-')) i++;
    std::string rest = s.substr(i);
    for (auto& ch : rest) ch = (char)std::tolower((unsigned char)ch);
    if (rest == "inf" || rest == "infinity" || rest == "nan") return true;
    size_t nd = 0;
    while (i < s.size() && std::isdigit((unsigned char)s[i])) { i++; nd++; }
    if (i < s.size() && s[i] == '.') { i++; while (i < s.size() && std::isdigit((unsigned char)s[i])) { i++; nd++; } }
    if (nd == 0) return false;
    if (i < s.size() && (s[i] == 'e' || s[i] == 'E')) {
        i++;
        if (i < s.size() && (s[i] == '+' || s[i] == '-')) i++;
        size_t ne = 0;
        while (i < s.size() && std::isdigit((unsigned char)s[i])) { i++; ne++; }
        if (ne == 0) return false;
    }
    return i == s.size();
}

Yaml string_node(const std::string& v) {
    Yaml y; y.kind = Yaml::String; y.s = v; return y;
}

// one flow-sequence or scalar value (already comment-free and stripped)
Yaml parse_value(const std::string& v, int line) {
    if (v.empty()) { Yaml y; y.kind = Yaml::Null; return y; }
    if (v[0] == '[') {
        if (v.back() != ']') throw YamlError{"unterminated flow sequence", line};
        Yaml y; y.kind = Yaml::Array;
        std::string cur;
        char q = 0;
        int depth = 0;
        const std::string body = v.substr(1, v.size() - 2);
        for (size_t i = 0; i <= body.size(); i++) {
            const char ch = i < body.size() ? body[i] : ',';
            if (q) { cur += ch; if (ch == q) q = 0; continue; }
            if (ch == '"' || ch == '\'') { q = ch; cur += ch; continue; }
            if (ch == '(' ) depth++;
            if (ch == ')' ) depth--;
            if (ch == ',' && depth <= 0) {
                const std::string item = strip(cur);
                if (!item.empty()) y.items.push_back(parse_value(item, line));
                cur.clear();
                continue;
            }
            cur += ch;
        }
        return y;
    }
    if ((v[0] == '"' || v[0] == '\'') && v.size() >= 2 && v.back() == v[0]) return string_node(v.substr(1, v.size() - 2));
    return Yaml::scalar(v);
}

}  // namespace

const Yaml& Yaml::operator[](const std::string& key) const {
    if (kind != Hash) return kBad;
    for (const auto& kv : map) if (kv.first == key) return kv.second;
    return kBad;
}

Yaml Yaml::scalar(const std::string& v) {
    Yaml y;
    if (v == "~" || v == "null") { y.kind = Null; return y; }
    if (v == "true" || v == "false") { y.kind = Boolean; y.b = (v == "true"); return y; }
    if (parse_i64(v, &y.i)) { y.kind = Integer; return y; }
    if (parse_f64(v)) { y.kind = Real; y.s = v; return y; }
    return string_node(v);
}

Yaml Yaml::load(const std::string& text) {
    std::vector<std::string> lines;
    {
        size_t a = 0;
        while (a <= text.size()) {
            size_t b = text.find('\n', a);
            if (b == std::string::npos) b = text.size();
            lines.push_back(rstrip(text.substr(a, b - a)));
            a = b + 1;
        }
    }
    Yaml root; root.kind = Hash;
    Yaml* section = nullptr;
    int section_indent = -1;
    for (size_t n = 0; n < lines.size(); n++) {
        const int line_no = (int)n + 1;
        std::string body = rstrip(cut_comment(lines[n]));
        if (strip(body).empty()) continue;
        if (strip(body) == "---" || strip(body) == "...") continue;
        const int ind = indent_of(body);
        body = strip(body);
        // split `key: value` at the first ':' that is followed by a blank or ends the line
        size_t colon = std::string::npos;
        for (size_t i = 0; i < body.size(); i++)
            if (body[i] == ':' && (i + 1 == body.size() || body[i + 1] == ' ')) { colon = i; break; }
        if (colon == std::string::npos) throw YamlError{"expected `key: value`", line_no};
        std::string key = strip(body.substr(0, colon));
        if (key.size() >= 2 && (key[0] == '"' || key[0] == '\'') && key.back() == key[0]) key = key.substr(1, key.size() - 2);
        std::string val = strip(body.substr(colon + 1));

        Yaml node;
        if (!val.empty() && (val[0] == '>' || val[0] == '|')) {
            // block scalar: the following lines indented deeper than this key; `>` folds newlines to blanks,
            // default chomping keeps one trailing newline
            const bool fold = val[0] == '>';
            const bool strip_chomp = val.find('-') != std::string::npos;
            std::string acc;
            size_t m = n + 1;
            int block_indent = -1;
            for (; m < lines.size(); m++) {
                if (strip(lines[m]).empty()) { if (!acc.empty()) acc += "\n"; continue; }
                const int li = indent_of(lines[m]);
                if (li <= ind) break;
                if (block_indent < 0) block_indent = li;
                const std::string piece = lines[m].substr(std::min(li, block_indent));
                if (!acc.empty() && acc.back() != '\n') acc += fold ? " " : "\n";
                acc += piece;
            }
            while (!acc.empty() && acc.back() == '\n') acc.pop_back();
            if (!strip_chomp) acc += "\n";
            node = string_node(acc);
            n = m - 1;
        } else if (!val.empty() && val[0] == '[' && val.back() != ']') {
            // flow sequence continued over several lines
            size_t m = n + 1;
            for (; m < lines.size(); m++) {
                val += " " + strip(cut_comment(lines[m]));
                if (!val.empty() && rstrip(val).back() == ']') break;
            }
            node = parse_value(rstrip(val), line_no);
            n = m;
        } else if (val.empty()) {
            // a section header (or an empty value): decided by the next non-blank line's indentation
            size_t m = n + 1;
            while (m < lines.size() && strip(cut_comment(lines[m])).empty()) m++;
            if (m < lines.size() && indent_of(lines[m]) > ind && strip(lines[m]) != "---") { node.kind = Hash; }
            else node.kind = Null;
        } else {
            node = parse_value(val, line_no);
        }

        if (section && ind > section_indent) {
            section->map.emplace_back(key, node);
        } else {
            root.map.emplace_back(key, node);
            if (node.kind == Hash) { section = &root.map.back().second; section_indent = ind; }
            else section = nullptr;
        }
    }
    return root;
}

}  // namespace opal_driver
