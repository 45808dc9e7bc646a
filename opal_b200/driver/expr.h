// expr.h -- arithmetic expressions of opal's input files.
//
// The reference parses every numeric field with the `meval 0.2.0` crate (setup.rs:170,220-225,298);
// meval is not vendored under /root/reference, so this follows its published grammar:
//   tokens : numbers (digits, '.', exponent), identifiers, + - * / % ^ ( ) , and postfix '!'
//   precedence (shunting yard): binary + - (1, left) < * / % (2, left) < unary + - (3) < ^ (4, right)
//           < postfix ! (5);  so  -a^2 == -(a^2)  and  2^-1 == 0.5
//   built-ins: sqrt abs exp ln sin cos tan asin acos atan sinh cosh tanh asinh acosh atanh floor ceil
//           round signum (1 arg), atan2 (2 args), max min (n >= 1 args), constants pi and e
// plus opal's own context (setup.rs:135-153), which overrides `e` with the elementary charge.
// Pinned by the reference's own parser tests setup.rs:446-484 (tests/test_driver_config.py).
#pragma once
#include <functional>
#include <map>
#include <string>
#include <vector>

namespace opal_driver {

struct ExprError {
    std::string what;
};

class Context {
public:
    using Fn = std::function<double(const double*, int)>;
    Context();                                   // meval's built-in context
    Context& var(const std::string& name, double v) { vars_[name] = v; return *this; }
    Context& func(const std::string& name, int nargs /* -1: variadic, >= 1 */, Fn f) { funcs_[name] = {nargs, std::move(f)}; return *this; }
    bool get_var(const std::string& name, double* v) const;
    // returns false when unknown; throws ExprError on a wrong argument count
    bool call(const std::string& name, const double* args, int n, double* out) const;

private:
    struct F { int nargs; Fn f; };
    std::map<std::string, double> vars_;
    std::map<std::string, F> funcs_;
};

class Expr {
public:
    static Expr parse(const std::string& text);          // throws ExprError
    // free variables are looked up in `names`/`values` first (the bound arguments of Config::func*), then in ctx
    double eval(const Context& ctx, const char* const* names = nullptr, const double* values = nullptr, int nbound = 0) const;

    struct Tok {
        enum Kind { Num, Var, Func, Bin, Un, Fact, LParen, Comma } kind;
        double num = 0.0;
        std::string name;
        char op = 0;
        int nargs = 0;
    };

private:
    std::vector<Tok> rpn_;
};

}  // namespace opal_driver
