// expr.cpp -- tokenizer, shunting yard and RPN evaluator (see expr.h for the grammar followed).
#include "expr.h"

#include <cctype>
#include <cmath>
#include <cstdlib>

namespace opal_driver {

Context::Context() {
    var("pi", 3.14159265358979323846);
    var("e", 2.71828182845904523536);
    auto f1 = [this](const char* n, double (*fn)(double)) { func(n, 1, [fn](const double* a, int) { return fn(a[0]); }); };
    f1("sqrt", std::sqrt); f1("abs", std::fabs); f1("exp", std::exp); f1("ln", std::log);
    f1("sin", std::sin); f1("cos", std::cos); f1("tan", std::tan);
    f1("asin", std::asin); f1("acos", std::acos); f1("atan", std::atan);
    f1("sinh", std::sinh); f1("cosh", std::cosh); f1("tanh", std::tanh);
    f1("asinh", std::asinh); f1("acosh", std::acosh); f1("atanh", std::atanh);
    f1("floor", std::floor); f1("ceil", std::ceil); f1("round", std::round);
    func("signum", 1, [](const double* a, int) { return std::isnan(a[0]) ? a[0] : (std::signbit(a[0]) ? -1.0 : 1.0); });   // f64::signum
    func("atan2", 2, [](const double* a, int) { return std::atan2(a[0], a[1]); });
    func("max", -1, [](const double* a, int n) { double m = a[0]; for (int i = 1; i < n; i++) m = std::fmax(m, a[i]); return m; });
    func("min", -1, [](const double* a, int n) { double m = a[0]; for (int i = 1; i < n; i++) m = std::fmin(m, a[i]); return m; });
}

bool Context::get_var(const std::string& name, double* v) const {
    auto it = vars_.find(name);
    if (it == vars_.end()) return false;
    *v = it->second;
    return true;
}

bool Context::call(const std::string& name, const double* args, int n, double* out) const {
    auto it = funcs_.find(name);
    if (it == funcs_.end()) return false;
    if ((it->second.nargs >= 0 && it->second.nargs != n) || n < 1) throw ExprError{"wrong number of arguments for " + name};
    *out = it->second.f(args, n);
    return true;
}

namespace {

using Tok = Expr::Tok;

int prec(const Tok& t) {
    if (t.kind == Tok::Bin) return (t.op == '+' || t.op == '-') ? 1 : (t.op == '^' ? 4 : 2);
    if (t.kind == Tok::Un) return 3;
    if (t.kind == Tok::Fact) return 5;
    return 0;
}
bool right_assoc(const Tok& t) { return t.kind == Tok::Bin && t.op == '^'; }

// infix token stream; `RParen` is represented as Bin with op ')' only inside this function
std::vector<Tok> tokenize(const std::string& s) {
    std::vector<Tok> out;
    size_t i = 0;
    bool expect_operand = true;      // decides unary vs binary +/-
    while (i < s.size()) {
        const char ch = s[i];
        if (std::isspace((unsigned char)ch)) { i++; continue; }
        if (std::isdigit((unsigned char)ch) || (ch == '.' && i + 1 < s.size() && std::isdigit((unsigned char)s[i + 1]))) {
            // digits [. digits] [e|E [+-] digits]
            size_t j = i;
            while (j < s.size() && std::isdigit((unsigned char)s[j])) j++;
            if (j < s.size() && s[j] == '.') { j++; while (j < s.size() && std::isdigit((unsigned char)s[j])) j++; }
            if (j < s.size() && (s[j] == 'e' || s[j] == 'E')) {
                size_t k = j + 1;
                if (k < s.size() && (s[k] == '+' || s[k] == '-')) k++;
                if (k < s.size() && std::isdigit((unsigned char)s[k])) { while (k < s.size() && std::isdigit((unsigned char)s[k])) k++; j = k; }
            }
            Tok t; t.kind = Tok::Num; t.num = std::strtod(s.substr(i, j - i).c_str(), nullptr);
            out.push_back(t); i = j; expect_operand = false; continue;
        }
        if (std::isalpha((unsigned char)ch) || ch == '_') {
            size_t j = i;
            while (j < s.size() && (std::isalnum((unsigned char)s[j]) || s[j] == '_')) j++;
            Tok t; t.name = s.substr(i, j - i);
            size_t k = j;
            while (k < s.size() && std::isspace((unsigned char)s[k])) k++;
            if (k < s.size() && s[k] == '(') { t.kind = Tok::Func; i = k + 1; expect_operand = true; }   // consumes the '('
            else { t.kind = Tok::Var; i = j; expect_operand = false; }
            out.push_back(t); continue;
        }
        Tok t;
        switch (ch) {
            case '+': case '-':
                t.kind = expect_operand ? Tok::Un : Tok::Bin; t.op = ch; expect_operand = true; break;
            case '*': case '/': case '%': case '^':
                if (expect_operand) throw ExprError{std::string("unexpected operator ") + ch};
                t.kind = Tok::Bin; t.op = ch; expect_operand = true; break;
            case '!':
                if (expect_operand) throw ExprError{"unexpected !"};
                t.kind = Tok::Fact; t.op = ch; break;
            case '(':
                if (!expect_operand) throw ExprError{"unexpected ("};
                t.kind = Tok::LParen; break;
            case ')':
                if (expect_operand) throw ExprError{"unexpected )"};
                t.kind = Tok::Bin; t.op = ')'; expect_operand = false; break;
            case ',':
                if (expect_operand) throw ExprError{"unexpected ,"};
                t.kind = Tok::Comma; expect_operand = true; break;
            default:
                throw ExprError{std::string("unexpected character ") + ch};
        }
        out.push_back(t); i++;
    }
    if (out.empty() || expect_operand) throw ExprError{"incomplete expression"};
    return out;
}

double factorial(double v) {
    if (v < 0.0 || std::floor(v) != v) return std::nan("");
    return std::tgamma(v + 1.0);
}

}  // namespace

Expr Expr::parse(const std::string& text) {
    const std::vector<Tok> in = tokenize(text);
    Expr e;
    std::vector<Tok> stack;             // operators, LParen, Func (a Func doubles as its own left parenthesis)
    std::vector<int> argc;              // argument counters of the open Func tokens
    for (const Tok& t : in) {
        switch (t.kind) {
            case Tok::Num: case Tok::Var: e.rpn_.push_back(t); break;
            case Tok::Func: stack.push_back(t); argc.push_back(1); break;
            case Tok::LParen: stack.push_back(t); break;
            case Tok::Comma:
                while (!stack.empty() && stack.back().kind != Tok::Func && stack.back().kind != Tok::LParen) { e.rpn_.push_back(stack.back()); stack.pop_back(); }
                if (stack.empty() || stack.back().kind != Tok::Func) throw ExprError{"comma outside a function call"};
                argc.back()++;
                break;
            case Tok::Un: stack.push_back(t); break;
            case Tok::Fact: e.rpn_.push_back(t); break;    // highest precedence, postfix: applies to what was just emitted
            case Tok::Bin:
                if (t.op == ')') {
                    while (!stack.empty() && stack.back().kind != Tok::Func && stack.back().kind != Tok::LParen) { e.rpn_.push_back(stack.back()); stack.pop_back(); }
                    if (stack.empty()) throw ExprError{"mismatched )"};
                    if (stack.back().kind == Tok::Func) { Tok f = stack.back(); f.nargs = argc.back(); argc.pop_back(); e.rpn_.push_back(f); }
                    stack.pop_back();
                    break;
                }
                while (!stack.empty() && (stack.back().kind == Tok::Bin || stack.back().kind == Tok::Un)) {
                    const int ps = prec(stack.back()), pt = prec(t);
                    if (ps > pt || (ps == pt && !right_assoc(t))) { e.rpn_.push_back(stack.back()); stack.pop_back(); }
                    else break;
                }
                stack.push_back(t);
                break;
        }
    }
    while (!stack.empty()) {
        if (stack.back().kind == Tok::LParen || stack.back().kind == Tok::Func) throw ExprError{"mismatched ("};
        e.rpn_.push_back(stack.back());
        stack.pop_back();
    }
    return e;
}

double Expr::eval(const Context& ctx, const char* const* names, const double* values, int nbound) const {
    double st[64];
    int sp = 0;
    for (const Tok& t : rpn_) {
        switch (t.kind) {
            case Tok::Num:
                if (sp >= 64) throw ExprError{"expression too deep"};
                st[sp++] = t.num; break;
            case Tok::Var: {
                double v = 0.0; bool ok = false;
                for (int k = 0; k < nbound && !ok; k++) if (t.name == names[k]) { v = values[k]; ok = true; }
                if (!ok && !ctx.get_var(t.name, &v)) throw ExprError{"unknown variable " + t.name};
                if (sp >= 64) throw ExprError{"expression too deep"};
                st[sp++] = v; break;
            }
            case Tok::Un:
                if (sp < 1) throw ExprError{"malformed expression"};
                if (t.op == '-') st[sp - 1] = -st[sp - 1];
                break;
            case Tok::Fact:
                if (sp < 1) throw ExprError{"malformed expression"};
                st[sp - 1] = factorial(st[sp - 1]); break;
            case Tok::Bin: {
                if (sp < 2) throw ExprError{"malformed expression"};
                const double b = st[--sp], a = st[sp - 1];
                double r = 0.0;
                switch (t.op) {
                    case '+': r = a + b; break;
                    case '-': r = a - b; break;
                    case '*': r = a * b; break;
                    case '/': r = a / b; break;
                    case '%': r = std::fmod(a, b); break;
                    case '^': r = std::pow(a, b); break;
                }
                st[sp - 1] = r; break;
            }
            case Tok::Func: {
                if (sp < t.nargs) throw ExprError{"malformed expression"};
                double r = 0.0;
                if (!ctx.call(t.name, st + sp - t.nargs, t.nargs, &r)) throw ExprError{"unknown function " + t.name};
                sp -= t.nargs;
                st[sp++] = r; break;
            }
            default: throw ExprError{"malformed expression"};
        }
    }
    if (sp != 1) throw ExprError{"malformed expression"};
    return st[0];
}

}  // namespace opal_driver
