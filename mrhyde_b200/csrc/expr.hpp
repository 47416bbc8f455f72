// Host-side expression front end: string -> branch tree -> postfix program (bytecode.h).
// Keeps the FunctionManager surface of the reference (addFunction(name, expr, location), lookup by name and
// location; /root/reference/src/managers/functionManager.cpp:49-69, :544-582) and its parsing rules
// (/root/reference/src/tools/interpreter.cpp:18-58 isScalar, :65-433 split, :441-526 isOperator;
//  leaf resolution order functionManager.cpp:112-479): a branch is "dep0; op1 dep1; op2 dep2 ..." applied
// left to right; '+'/'-' split first, then '*' '/' '<' '>', then '^', then parentheses / name(...).
// Quirks kept: leading '-' becomes "0.0-..."; a sign inside a literal such as 1.0e-3*x splits the literal.
#pragma once
#include <cctype>
#include <cmath>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>
#include "bytecode.h"

namespace mrh {

struct FunctionTable {
  std::string location;                                   // "ip" | "side ip" | "point"
  std::vector<std::pair<std::string, std::string>> fns;   // name -> expression (spaces stripped)
  int find(const std::string& n) const { for (size_t i = 0; i < fns.size(); ++i) if (fns[i].first == n) return (int)i; return -1; }
};

class ExprCompiler {
 public:
  std::vector<std::string> varlist;
  std::vector<std::pair<std::string, std::vector<double>>> parameters;     // scalar / vector parameters: leaves `name`, `name(k)`
  int dim = 3;
  std::vector<FunctionTable> tables;       // one per location
  std::vector<Instr> code;                 // all compiled programs, concatenated
  std::vector<double> cpool;

  ExprCompiler() { tables.resize(3); tables[0].location = "ip"; tables[1].location = "side ip"; tables[2].location = "point"; }

  // first registration wins, like FunctionManager::addFunction
  int addFunction(const std::string& name, const std::string& expr, const std::string& location) {
    for (auto& T : tables) if (T.location == location) {
      int i = T.find(name);
      if (i >= 0) return i;
      std::string e; for (char c : expr) if (c != ' ') e += c;
      T.fns.push_back({name, e});
      return (int)T.fns.size() - 1;
    }
    throw std::runtime_error("unknown function location: " + location);
  }
  bool has(const std::string& name, const std::string& location) const {
    for (auto& T : tables) if (T.location == location) return T.find(name) >= 0;
    return false;
  }

  // compile function `name` at `location` into a program appended to `code`
  Prog compile(const std::string& name, const std::string& location) {
    const FunctionTable* T = nullptr;
    for (auto& t : tables) if (t.location == location) T = &t;
    if (!T) throw std::runtime_error("unknown function location: " + location);
    const int fi = T->find(name);
    if (fi < 0) throw std::runtime_error("Error: function manager could not evaluate: " + name + " at " + location);
    std::vector<Instr> out; bool state = false;
    emit(*T, T->fns[fi].second, out, state, 0);
    fold(out);
    fuse(out);
    int depth = 0, sp = 0;
    for (auto& in : out) { sp += (in.op >= OP_ADD && in.op <= OP_MEAN && in.arg != 0) ? 0 : delta(in.op); if (sp + 1 > depth) depth = sp + 1; }
    if (depth > MRH_MAXSTACK) throw std::runtime_error("expression too deep for the device evaluator: " + name);
    Prog p; p.pad_ = 0; p.cval = 0.0;
    if (out.size() == 1 && out[0].op == OP_CONST) { p.kind = 0; p.cval = cpool[out[0].cidx]; p.start = 0; p.len = 0; return p; }
    p.kind = state ? 2 : 1; p.start = (int)code.size(); p.len = (int)out.size();
    code.insert(code.end(), out.begin(), out.end());
    return p;
  }

  // Literal test of the reference (src/tools/interpreter.cpp:18-58), restated as a scan over character classes: only
  // digits, at most one '.', at most one exponent marker, and a sign only in front or right behind the exponent marker.
  // (Loose on purpose, like the reference: "", "." and "e" pass here and fail later in the conversion.)
  static bool isScalar(const std::string& s) {
    int dots = 0, exps = 0;
    char prev = 'e';                                  // makes a leading sign acceptable
    for (const char c : s) {
      const bool expo = (c == 'e' || c == 'E');
      const bool sign = (c == '+' || c == '-');
      if (c == '.') ++dots;
      else if (expo) ++exps;
      else if (sign) { if (!(prev == 'e' || prev == 'E')) return false; }
      else if (!isdigit((unsigned char)c)) return false;
      if (dots > 1 || exps > 1) return false;
      prev = c;
    }
    return true;
  }

 private:
  struct Dep { std::string expr, op; };

  static int delta(uint8_t op) {
    if (op <= OP_FIELD) return 1;
    if (op >= OP_ADD && op <= OP_MEAN) return -1;
    if (op == OP_NIP) return -1;
    return 0;
  }
  bool findParameter(const std::string& expr, double& val) const {
    for (auto& pr : parameters) {
      const std::string& nm = pr.first; const size_t L = nm.length(), n = expr.length();
      int idx = -1;
      if (expr == nm) idx = 0;
      else if (n == L + 3 && expr.compare(0, L, nm) == 0 && expr[n - 3] == '(' && expr[n - 1] == ')' && isdigit((unsigned char)expr[n - 2])) idx = expr[n - 2] - '0';
      else if (n == L + 4 && expr.compare(0, L, nm) == 0 && expr[n - 4] == '(' && expr[n - 1] == ')' && isdigit((unsigned char)expr[n - 3]) && isdigit((unsigned char)expr[n - 2]))
        idx = (expr[n - 3] - '0') * 10 + (expr[n - 2] - '0');
      if (idx >= 0) {
        if (idx >= (int)pr.second.size()) throw std::runtime_error("parameter component out of range: " + expr);
        val = pr.second[idx]; return true;
      }
    }
    return false;
  }
  int constIndex(double v) { cpool.push_back(v); return (int)cpool.size() - 1; }

  int fieldId(const std::string& e) const {
    static const char* xyz[3] = {"x", "y", "z"};
    for (size_t v = 0; v < varlist.size(); ++v) {
      if (e == varlist[v]) return (int)v * FIELD_KINDS + 0;
      for (int c = 0; c < 3; ++c) if (e == "grad(" + varlist[v] + ")[" + xyz[c] + "]") return (int)v * FIELD_KINDS + 1 + c;
      if (e == varlist[v] + "_t") return (int)v * FIELD_KINDS + 4;
    }
    return -1;
  }

  static int binaryOp(const std::string& op) {
    static const std::map<std::string, int> m = {{"plus", OP_ADD}, {"minus", OP_SUB}, {"times", OP_MUL}, {"divide", OP_DIV},
      {"power", OP_POW}, {"lt", OP_LT}, {"lte", OP_LTE}, {"gt", OP_GT}, {"gte", OP_GTE}, {"max", OP_MAX}, {"min", OP_MIN}, {"mean", OP_MEAN}};
    auto it = m.find(op); return it == m.end() ? -1 : it->second;
  }
  static int unaryOp(const std::string& op) {
    static const std::map<std::string, int> m = {{"sin", OP_SIN}, {"cos", OP_COS}, {"tan", OP_TAN}, {"exp", OP_EXP}, {"log", OP_LOG},
      {"sqrt", OP_SQRT}, {"sinh", OP_SINH}, {"cosh", OP_COSH}, {"abs", OP_ABS}};
    auto it = m.find(op); return it == m.end() ? -1 : it->second;
  }

  void emit(const FunctionTable& T, const std::string& expr, std::vector<Instr>& out, bool& state, int depth) {
    if (depth > 64) throw std::runtime_error("Error: MrHyDE was not able to decompose (recursion): " + expr);
    const bool side = (T.location == "side ip");
    // 1. solution fields of the workset
    { const int f = fieldId(expr); if (f >= 0) { out.push_back({(uint8_t)OP_FIELD, (uint8_t)f, 0}); state = true; return; } }
    // 2. scalar fields
    if (expr == "x") { out.push_back({(uint8_t)OP_X, 0, 0}); return; }
    if (expr == "y") { out.push_back({(uint8_t)OP_Y, 0, 0}); return; }
    if (expr == "z") { out.push_back({(uint8_t)OP_Z, 0, 0}); return; }
    if (side) {
      static const char* nn[6] = {"n[x]", "n[y]", "n[z]", "t[x]", "t[y]", "t[z]"};
      for (int c = 0; c < 6; ++c) if (expr == nn[c]) { out.push_back({(uint8_t)(OP_NX + c), 0, 0}); return; }
    }
    // 2b. parameters: `name` is component 0, `name(k)` / `name(kk)` component k (functionManager.cpp:270-355); constants of the pass
    { double pv = 0.0; if (findParameter(expr, pv)) { out.push_back({(uint8_t)OP_CONST, 0, (uint16_t)constIndex(pv)}); return; } }
    // 3. another function of the same location
    { const int j = T.find(expr); if (j >= 0) { emit(T, T.fns[j].second, out, state, depth + 1); return; } }
    // 4. literal, time, pi
    if (isScalar(expr)) {
      double v; try { v = std::stod(expr); } catch (...) { throw std::runtime_error("Error: MrHyDE was not able to decompose " + expr); }
      out.push_back({(uint8_t)OP_CONST, 0, (uint16_t)constIndex(v)}); return; }
    if (expr == "t") { out.push_back({(uint8_t)OP_TIME, 0, 0}); return; }
    if (expr == "pi") { out.push_back({(uint8_t)OP_CONST, 0, (uint16_t)constIndex(M_PI)}); return; }
    // 5. known operator / split
    std::vector<Dep> deps;
    if (!isOperator(expr, deps)) split(expr, deps);
    if (deps.empty()) throw std::runtime_error("Error: MrHyDE was not able to decompose " + expr);
    for (size_t k = 0; k < deps.size(); ++k) {
      if (deps[k].expr == expr && deps.size() == 1 && deps[k].op.empty()) throw std::runtime_error("Error: MrHyDE was not able to decompose " + expr);
      emit(T, deps[k].expr, out, state, depth + 1);
      const std::string& op = deps[k].op;
      const int b = binaryOp(op), u = unaryOp(op);
      if (k > 0 && b >= 0) { out.push_back({(uint8_t)b, 0, 0}); continue; }
      if (u >= 0) out.push_back({(uint8_t)u, 0, 0});
      else if (!op.empty() && !(k == 0 && b >= 0)) {
        if (op == "emax" || op == "emin" || op == "emean")
          throw std::runtime_error("element-wide operator '" + op + "' is not supported inside the fused kernel");
        throw std::runtime_error("Error: MrHyDE was not able to decompose (unknown operator '" + op + "') " + expr);
      }
      if (k > 0) out.push_back({(uint8_t)OP_NIP, 0, 0});
    }
  }

  static void split(const std::string& s0, std::vector<Dep>& deps) {
    std::string s = s0;
    if (!s.empty() && s[0] == '-') s = "0.0" + s;
    if (s.empty()) return;
    if (s.length() == 1) { deps.push_back({s, ""}); return; }
    size_t num_pm = 0, num_mdp = 0, num_pow = 0; int paren = 0;
    for (size_t i = 0; i < s.length(); i++) {
      if (s[i] == '(') paren += 1;
      else if (s[i] == ')') paren -= 1;
      else if (paren == 0) {
        if (s[i] == '+' || s[i] == '-') num_pm += 1;
        if (s[i] == '*' || s[i] == '/' || s[i] == '<' || s[i] == '>') num_mdp += 1;
        if (s[i] == '^') num_pow += 1;
      }
    }
    paren = 0;
    std::string cur, op;
    auto flush = [&](const std::string& next) { deps.push_back({cur, op}); cur.clear(); op = next; };
    if (num_pm > 0) {
      for (size_t i = 0; i < s.length(); i++) {
        const char c = s[i];
        if (c == ' ' || c == '=') {}
        else if (c == '(') { paren += 1; cur += c; }
        else if (c == ')') { paren -= 1; cur += c; }
        else if (paren == 0 && c == '+' && !cur.empty()) flush("plus");
        else if (paren == 0 && c == '-' && !cur.empty()) flush("minus");
        else cur += c;
        if (i == s.length() - 1 && !cur.empty() && paren == 0) deps.push_back({cur, op});
      }
    } else if (num_mdp > 0) {
      for (size_t i = 0; i < s.length(); i++) {
        const char c = s[i];
        if (c == ' ') {}
        else if (c == '(') { paren += 1; cur += c; }
        else if (c == ')') { paren -= 1; cur += c; }
        else if (paren == 0 && c == '*') flush("times");
        else if (paren == 0 && c == '/') flush("divide");
        else if (paren == 0 && c == '<') { if (i + 1 < s.length() && s[i + 1] == '=') { flush("lte"); ++i; } else flush("lt"); }
        else if (paren == 0 && c == '>') { if (i + 1 < s.length() && s[i + 1] == '=') { flush("gte"); ++i; } else flush("gt"); }
        else cur += c;
        if (i == s.length() - 1) deps.push_back({cur, op});
      }
    } else if (num_pow > 0) {
      for (size_t i = 0; i < s.length(); i++) {
        const char c = s[i];
        if (c == '(') { paren += 1; cur += c; }
        else if (c == ')') { paren -= 1; cur += c; }
        else if (paren == 0 && c == '^') flush("power");
        else cur += c;
        if (i == s.length() - 1) deps.push_back({cur, op});
      }
    } else {
      if (s[0] == '(' && s[s.length() - 1] == ')') deps.push_back({s.substr(1, s.length() - 2), ""});
      else {
        bool found = false; size_t pindex = 0;
        for (size_t k = 1; k + 1 < s.length(); k++) if (s[k] == '(' && !found) { found = true; pindex = k; }
        if (found && s[s.length() - 1] == ')') deps.push_back({s.substr(pindex + 1, s.length() - pindex - 2), s.substr(0, pindex)});
      }
    }
  }

  static bool isOperator(const std::string& s, std::vector<Dep>& deps) {
    static const char* known[] = {"sin", "cos", "exp", "log", "tan", "abs", "max", "min", "mean", "emax", "emin", "emean", "sqrt", "sinh", "cosh"};
    std::string oper, argument; bool found = false;
    for (const char* kop : known) {
      if (found) break;
      const std::string ko(kop); const size_t L = ko.length();
      if (s.length() > L + 1 && s.compare(0, L, ko) == 0) {
        bool iseq = (s[L] == '(' && s[s.length() - 1] == ')');
        if (iseq) for (size_t j = L + 1; j + 1 < s.length(); j++) if (s[j] == ')') iseq = false;
        if (iseq) { found = true; oper = ko; argument = s.substr(L + 1, s.length() - L - 2); }
      }
    }
    if (!found) return false;
    bool has_comma = false; size_t comma = 0;
    for (size_t i = 0; i + 1 < argument.length(); ++i) if (argument[i] == ',') { has_comma = true; comma = i; }
    if (has_comma) { deps.push_back({argument.substr(0, comma), ""}); deps.push_back({argument.substr(comma + 1), oper}); }
    else deps.push_back({argument, oper});
    return true;
  }

  // [leaf][binary] -> binary with fused right operand (bytecode.h); fields are not fused (arg holds the field id)
  static void fuse(std::vector<Instr>& p) {
    for (size_t i = 1; i < p.size(); ++i) {
      const uint8_t op = p[i].op, lop = p[i - 1].op;
      if (op >= OP_ADD && op <= OP_MEAN && p[i].arg == 0 && lop < OP_FIELD && i >= 2) {
        // the leaf must be the right operand: the instruction before it must leave the left operand on the stack
        p[i].arg = (uint8_t)(lop + 1); p[i].cidx = p[i - 1].cidx;
        p.erase(p.begin() + i - 1); --i;
      }
    }
  }
  // fold constant sub-programs exactly as the device would evaluate them (same operations, same order)
  void fold(std::vector<Instr>& p) {
    bool changed = true;
    while (changed) {
      changed = false;
      for (size_t i = 0; i < p.size() && !changed; ++i) {
        const uint8_t op = p[i].op;
        if (op >= OP_ADD && op <= OP_MEAN && i >= 2 && p[i - 1].op == OP_CONST && p[i - 2].op == OP_CONST) {
          Instr tmp[3] = {p[i - 2], p[i - 1], p[i]};
          PointState<double> ps{}; ps.fields = nullptr;
          const double v = eval_code<double>(tmp, cpool.data(), 0, 3, ps);
          p[i - 2] = {(uint8_t)OP_CONST, 0, (uint16_t)constIndex(v)};
          p.erase(p.begin() + i - 1, p.begin() + i + 1); changed = true;
        } else if (op >= OP_SIN && op <= OP_ABS && i >= 1 && p[i - 1].op == OP_CONST) {
          Instr tmp[2] = {p[i - 1], p[i]};
          PointState<double> ps{}; ps.fields = nullptr;
          const double v = eval_code<double>(tmp, cpool.data(), 0, 2, ps);
          p[i - 1] = {(uint8_t)OP_CONST, 0, (uint16_t)constIndex(v)};
          p.erase(p.begin() + i); changed = true;
        } else if (op == OP_NIP && i >= 2 && p[i - 1].op == OP_CONST && p[i - 2].op == OP_CONST) {
          p[i - 2] = p[i - 1]; p.erase(p.begin() + i - 1, p.begin() + i + 1); changed = true;
        }
      }
    }
    if (cpool.size() > 65000) throw std::runtime_error("constant pool overflow");
  }
};

}  // namespace mrh
