// ORACLE (test infrastructure, NOT product code).
// Run-time string-expression evaluation restating
//   /root/reference/src/managers/functionManager.cpp:49-69 (addFunction), :112-482 (decomposeFunctions),
//   :512-537 (checkDepDataType), :544-582 (evaluate by name), :588-760 (recursive evaluate),
//   :762-1900 (evaluateOp*: "", plus, minus, times, divide, power, sin.., lt/gt/lte/gte, max/min/mean, emax/emin/emean)
//   /root/reference/src/tools/interpreter.cpp:18-58 (isScalar), :65-433 (split), :441-526 (isOperator)
//   /root/reference/src/tools/dag.hpp (Branch / Tree / Forest).
// Branch data is "dep0; op1 dep1; op2 dep2; ..." applied left to right.
#pragma once
#include <cctype>
#include <string>
#include <vector>
#include "workset.hpp"

namespace orc {

inline bool interp_isScalar(const std::string& s) {
  bool isnum = true; int numdots = 0, nume = 0;
  for (size_t k = 0; k < s.length(); k++) {
    if (!isdigit((unsigned char)s[k])) {
      if (s[k] == '.') { if (numdots == 0) numdots += 1; else isnum = false; }
      else if (s[k] == 'e' || s[k] == 'E') { if (nume == 0) nume += 1; else isnum = false; }
      else if (s[k] == '+' || s[k] == '-') { if (k > 0 && !(s[k - 1] == 'e' || s[k - 1] == 'E')) isnum = false; }
      else isnum = false;
    }
  }
  return isnum;
}

template <class EvalT>
struct Branch {
  std::string expr;
  bool is_leaf = false, is_decomposed = false, is_func = false, is_view = false, is_AD = false;
  bool is_constant = false, is_workset_data = false, is_time = false;
  int workset_data_index = 0, func_index = 0;
  std::vector<int> dep_list; std::vector<std::string> dep_ops;
  double sSc = 0.0; EvalT sAD = EvalT(0.0);
  Arr2<double> vSc; Arr2<EvalT> vAD;
  Branch() {}
  explicit Branch(const std::string& e) : expr(e) {}
};

template <class EvalT>
struct Tree { std::string name; std::vector<Branch<EvalT>> branches; };

template <class EvalT>
struct Vista {
  bool is_view = false, is_AD = false;
  double sSc = 0.0; EvalT sAD = EvalT(0.0);
  const Arr2<double>* vSc = nullptr; const Arr2<EvalT>* vAD = nullptr;
  EvalT operator()(int e, int q) const {
    if (is_view) return is_AD ? (*vAD)(e, q) : EvalT((*vSc)(e, q));
    return is_AD ? sAD : EvalT(sSc);
  }
};

template <class EvalT>
struct Forest { std::string location; int dim0 = 0, dim1 = 0; std::vector<Tree<EvalT>> trees; };

template <class EvalT>
class FunctionManager {
 public:
  Workset<EvalT>* wkset = nullptr;
  std::vector<Forest<EvalT>> forests;
  std::vector<std::string> known_ops{"sin", "cos", "exp", "log", "tan", "abs", "max", "min", "mean",
                                     "emax", "emin", "emean", "sqrt", "sinh", "cosh"};

  std::vector<std::pair<std::string, std::vector<double>>> parameters;
  bool findParameter(const std::string& expr, double& val) const {
    for (auto& pr : parameters) {
      const std::string& nm = pr.first; const size_t L = nm.length(), n = expr.length();
      int idx = -1;
      if (expr == nm) idx = 0;
      else if (n == L + 3 && expr.compare(0, L, nm) == 0 && expr[n - 3] == '(' && expr[n - 1] == ')' && isdigit((unsigned char)expr[n - 2])) idx = expr[n - 2] - '0';
      else if (n == L + 4 && expr.compare(0, L, nm) == 0 && expr[n - 4] == '(' && expr[n - 1] == ')' && isdigit((unsigned char)expr[n - 3]) && isdigit((unsigned char)expr[n - 2]))
        idx = (expr[n - 3] - '0') * 10 + (expr[n - 2] - '0');
      if (idx >= 0) {
        if (idx >= (int)pr.second.size()) throw std::runtime_error("oracle funcman: parameter component out of range: " + expr);
        val = pr.second[idx]; return true;
      }
    }
    return false;
  }

  FunctionManager(int num_elem, int num_ip, int num_ip_side) {
    forests.resize(2);
    forests[0].location = "ip"; forests[0].dim0 = num_elem; forests[0].dim1 = num_ip;
    forests[1].location = "side ip"; forests[1].dim0 = num_elem; forests[1].dim1 = num_ip_side;
  }

  int addFunction(const std::string& fname, const std::string& expression, const std::string& location) {
    int findex = 0;
    for (auto& F : forests) if (F.location == location) {
      bool found = false;
      for (size_t j = 0; j < F.trees.size(); ++j) if (F.trees[j].name == fname) { found = true; findex = (int)j; }
      if (!found) {
        Tree<EvalT> t; t.name = fname;
        std::string e; for (char c : expression) if (c != ' ') e += c;   // Tree ctor strips spaces (dag.hpp)
        t.branches.push_back(Branch<EvalT>(e));
        F.trees.push_back(t); findex = (int)F.trees.size() - 1;
      }
    }
    return findex;
  }

  // ---- interpreter ------------------------------------------------------------------------------
  static void pushDep(std::vector<Branch<EvalT>>& br, size_t index, const std::string& e, const std::string& op) {
    br.push_back(Branch<EvalT>(e));
    br[index].dep_list.push_back((int)br.size() - 1);
    br[index].dep_ops.push_back(op);
  }
  static void split(std::vector<Branch<EvalT>>& br, size_t index) {
    std::string s = br[index].expr;
    if (!s.empty() && s[0] == '-') s = "0.0" + s;
    if (s.empty()) return;
    if (s.length() == 1) { pushDep(br, index, s, ""); return; }
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
    if (num_pm > 0) {
      for (size_t i = 0; i < s.length(); i++) {
        const char c = s[i];
        if (c == ' ' || c == '=') {}
        else if (c == '(') { paren += 1; cur += c; }
        else if (c == ')') { paren -= 1; cur += c; }
        else if (paren == 0 && c == '+' && cur.length() > 0) { pushDep(br, index, cur, op); cur = ""; op = "plus"; }
        else if (paren == 0 && c == '-' && cur.length() > 0) { pushDep(br, index, cur, op); cur = ""; op = "minus"; }
        else cur += c;
        if (i == s.length() - 1 && cur.length() > 0 && paren == 0) pushDep(br, index, cur, op);
      }
    } else if (num_mdp > 0) {
      for (size_t i = 0; i < s.length(); i++) {
        const char c = s[i];
        if (c == ' ') {}
        else if (c == '(') { paren += 1; cur += c; }
        else if (c == ')') { paren -= 1; cur += c; }
        else if (paren == 0 && c == '*') { pushDep(br, index, cur, op); cur = ""; op = "times"; }
        else if (paren == 0 && c == '/') { pushDep(br, index, cur, op); cur = ""; op = "divide"; }
        else if (paren == 0 && c == '<') { pushDep(br, index, cur, op); cur = ""; if (i + 1 < s.length() && s[i + 1] == '=') { op = "lte"; ++i; } else op = "lt"; }
        else if (paren == 0 && c == '>') { pushDep(br, index, cur, op); cur = ""; if (i + 1 < s.length() && s[i + 1] == '=') { op = "gte"; ++i; } else op = "gt"; }
        else cur += c;
        if (i == s.length() - 1) pushDep(br, index, cur, op);
      }
    } else if (num_pow > 0) {
      for (size_t i = 0; i < s.length(); i++) {
        const char c = s[i];
        if (c == '(') { paren += 1; cur += c; }
        else if (c == ')') { paren -= 1; cur += c; }
        else if (paren == 0 && c == '^') { pushDep(br, index, cur, op); cur = ""; op = "power"; }
        else cur += c;
        if (i == s.length() - 1) pushDep(br, index, cur, op);
      }
    } else {
      if (s[0] == '(' && s[s.length() - 1] == ')') pushDep(br, index, s.substr(1, s.length() - 2), "");
      else {
        bool foundparen = false; size_t pindex = 0;
        for (size_t k = 1; k + 1 < s.length(); k++) if (s[k] == '(' && !foundparen) { foundparen = true; pindex = k; }
        if (foundparen && s[s.length() - 1] == ')')
          pushDep(br, index, s.substr(pindex + 1, s.length() - pindex - 2), s.substr(0, pindex));
      }
    }
  }
  bool isOperator(std::vector<Branch<EvalT>>& br, size_t index) {
    const std::string s = br[index].expr;
    std::string oper, argument; bool found = false;
    for (size_t k = 0; !found && k < known_ops.size(); ++k) {
      const size_t L = known_ops[k].length();
      if (s.length() > L + 1 && s.compare(0, L, known_ops[k]) == 0) {
        bool iseq = (s[L] == '(' && s[s.length() - 1] == ')');
        if (iseq) for (size_t j = L + 1; j + 1 < s.length(); j++) if (s[j] == ')') iseq = false;
        if (iseq) { found = true; oper = known_ops[k]; argument = s.substr(L + 1, s.length() - L - 2); }
      }
    }
    if (!found) return false;
    bool has_comma = false; size_t comma = 0;
    for (size_t i = 0; i + 1 < argument.length(); ++i) if (argument[i] == ',') { has_comma = true; comma = i; }
    if (has_comma) {
      pushDep(br, index, argument.substr(0, comma), "");
      pushDep(br, index, argument.substr(comma + 1), oper);
    } else pushDep(br, index, argument, oper);
    br[index].is_decomposed = true;
    return true;
  }

  void decomposeFunctions() {
    for (size_t fi = 0; fi < forests.size(); fi++) {
      auto& F = forests[fi];
      const bool side = (F.location == "side ip");
      for (size_t ti = 0; ti < F.trees.size(); ti++) {
        auto& br = F.trees[ti].branches;
        bool done = false; int iter = 0; const int maxiter = 20;
        while (!done && iter < maxiter) {
          iter++;
          const size_t Nb = br.size();
          for (size_t k = 0; k < Nb; k++) {
            if (br[k].is_leaf || br[k].is_decomposed) continue;
            const std::string expr = br[k].expr;
            bool decompose = true;
            { const int j = wkset->findField(expr, side);
              if (j >= 0) { decompose = false; br[k].is_leaf = br[k].is_decomposed = br[k].is_view = br[k].is_AD = br[k].is_workset_data = true; br[k].workset_data_index = j; } }
            if (decompose) {
              auto& L = side ? wkset->side_scalar_fields : wkset->scalar_fields;
              for (size_t j = 0; j < L.size() && decompose; ++j) if (expr == L[j].expr) {
                decompose = false; br[k].is_leaf = br[k].is_decomposed = br[k].is_view = br[k].is_workset_data = true; br[k].is_AD = false; br[k].workset_data_index = (int)j; }
            }
            // parameters: `name` is component 0, `name(k)` / `name(kk)` component k (functionManager.cpp:270-355).  During
            // assembleJacRes the parameters are not seeded with respect to the solution, so the leaf is a constant of the pass
            if (decompose) {
              double pv = 0.0;
              if (findParameter(expr, pv)) { br[k].is_leaf = br[k].is_decomposed = br[k].is_constant = true; br[k].sSc = pv; decompose = false; }
            }
            if (decompose) for (size_t j = 0; j < F.trees.size(); j++) if (expr == F.trees[j].name) { br[k].is_decomposed = true; br[k].is_func = true; br[k].func_index = (int)j; decompose = false; }
            if (decompose && interp_isScalar(expr)) { br[k].is_leaf = br[k].is_decomposed = br[k].is_constant = true; br[k].sSc = std::stod(expr); decompose = false; }
            if (decompose && expr == "t") { decompose = false; br[k].is_leaf = br[k].is_decomposed = br[k].is_time = true; br[k].sSc = wkset->time; }
            if (decompose && expr == "pi") { decompose = false; br[k].is_leaf = br[k].is_decomposed = br[k].is_constant = true; br[k].sSc = M_PI; }
            if (decompose && isOperator(br, k)) decompose = false;
            if (decompose) { split(br, k); br[k].is_decomposed = true; }
          }
          done = true;
          for (auto& b : br) if (!b.is_leaf && !b.is_decomposed) done = false;
        }
        if (!done) throw std::runtime_error("Error: MrHyDE was not able to decompose " + F.trees[ti].name);
      }
    }
    for (size_t f = 0; f < forests.size(); ++f)
      for (size_t k = 0; k < forests[f].trees.size(); k++)
        for (size_t j = 0; j < forests[f].trees[k].branches.size(); j++) {
          bool isConst = true, isView = false, isAD = false;
          checkDepDataType(f, k, j, isConst, isView, isAD, 0);
          auto& b = forests[f].trees[k].branches[j];
          b.is_constant = isConst; b.is_view = isView; b.is_AD = isAD;
          if (isView && !b.is_workset_data) { if (isAD) b.vAD.resize(forests[f].dim0, forests[f].dim1); else b.vSc.resize(forests[f].dim0, forests[f].dim1); }
        }
    for (size_t f = 0; f < forests.size(); ++f)
      for (size_t k = 0; k < forests[f].trees.size(); k++)
        for (size_t j = 0; j < forests[f].trees[k].branches.size(); j++) {
          auto& b = forests[f].trees[k].branches[j];
          if (b.is_constant && !b.is_leaf) evaluate(f, k, j);
        }
    decomposed = true;
  }

  void checkDepDataType(size_t f, size_t t, size_t bi, bool& isConst, bool& isView, bool& isAD, int depth) {
    if (depth > 64) throw std::runtime_error("oracle funcman: function recursion too deep");
    auto& b = forests[f].trees[t].branches[bi];
    if (b.is_leaf) { if (!b.is_constant) isConst = false; if (b.is_view) isView = true; if (b.is_AD) isAD = true; }
    else if (b.is_func) checkDepDataType(f, b.func_index, 0, isConst, isView, isAD, depth + 1);
    else for (int d : b.dep_list) checkDepDataType(f, t, d, isConst, isView, isAD, depth + 1);
  }

  Vista<EvalT> evaluate(const std::string& fname, const std::string& location) {
    if (!decomposed) decomposeFunctions();
    for (size_t f = 0; f < forests.size(); ++f) if (forests[f].location == location)
      for (size_t t = 0; t < forests[f].trees.size(); ++t) if (forests[f].trees[t].name == fname) {
        auto& b0 = forests[f].trees[t].branches[0];
        if (!b0.is_constant) evaluate(f, t, 0);
        Vista<EvalT> v; v.is_view = b0.is_view; v.is_AD = b0.is_AD;
        v.sSc = b0.sSc; v.sAD = b0.sAD; v.vSc = viewSc(b0, location == "side ip"); v.vAD = viewAD(b0, location == "side ip");
        return v;
      }
    throw std::runtime_error("Error: function manager could not evaluate: " + fname + " at " + location);
  }

 private:
  bool decomposed = false;

  const Arr2<double>* viewSc(const Branch<EvalT>& b, bool side) const {
    if (b.is_workset_data && !b.is_AD) return &(side ? wkset->side_scalar_fields : wkset->scalar_fields)[b.workset_data_index].data;
    return &b.vSc;
  }
  const Arr2<EvalT>* viewAD(const Branch<EvalT>& b, bool side) const {
    if (b.is_workset_data && b.is_AD) return &(side ? wkset->side_soln_fields : wkset->soln_fields)[b.workset_data_index].data;
    return &b.vAD;
  }

  template <class D, class S>
  static void op_point(D& data, const S& t, const std::string& op) {
    if (op == "") data = D(t);
    else if (op == "plus") data += t;
    else if (op == "minus") data += -t;
    else if (op == "times") data *= t;
    else if (op == "divide") data /= t;
    else if (op == "power") data = pow(data, t);
    else if (op == "sin") data = D(sin(t));
    else if (op == "cos") data = D(cos(t));
    else if (op == "tan") data = D(tan(t));
    else if (op == "exp") data = D(exp(t));
    else if (op == "log") data = D(log(t));
    else if (op == "sqrt") { if (t <= 0.0) data = D(0.0); else data = D(sqrt(t)); }   // functionManager.cpp:1062-1067, 1258-1262: 0 for arguments <= 0
    else if (op == "sinh") data = D(sinh(t));
    else if (op == "cosh") data = D(cosh(t));
    else if (op == "abs") { if (t < 0.0) data = D(-t); else data = D(t); }
    else if (op == "max") { if (data < t) data = D(t); }
    else if (op == "min") { if (data > t) data = D(t); }
    else if (op == "mean") data = 0.5 * data + 0.5 * D(t);
    else if (op == "lt") data = (data < t) ? D(1.0) : D(0.0);
    else if (op == "lte") data = (data <= t) ? D(1.0) : D(0.0);
    else if (op == "gt") data = (data > t) ? D(1.0) : D(0.0);
    else if (op == "gte") data = (data >= t) ? D(1.0) : D(0.0);
    else throw std::runtime_error("oracle funcman: unknown operator " + op);
  }
  // element-wide operators on views (functionManager.cpp "emax"/"emin"/"emean"), quirks kept
  template <class AD_, class AS_>
  static void op_elem(AD_& data, const AS_& t, int ne, int np, const std::string& op) {
    for (int e = 0; e < ne; ++e) {
      if (op == "emax") { data(e, 0) = t(e, 0); for (int n = 0; n < np; ++n) if (t(e, n) > t(e, 0)) data(e, 0) = t(e, n); }
      else if (op == "emin") { data(e, 0) = t(e, 0); for (int n = 0; n < np; ++n) if (t(e, n) < t(e, 0)) data(e, 0) = t(e, n); }
      else { const double scale = (double)np; data(e, 0) = t(e, 0) / scale; for (int n = 0; n < np; ++n) data(e, 0) += t(e, n) / scale; }
      for (int n = 1; n < np; ++n) data(e, n) = data(e, 0);
    }
  }

  void evaluate(size_t f, size_t t, size_t bi) {
    auto& br = forests[f].trees[t].branches;
    const bool side = (forests[f].location == "side ip");
    Branch<EvalT>& b = br[bi];
    if (b.is_leaf) {
      if (b.is_workset_data && b.is_AD) {
        auto& L = side ? wkset->side_soln_fields : wkset->soln_fields;
        if (!L[b.workset_data_index].updated) wkset->evaluateField(L[b.workset_data_index]);
      } else if (b.is_time) b.sSc = wkset->time;
      return;
    }
    if (b.is_func) {
      evaluate(f, b.func_index, 0);
      const auto& r = forests[f].trees[b.func_index].branches[0];
      if (b.is_AD) { if (b.is_view) b.vAD = *viewAD(r, side); else b.sAD = r.sAD; }
      else { if (b.is_view) b.vSc = *viewSc(r, side); else b.sSc = r.sSc; }
      return;
    }
    const int ne = wkset->numElem, np = forests[f].dim1;
    for (size_t k = 0; k < b.dep_list.size(); k++) {
      const int dep = b.dep_list[k];
      evaluate(f, t, dep);
      const Branch<EvalT>& d = forests[f].trees[t].branches[dep];
      Branch<EvalT>& bb = forests[f].trees[t].branches[bi];   // re-fetch (vector may not move, but be safe)
      const std::string& op = bb.dep_ops[k];
      const bool elemop = (op == "emax" || op == "emin" || op == "emean");
      if (bb.is_view) {
        if (d.is_view) {
          if (bb.is_AD) {
            if (d.is_AD) { const auto& src = *viewAD(d, side); if (elemop) op_elem(bb.vAD, src, ne, np, op); else for (int e = 0; e < ne; ++e) for (int q = 0; q < np; ++q) op_point(bb.vAD(e, q), src(e, q), op); }
            else { const auto& src = *viewSc(d, side); if (elemop) op_elem(bb.vAD, src, ne, np, op); else for (int e = 0; e < ne; ++e) for (int q = 0; q < np; ++q) op_point(bb.vAD(e, q), src(e, q), op); }
          } else if (!d.is_AD) {
            const auto& src = *viewSc(d, side);
            if (elemop) op_elem(bb.vSc, src, ne, np, op); else for (int e = 0; e < ne; ++e) for (int q = 0; q < np; ++q) op_point(bb.vSc(e, q), src(e, q), op);
          }
        } else {
          if (bb.is_AD) {
            if (d.is_AD) { for (int e = 0; e < ne; ++e) for (int q = 0; q < np; ++q) op_point(bb.vAD(e, q), d.sAD, op); }
            else { for (int e = 0; e < ne; ++e) for (int q = 0; q < np; ++q) op_point(bb.vAD(e, q), d.sSc, op); }
          } else if (!d.is_AD) { for (int e = 0; e < ne; ++e) for (int q = 0; q < np; ++q) op_point(bb.vSc(e, q), d.sSc, op); }
        }
      } else if (!d.is_view) {
        if (bb.is_AD) { if (d.is_AD) op_point(bb.sAD, d.sAD, op); else op_point(bb.sAD, d.sSc, op); }
        else if (!d.is_AD) op_point(bb.sSc, d.sSc, op);
      }
    }
  }
};

}  // namespace orc
