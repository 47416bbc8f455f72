// ORACLE (test infrastructure, NOT product code).
// AssemblyManager restatement: element batches ("groups"), gather -> seed -> physics -> scatter, DBC fix-up.
//   createGroups            /root/reference/src/managers/assemblyManager.cpp:272-712
//   allocateGroupStorage    :720-760 -> Group::computeBasis /root/reference/src/tools/group.cpp:137-245 (storeAll)
//   createFixedDOFs         :185-265
//   assembleJacRes<EvalT>   :2150-2665      assembleRes :2946-3150
//   performGather           :3598-3643      scatter :4031-4145     scatterRes :4154-4202
//   dofConstraints          :3158-3188      updateJacDBC :1167-1180   fix-zero-rows :2644-2663
//   type_AD selection       :131-160
//   BC function names       /root/reference/src/interfaces/physicsInterface.cpp:156-478
//   var_bcs strings         /root/reference/src/interfaces/discretizationInterface.cpp:2556-2593
#pragma once
#include <atomic>
#include <chrono>
#include <mutex>
#include <sstream>
#include <thread>
#include "basis.hpp"
#include "mesh.hpp"
#include "physics.hpp"

namespace orc {

struct Config {
  int dim = 3; int N[3] = {1, 1, 1}; double lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
  std::string physics = "thermal";
  int order = 1, quadrature = -1, side_quadrature = -1, workset = 100;
  bool transient = false, weak_dbc = false, strong_dbc = true, fix_zero_rows = false, lump_mass = false;
  StrMap functions, settings;
  std::vector<std::pair<std::string, std::vector<double>>> parameters;      // "parameter.<name>: v0 v1 ..." (scalar / vector parameters, constants of the pass)
  std::map<std::string, StrMap> dirichlet, neumann, flux;   // [var][side or "all boundaries"] -> expr

  static std::string trim(const std::string& s) {
    size_t a = s.find_first_not_of(" \t\r"), b = s.find_last_not_of(" \t\r");
    return (a == std::string::npos) ? "" : s.substr(a, b - a + 1);
  }
  void parse(const std::string& text) {
    std::istringstream in(text); std::string line;
    while (std::getline(in, line)) {
      const size_t c = line.find(':'); if (c == std::string::npos) continue;
      const std::string k = trim(line.substr(0, c)), v = trim(line.substr(c + 1));
      if (k.empty() || k[0] == '#') continue;
      auto pre = [&](const char* p) { return k.compare(0, strlen(p), p) == 0; };
      if (k == "dim") dim = std::stoi(v);
      else if (k == "NX") N[0] = std::stoi(v); else if (k == "NY") N[1] = std::stoi(v); else if (k == "NZ") N[2] = std::stoi(v);
      else if (k == "xmin") lo[0] = std::stod(v); else if (k == "xmax") hi[0] = std::stod(v);
      else if (k == "ymin") lo[1] = std::stod(v); else if (k == "ymax") hi[1] = std::stod(v);
      else if (k == "zmin") lo[2] = std::stod(v); else if (k == "zmax") hi[2] = std::stod(v);
      else if (k == "physics") physics = v;
      else if (k == "order") order = std::stoi(v);
      else if (k == "quadrature") quadrature = std::stoi(v);
      else if (k == "side quadrature") side_quadrature = std::stoi(v);
      else if (k == "workset size") workset = std::stoi(v);
      else if (k == "transient") transient = (v == "true");
      else if (k == "use weak Dirichlet") weak_dbc = (v == "true");
      else if (k == "use strong DBCs") strong_dbc = (v == "true");
      else if (k == "fix zero rows") fix_zero_rows = (v == "true");
      else if (k == "lump mass") lump_mass = (v == "true");
      else if (pre("function.")) functions[k.substr(9)] = v;
      else if (pre("setting.")) settings[k.substr(8)] = v;
      else if (pre("parameter.")) {      // (strtod, not operator>>: a process that loaded another libstdc++ first must not resolve our stream extraction into it)
        std::vector<double> vals; const char* s = v.c_str(); char* end = nullptr;
        for (double x = strtod(s, &end); end != s; x = strtod(s, &end)) { vals.push_back(x); s = end; }
        parameters.push_back({k.substr(10), vals}); }
      else if (pre("dirichlet.") || pre("neumann.") || pre("flux.")) {
        const size_t d1 = k.find('.'), d2 = k.find('.', d1 + 1);
        const std::string var = k.substr(d1 + 1, d2 - d1 - 1), side = k.substr(d2 + 1);
        if (pre("dirichlet.")) dirichlet[var][side] = v; else if (pre("neumann.")) neumann[var][side] = v; else flux[var][side] = v;
      }
      else throw std::runtime_error("oracle config: unknown key '" + k + "'");
    }
    if (quadrature < 0) quadrature = 2 * order;
    if (side_quadrature < 0) side_quadrature = 2 * order;
  }
};

inline void atomic_add(double* p, double v) {      // Kokkos::atomic_add on the host
  uint64_t* q = reinterpret_cast<uint64_t*>(p);
  uint64_t old = __atomic_load_n(q, __ATOMIC_RELAXED), want;
  do { double d; memcpy(&d, &old, 8); d += v; memcpy(&want, &d, 8); }
  while (!__atomic_compare_exchange_n(q, &old, want, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED));
}

struct TimeIntegration {
  int nstages = 1, nsteps = 1, stage = 0;
  std::vector<double> A{1.0}, b{1.0}, bdf{1.0, -1.0};   // backward Euler
  double dt = 1.0, time = 0.0;
};

class Oracle {
 public:
  Config cfg; BrickMesh mesh; LocalLayout lay; DofMap dofs; CrsGraph crs; RefData ref;
  std::vector<std::string> varlist;
  std::vector<std::vector<std::string>> var_bcs;   // [var][side]
  std::vector<uint8_t> isFixed;
  std::vector<std::vector<int>> dbc_dofs;           // [var] sorted unique LIDs
  std::vector<GroupData> groups;                    // volume batches
  struct BGroup { GroupData g; int side; int local_side; };
  std::vector<BGroup> bgroups;
  StrMap allFunctions;                               // user + BC functions: name -> expr ; side-ip BC functions
  std::vector<std::pair<std::string, std::string>> sideFunctions;
  TimeIntegration ti;
  int type_AD = 0;
  double last_seconds[4] = {0, 0, 0, 0};            // gather+seed, physics, scatter, total

  explicit Oracle(const std::string& text) {
    cfg.parse(text);
    mesh.build(cfg.dim, cfg.N, cfg.lo, cfg.hi);
    lay.build(cfg.dim, cfg.order);
    { auto m = importPhysics<double>(cfg.physics, cfg.settings, cfg.dim); varlist = m->myvars; }
    dofs.build(mesh, lay, (int)varlist.size());
    crs.build(dofs, mesh.nelem);
    ref.build(cfg.dim, cfg.order, cfg.quadrature, cfg.side_quadrature);
    setBCData(); createFixedDOFs(); createGroups();
    const int ndr = dofs.ndof_elem;
    const int sizes[7] = {2, 4, 8, 16, 18, 24, 32};
    type_AD = -1; for (int s : sizes) if (ndr <= s) { type_AD = s; break; }
  }

  // Replace the vertex coordinates (same topology): distorted / non-affine test meshes.  Stored physical bases are rebuilt,
  // as Group::computeBasis would after a mesh update (/root/reference/src/tools/group.cpp:137-245).
  void setNodes(const double* xyz) {
    mesh.nodes.assign(xyz, xyz + (size_t)mesh.nnode * mesh.dim);
    groups.clear(); bgroups.clear(); createGroups();
  }

  int varIndex(const std::string& v) const { for (size_t i = 0; i < varlist.size(); ++i) if (varlist[i] == v) return (int)i; return -1; }

  void setBCData() {
    var_bcs.assign(varlist.size(), std::vector<std::string>(mesh.sides.size(), "none"));
    for (size_t j = 0; j < varlist.size(); ++j)
      for (size_t s = 0; s < mesh.sides.size(); ++s) {
        const std::string& sn = mesh.sides[s].name; const std::string& var = varlist[j];
        auto has = [&](const std::map<std::string, StrMap>& m, std::string& expr) {
          auto it = m.find(var); if (it == m.end()) return false;
          auto a = it->second.find(sn); if (a != it->second.end()) { expr = a->second; return true; }
          a = it->second.find("all boundaries"); if (a != it->second.end()) { expr = a->second; return true; }
          return false;
        };
        std::string ex;
        if (has(cfg.dirichlet, ex)) { var_bcs[j][s] = cfg.weak_dbc ? "weak Dirichlet" : "Dirichlet"; sideFunctions.push_back({"Dirichlet " + var + " " + sn, ex}); }
        if (has(cfg.neumann, ex)) { var_bcs[j][s] = "Neumann"; sideFunctions.push_back({"Neumann " + var + " " + sn, ex}); }
        if (has(cfg.flux, ex)) { var_bcs[j][s] = "Flux"; sideFunctions.push_back({"Flux " + var + " " + sn, ex}); }
      }
  }

  void createFixedDOFs() {
    isFixed.assign((size_t)dofs.nlocal, 0);
    dbc_dofs.assign(varlist.size(), {});
    if (!cfg.strong_dbc) return;
    for (size_t j = 0; j < varlist.size(); ++j) {
      auto& vd = dbc_dofs[j];
      for (size_t s = 0; s < mesh.sides.size(); ++s) {
        if (var_bcs[j][s] != "Dirichlet") continue;
        const auto sd = lay.side_dofs(mesh.sides[s].local_side);
        for (int e : mesh.sides[s].elems)
          for (int o : sd) vd.push_back(dofs.lids[(size_t)e * dofs.ndof_elem + dofs.offsets[j * dofs.nb + o]]);
      }
      std::sort(vd.begin(), vd.end()); vd.erase(std::unique(vd.begin(), vd.end()), vd.end());
      for (int l : vd) isFixed[l] = 1;
    }
  }

  // physical basis, weights, ips of a batch (volume: side<0)
  void fillGroup(GroupData& g, const std::vector<int>& elems, int side) {
    const int dim = cfg.dim, nb = ref.nb, nv = ref.nv;
    const int npt = (side < 0) ? ref.nq : ref.nqs;
    const std::vector<double>& B = (side < 0) ? ref.B : ref.sB[side];
    const std::vector<double>& G = (side < 0) ? ref.G : ref.sG[side];
    const std::vector<double>& VB = (side < 0) ? ref.VB : ref.sVB[side];
    const std::vector<double>& VG = (side < 0) ? ref.VG : ref.sVG[side];
    g.numElem = (int)elems.size(); g.nb = nb; g.nq = npt; g.dim = dim; g.elems = elems;
    g.basis.resize((size_t)g.numElem * nb * npt); g.basis_grad.resize((size_t)g.numElem * nb * npt * dim);
    g.wts.resize((size_t)g.numElem * npt);
    for (int c = 0; c < dim; ++c) { g.ipx[c].resize((size_t)g.numElem * npt); if (side >= 0) { g.nrm[c].resize((size_t)g.numElem * npt); g.tng[c].resize((size_t)g.numElem * npt); } }
    for (int le = 0; le < g.numElem; ++le) {
      const int e = elems[le];
      double xv[24];
      for (int a = 0; a < nv; ++a) for (int c = 0; c < dim; ++c) xv[a * dim + c] = mesh.nodes[(size_t)mesh.elem2node[(size_t)e * nv + a] * dim + c];
      for (int q = 0; q < npt; ++q) {
        double J[9], Ji[9];
        jacobian_at(dim, nv, npt, q, xv, VG.data(), J);
        const double det = det_inv(dim, J, Ji);
        for (int c = 0; c < dim; ++c) {            // mapToPhysicalFrame
          double x = 0.0; for (int a = 0; a < nv; ++a) x += xv[a * dim + c] * VB[(size_t)a * npt + q];
          g.ipx[c][(size_t)le * npt + q] = x;
        }
        for (int i = 0; i < nb; ++i) {
          g.basis[((size_t)le * nb + i) * npt + q] = B[(size_t)i * npt + q];
          for (int c = 0; c < dim; ++c) {            // grad = J^{-T} ref_grad
            double s = 0.0; for (int m = 0; m < dim; ++m) s += Ji[m * dim + c] * G[((size_t)i * npt + q) * dim + m];
            g.basis_grad[(((size_t)le * nb + i) * npt + q) * dim + c] = s;
          }
        }
        if (side < 0) g.wts[(size_t)le * npt + q] = std::fabs(det) * ref.wts[q];
        else {
          double n[3] = {0, 0, 0}, t[3] = {0, 0, 0}, measure;
          if (dim == 2) {
            for (int i = 0; i < 2; ++i) { t[i] = 0.0; for (int j = 0; j < 2; ++j) t[i] += J[i * 2 + j] * ref.tanU[side][j]; }
            n[0] = t[1]; n[1] = -t[0];
            measure = std::sqrt(t[0] * t[0] + t[1] * t[1]);
          } else {
            double tu[3], tv[3];
            for (int i = 0; i < 3; ++i) { tu[i] = tv[i] = 0.0; for (int j = 0; j < 3; ++j) { tu[i] += J[i * 3 + j] * ref.tanU[side][j]; tv[i] += J[i * 3 + j] * ref.tanV[side][j]; } }
            n[0] = tu[1] * tv[2] - tu[2] * tv[1]; n[1] = tu[2] * tv[0] - tu[0] * tv[2]; n[2] = tu[0] * tv[1] - tu[1] * tv[0];
            measure = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
          }
          g.wts[(size_t)le * npt + q] = measure * ref.swts[q];
          double nl = 0.0; for (int c = 0; c < dim; ++c) nl += n[c] * n[c]; nl = std::sqrt(nl);
          for (int c = 0; c < dim; ++c) { g.nrm[c][(size_t)le * npt + q] = n[c] * (1.0 / nl); g.tng[c][(size_t)le * npt + q] = t[c]; }
        }
      }
    }
  }

  void createGroups() {
    const int per = (cfg.workset <= 0) ? (int)mesh.nelem : cfg.workset;
    for (int64_t prog = 0; prog < mesh.nelem; prog += per) {
      std::vector<int> el; for (int64_t e = prog; e < std::min<int64_t>(mesh.nelem, prog + per); ++e) el.push_back((int)e);
      groups.emplace_back(); fillGroup(groups.back(), el, -1);
    }
    for (size_t s = 0; s < mesh.sides.size(); ++s) {
      const auto& S = mesh.sides[s];
      for (size_t prog = 0; prog < S.elems.size(); prog += per) {
        std::vector<int> el(S.elems.begin() + prog, S.elems.begin() + std::min(S.elems.size(), prog + (size_t)per));
        bgroups.emplace_back(); bgroups.back().side = (int)s; bgroups.back().local_side = S.local_side;
        fillGroup(bgroups.back().g, el, S.local_side);
      }
    }
  }

  // ------------------------------------------------------------------------------------------------
  template <class EvalT>
  void assembleImpl(const double* u, const double* u_prev, const double* u_stage, bool compute_jacobian,
                    double* res_out, double* jac_out, int nthreads) {
    using clk = std::chrono::steady_clock;
    const auto T0 = clk::now();
    double tg_all = 0, tp_all = 0, ts_all = 0;
    if (nthreads < 1) nthreads = 1;
    const bool atomic = nthreads > 1;          // reference: useAtomics when the execution space is concurrent (assemblyManager.cpp:4058-4061)
    std::string fail;
    // One Workset / FunctionManager / module per thread (the reference has one per block and lets Kokkos::OpenMP thread over the
    // elements inside each kernel; threading over whole groups keeps the per-element operation sequence identical).
    // (std::thread, not OpenMP: the image has no libgomp.)
    std::atomic<size_t> next_group{0}, next_bgroup{0};
    std::mutex mu;
    auto worker = [&]() {
    try {
    double tg = 0, tp = 0, ts = 0;
    const int nvar = (int)varlist.size(), nb = ref.nb, N = dofs.ndof_elem;
    const int maxElem = (cfg.workset <= 0) ? (int)mesh.nelem : cfg.workset;
    Workset<EvalT> wk; wk.setup(cfg.dim, varlist, nb, dofs.offsets, maxElem, ref.nq, ref.nqs, cfg.transient);
    wk.var_bcs = var_bcs; wk.time = ti.time; wk.deltat = ti.dt; wk.current_stage = ti.stage;
    wk.butcher_A = ti.A; wk.butcher_b = ti.b; wk.BDF_wts = ti.bdf; wk.nstages = ti.nstages;
    FunctionManager<EvalT> fm(maxElem, ref.nq, ref.nqs); fm.wkset = &wk;
    auto module = importPhysics<EvalT>(cfg.physics, cfg.settings, cfg.dim);
    fm.parameters = cfg.parameters;
    module->defineFunctions(cfg.functions, &fm);
    for (auto& f : sideFunctions) fm.addFunction(f.first, f.second, "side ip");
    // every entry of the user "Functions" list, both locations (physicsInterface.cpp:466-483)
    for (auto& f : cfg.functions) { fm.addFunction(f.first, f.second, "ip"); fm.addFunction(f.first, f.second, "side ip"); }
    module->setWorkset(&wk);
    fm.decomposeFunctions();
    const int seedwhat = compute_jacobian ? 1 : 0;
    std::vector<double> gu((size_t)maxElem * N), gprev, gstage;
    if (cfg.transient) { gprev.resize((size_t)maxElem * N * ti.nsteps); gstage.resize((size_t)maxElem * N * ti.nstages); }

    auto gather = [&](const GroupData& g) {
      for (int le = 0; le < g.numElem; ++le) {
        const int* l = &dofs.lids[(size_t)g.elems[le] * N];
        for (int v = 0; v < nvar; ++v) for (int k = 0; k < nb; ++k) {
          const int lid = l[dofs.offsets[v * nb + k]];
          const size_t id = ((size_t)le * nvar + v) * nb + k;
          gu[id] = u[lid];
          if (cfg.transient) {
            for (int s = 0; s < ti.nsteps; ++s) gprev[id * ti.nsteps + s] = u_prev[(size_t)lid * ti.nsteps + s];
            for (int s = 0; s < ti.nstages; ++s) gstage[id * ti.nstages + s] = u_stage[(size_t)lid * ti.nstages + s];
          }
        }
      }
    };
    auto seed = [&]() {
      if (cfg.transient) wk.computeSolnTransientSeeded(gu, gprev, ti.nsteps, gstage, seedwhat);
      else wk.computeSolnSteadySeeded(gu, seedwhat);
    };
    auto scatter = [&](const GroupData& g) {
      for (int le = 0; le < g.numElem; ++le) {
        const int* l = &dofs.lids[(size_t)g.elems[le] * N];
        for (int n = 0; n < nvar; ++n) for (int j = 0; j < nb; ++j) {
          const int row = dofs.offsets[n * nb + j], rowIndex = l[row];
          if (isFixed[rowIndex]) continue;
          const double rv = -value_of(wk.res(le, row));
          if (atomic) atomic_add(&res_out[rowIndex], rv); else res_out[rowIndex] += rv;
        }
        if constexpr (ad_traits<EvalT>::lanes > 0) {
          if (compute_jacobian && jac_out) {
            for (int n = 0; n < nvar; ++n) for (int j = 0; j < nb; ++j) {
              const int row = dofs.offsets[n * nb + j], rowIndex = l[row];
              if (isFixed[rowIndex]) continue;
              // sumIntoValues(rowIndex, cols[], N, vals[], unsorted): linear search per entry
              const int64_t rb = crs.rowptr[rowIndex], re = crs.rowptr[rowIndex + 1];
              for (int m = 0; m < nvar; ++m) for (int k = 0; k < nb; ++k) {
                const int col = dofs.offsets[m * nb + k];
                const int cidx = cfg.lump_mass ? rowIndex : l[col];
                const double val = wk.res(le, row).fastAccessDx(col);
                for (int64_t p = rb; p < re; ++p) if (crs.colind[p] == cidx) {
                  if (atomic) atomic_add(&jac_out[p], val); else jac_out[p] += val;
                  break;
                }
              }
            }
          }
        }
      }
    };

    for (size_t gi = next_group++; gi < groups.size(); gi = next_group++) {
      const auto& g = groups[gi];
      auto t0 = clk::now();
      wk.attach(&g, false); gather(g); wk.resetResidual(); seed();
      auto t1 = clk::now();
      module->volumeResidual();
      auto t2 = clk::now();
      scatter(g);
      auto t3 = clk::now();
      tg += std::chrono::duration<double>(t1 - t0).count(); tp += std::chrono::duration<double>(t2 - t1).count(); ts += std::chrono::duration<double>(t3 - t2).count();
    }
    for (size_t bi = next_bgroup++; bi < bgroups.size(); bi = next_bgroup++) {
      const auto& bg = bgroups[bi];
      auto t0 = clk::now();
      wk.attach(&bg.g, true); wk.sidename = mesh.sides[bg.side].name; wk.currentside = bg.side;
      gather(bg.g); wk.resetResidual(); seed();
      auto t1 = clk::now();
      module->boundaryResidual();
      fluxConditions(wk, fm);
      auto t2 = clk::now();
      scatter(bg.g);
      auto t3 = clk::now();
      tg += std::chrono::duration<double>(t1 - t0).count(); tp += std::chrono::duration<double>(t2 - t1).count(); ts += std::chrono::duration<double>(t3 - t2).count();
    }
    std::lock_guard<std::mutex> lk(mu);
    tg_all += tg; tp_all += tp; ts_all += ts;
    } catch (std::exception& e) { std::lock_guard<std::mutex> lk(mu); fail = e.what(); }
    };
    if (nthreads == 1) worker();
    else { std::vector<std::thread> th; for (int t = 0; t < nthreads; ++t) th.emplace_back(worker); for (auto& x : th) x.join(); }
    if (!fail.empty()) throw std::runtime_error(fail);
    const double tg = tg_all / nthreads, tp = tp_all / nthreads, ts = ts_all / nthreads;
    if (compute_jacobian && jac_out) {
      if (cfg.strong_dbc) for (auto& vd : dbc_dofs) for (int d : vd) { const int64_t p = crs.find(d, d); if (p >= 0) jac_out[p] = 1.0; }
      if (cfg.fix_zero_rows)
        for (int64_t r = 0; r < crs.nrows; ++r) {
          double abssum = 0.0; for (int64_t p = crs.rowptr[r]; p < crs.rowptr[r + 1]; ++p) abssum += std::fabs(jac_out[p]);
          if (abssum < 1.0e-14) { const int64_t p = crs.find((int)r, (int)r); if (p >= 0) jac_out[p] = 1.0; }
        }
    }
    last_seconds[0] = tg; last_seconds[1] = tp; last_seconds[2] = ts;
    last_seconds[3] = std::chrono::duration<double>(clk::now() - T0).count();
  }

  // AssemblyManager::setDirichlet (/root/reference/src/managers/assemblyManager.cpp:1855-1948) with getDirichletBoundary (:6288-6357)
  // and getMassBoundary (:6360-6420): the L2-projection system of the strong Dirichlet data.  Rows of fixed DOFs collect the side
  // mass matrix and the side moments of the Dirichlet function, every other row is the identity; rhs / mass are accumulated into
  // (caller zeroes).  The solve (SolverManager::projectDirichlet, solverManager.cpp:1060-1098) is the caller's.
  void assembleDirichlet(double time, double* rhs, double* mass) {
    const int nvar = (int)varlist.size(), nb = ref.nb, N = dofs.ndof_elem;
    const int maxElem = (cfg.workset <= 0) ? (int)mesh.nelem : cfg.workset;
    Workset<double> wk; wk.setup(cfg.dim, varlist, nb, dofs.offsets, maxElem, ref.nq, ref.nqs, cfg.transient);
    wk.var_bcs = var_bcs; wk.time = time;
    FunctionManager<double> fm(maxElem, ref.nq, ref.nqs); fm.wkset = &wk;
    auto module = importPhysics<double>(cfg.physics, cfg.settings, cfg.dim);
    fm.parameters = cfg.parameters;
    module->defineFunctions(cfg.functions, &fm);
    for (auto& f : sideFunctions) fm.addFunction(f.first, f.second, "side ip");
    for (auto& f : cfg.functions) { fm.addFunction(f.first, f.second, "ip"); fm.addFunction(f.first, f.second, "side ip"); }
    module->setWorkset(&wk);
    fm.decomposeFunctions();
    std::vector<double> zero((size_t)maxElem * N, 0.0);
    for (const auto& bg : bgroups) {
      const GroupData& g = bg.g;
      wk.attach(&g, true); wk.sidename = mesh.sides[bg.side].name; wk.currentside = bg.side;
      wk.computeSolnSteadySeeded(zero, 0);
      const int nq = g.nq;
      for (int n = 0; n < nvar; ++n) {
        if (var_bcs[n][bg.side] != "Dirichlet") continue;
        auto dip = fm.evaluate("Dirichlet " + varlist[n] + " " + wk.sidename, "side ip");
        for (int e = 0; e < g.numElem; ++e) {
          const int* l = &dofs.lids[(size_t)g.elems[e] * N];
          for (int i = 0; i < nb; ++i) {
            const int rowIndex = l[dofs.offsets[n * nb + i]];
            if (!isFixed[rowIndex]) continue;
            double dval = 0.0;
            for (int j = 0; j < nq; ++j) dval += dip(e, j) * g.B(e, i, j) * g.wts[(size_t)e * nq + j];
            rhs[rowIndex] += dval;
            for (int k = 0; k < nb; ++k) {
              double m = 0.0;
              for (int j = 0; j < nq; ++j) m += g.B(e, i, j) * g.B(e, k, j) * g.wts[(size_t)e * nq + j];
              const int64_t pos = crs.find(rowIndex, l[dofs.offsets[n * nb + k]]);
              if (pos >= 0) mass[pos] += m;
            }
          }
        }
      }
    }
    for (int64_t r = 0; r < crs.nrows; ++r) if (!isFixed[r]) { const int64_t pos = crs.find((int)r, (int)r); if (pos >= 0) mass[pos] = 1.0; }
  }

  template <class EvalT>
  void fluxConditions(Workset<EvalT>& wk, FunctionManager<EvalT>& fm) {
    const int nb = wk.nb, nq = wk.gd->nq;
    for (size_t var = 0; var < varlist.size(); ++var) {
      if (wk.var_bcs[var][wk.currentside] != "Flux") continue;
      auto fluxvals = fm.evaluate("Flux " + varlist[var] + " " + wk.sidename, "side ip");
      for (int elem = 0; elem < wk.numElem; ++elem)
        for (int dof = 0; dof < nb; ++dof)
          for (int pt = 0; pt < nq; ++pt)
            wk.res(elem, wk.off((int)var, dof)) += -fluxvals(elem, pt) * wk.wt(elem, pt) * wk.gd->B(elem, dof, pt);
    }
  }

  // res_out / jac_out are ACCUMULATED into (caller zeroes), like the reference's overlapped vector/matrix
  void assemble(const double* u, const double* u_prev, const double* u_stage, bool compute_jacobian,
                double* res_out, double* jac_out, int nthreads) {
    if (!compute_jacobian) { assembleImpl<double>(u, u_prev, u_stage, false, res_out, nullptr, nthreads); return; }
    switch (type_AD) {
      case 2: case 4: assembleImpl<SFad<4>>(u, u_prev, u_stage, true, res_out, jac_out, nthreads); break;
      case 8: assembleImpl<SFad<8>>(u, u_prev, u_stage, true, res_out, jac_out, nthreads); break;
      case 16: assembleImpl<SFad<16>>(u, u_prev, u_stage, true, res_out, jac_out, nthreads); break;
      case 18: case 24: assembleImpl<SFad<24>>(u, u_prev, u_stage, true, res_out, jac_out, nthreads); break;
      case 32: assembleImpl<SFad<32>>(u, u_prev, u_stage, true, res_out, jac_out, nthreads); break;
      default: assembleImpl<SFad<64>>(u, u_prev, u_stage, true, res_out, jac_out, nthreads); break;
    }
  }
};

}  // namespace orc
