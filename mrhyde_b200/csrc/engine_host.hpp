// Host-side state of one engine handle: configuration, mesh/DOF/graph arrays, boundary-condition table,
// FunctionManager-style registry and the compiled coefficient programs.  Mirrors what the reference spreads
// over AssemblyManager (createGroups / createFixedDOFs / finalizeFunctions), PhysicsInterface (module
// defineFunctions + BC function names, /root/reference/src/interfaces/physicsInterface.cpp:156-478) and
// DiscretizationInterface::setBCData (/root/reference/src/interfaces/discretizationInterface.cpp:2481-2739).
// Contains no CUDA; engine.cu adds the device side.
#pragma once
#include <array>
#include <map>
#include <memory>
#include "../../include/mrhyde_b200.h"
#include "brick.hpp"
#include "partition.hpp"
#include "expr.hpp"
#include "params.h"
#include "refdata.hpp"

namespace mrh {

struct SideLaunch {
  int side = 0, local_side = 0;
  bool active = false;
  int bc[4] = {0, 0, 0, 0};
  Prog progs[MRH_MAXPROG];
};

struct BasePtrs {            // where the arrays of a launch live (device pointers in the product)
  const double* nodes = nullptr; const int* e2n = nullptr; const int* lids = nullptr; const uint8_t* fixed = nullptr;
  const int64_t* rowptr = nullptr; const uint16_t* pat = nullptr; const int* pat_id = nullptr;
  const Instr* code = nullptr; const double* cpool = nullptr;
  RefTab vol_tab{}, side_tab[6]{};   // table pointers + nq + tangents, filled by whoever owns the memory
};

class EngineHost {
 public:
  mrh_config cfg{};
  std::string physics, err;
  int phys = 0, dim = 3, order = 1, nvar = 1, nb = 8, nv = 8, N = 8;
  std::vector<std::string> varlist;
  HostMesh mesh; LocalLayout lay; HostDofs dofs; HostCrs crs; RefElement ref;
  Partition part; ExchangePlan plan;       // multi-GPU: block partition and interface exchange (partition.hpp)
  bool have_mesh = false, have_dofs = false, have_crs = false, host_final = false;
  std::map<std::string, std::map<std::string, std::string>> dirichlet, weakd, neumann, flux;   // [var][side] -> expr
  std::vector<std::pair<std::string, std::string>> user_fns;    // name -> expr, first registration wins
  std::vector<std::vector<int>> var_bcs;                         // [var][side]
  std::vector<uint8_t> isFixed;
  std::vector<int> dbc_dofs;                                     // all strong-Dirichlet LIDs, sorted unique
  ExprCompiler ec;
  Prog vol_progs[MRH_MAXPROG];
  std::vector<SideLaunch> side_launches;
  std::vector<SideLaunch> dirichlet_launches;       // sides with a strong Dirichlet variable: L2 projection of the data (setDirichlet)
  Prog init_progs[MRH_MAXPROG];                     // "initial <var>" at "ip" (setInitial), slots VOL_INIT + var
  std::string proj_unsupported;                     // non-empty: why the projection launches cannot run (state-dependent data)

  void configure(const mrh_config& c) {
    cfg = c; physics = c.physics ? c.physics : "";
    dim = c.dim; order = c.order;
    if (dim != 2 && dim != 3) throw std::runtime_error("dim must be 2 or 3");
    if (order != 1 && order != 2) throw std::runtime_error("HGRAD order must be 1 or 2");
    if (cfg.quadrature < 0) cfg.quadrature = 2 * order;
    if (cfg.side_quadrature < 0) cfg.side_quadrature = 2 * order;
    if (physics == "thermal") { phys = PHYS_THERMAL; varlist = {"e"}; }
    else if (physics == "cdr") { phys = PHYS_CDR; varlist = {"c"}; }
    else if (physics == "linearelasticity") { phys = PHYS_LE; varlist = {"dx", "dy", "dz"}; varlist.resize(dim); }
    else if (physics == "navierstokes") { phys = PHYS_NS; varlist = {"ux", "pr", "uy", "uz"}; varlist.resize(dim + 1); }
    else if (physics == "stokes" || physics == "Stokes") { phys = PHYS_STOKES; varlist = {"ux", "pr", "uy", "uz"}; varlist.resize(dim + 1); }
    else if (physics == "helmholtz") { phys = PHYS_HELMHOLTZ; varlist = {"ureal", "uimag"}; }
    else if (physics == "burgers" || physics == "Burgers") { phys = PHYS_BURGERS; varlist = {"u"}; }
    else throw std::runtime_error("physics module not on the hot path: '" + physics + "'");
    nvar = (int)varlist.size();
    lay.build(dim, order);
    nb = lay.nb; nv = (dim == 3) ? 8 : 4; N = nvar * nb;
    ref.build(dim, order, cfg.quadrature, cfg.side_quadrature);
    ec.varlist = varlist; ec.dim = dim;
  }
  int varIndex(const std::string& v) const { for (int i = 0; i < nvar; ++i) if (varlist[i] == v) return i; return -1; }

  void buildBrick(const int* nxyz, const double* lo, const double* hi) {
    build_brick(mesh, dim, nxyz, lo, hi); have_mesh = true;
    build_brick_dofs(dofs, mesh, lay, nvar); have_dofs = true;
    build_crs(crs, dofs.lids.data(), mesh.nelem, N, dofs.nlocal, nvar); have_crs = true;
    host_final = false;
  }

  // block `rank` of a procs[0] x procs[1] x procs[2] decomposition of the global brick (meshInterface.cpp:140-155 X/Y/Z Procs)
  void buildBrickPart(const int* nxyz, const double* lo, const double* hi, const int* procs, int rank) {
    build_brick_part(mesh, dofs, crs, part, lay, dim, nvar, nxyz, lo, hi, procs, rank);
    have_mesh = have_dofs = have_crs = true; host_final = false;
    plan.build_send(part, crs);
  }
  void setPartition(const int64_t* gids, const int* owner, int rank, int world) {
    if (!have_dofs || !have_crs) throw std::runtime_error("dofmap and CRS graph must be set before the partition");
    part.rank = rank; part.world = world;
    part.gids.assign(gids, gids + dofs.nlocal); part.owner.assign(owner, owner + dofs.nlocal);
    for (int o : part.owner) if (o < 0 || o >= world) throw std::runtime_error("owner rank out of range");
    plan.build_send(part, crs);
  }

  // scalar / vector parameter usable as a leaf of coefficient expressions (`name`, `name(k)`; the reference's Parameters block,
  // functionManager.cpp:270-355).  Its values are constants of the assembly pass: set again + mrh_finalize to change them.
  std::vector<std::pair<std::string, std::vector<double>>> parameters;
  void setParameter(const std::string& name, const double* vals, int n) {
    for (auto& pr : parameters) if (pr.first == name) { pr.second.assign(vals, vals + n); return; }
    parameters.push_back({name, std::vector<double>(vals, vals + n)});
  }
  void addUserFunction(const std::string& name, const std::string& expr, const std::string& location) {
    bool seen = false;
    for (auto& f : user_fns) if (f.first == name) seen = true;
    if (!seen) user_fns.push_back({name, expr});
    pending.push_back({name, expr, location});
  }

  void setBC(const std::string& var, const std::string& side, int code, const std::string& expr) {
    if (varIndex(var) < 0) throw std::runtime_error("unknown variable '" + var + "'");
    auto& m = (code == MRH_BC_DIRICHLET) ? dirichlet : (code == MRH_BC_WEAK_DIRICHLET) ? weakd : (code == MRH_BC_NEUMANN) ? neumann : flux;
    if (code < MRH_BC_DIRICHLET || code > MRH_BC_FLUX) throw std::runtime_error("unsupported BC code");
    m[var][side] = expr;
    host_final = false;
  }

  // ---- everything that needs no GPU -----------------------------------------------------------------
  void finalizeHost() {
    if (!have_mesh || !have_dofs || !have_crs) throw std::runtime_error("mesh, dofmap and CRS graph must be set before finalize");
    // fresh registry: module functions first, then BC functions, then the user's list (first wins)
    ec = ExprCompiler(); ec.varlist = varlist; ec.dim = dim; ec.parameters = parameters;
    defineModuleFunctions();
    const int nsides = (int)mesh.sides.size();
    var_bcs.assign(nvar, std::vector<int>(nsides, MRH_BC_NONE));
    for (int j = 0; j < nvar; ++j)
      for (int s = 0; s < nsides; ++s) {
        const std::string& sn = mesh.sides[s].name; const std::string& var = varlist[j];
        std::string ex;
        if (lookup(dirichlet, var, sn, ex)) { var_bcs[j][s] = MRH_BC_DIRICHLET; ec.addFunction("Dirichlet " + var + " " + sn, ex, "side ip"); }
        if (lookup(weakd, var, sn, ex)) { var_bcs[j][s] = MRH_BC_WEAK_DIRICHLET; ec.addFunction("Dirichlet " + var + " " + sn, ex, "side ip"); }
        if (lookup(neumann, var, sn, ex)) { var_bcs[j][s] = MRH_BC_NEUMANN; ec.addFunction("Neumann " + var + " " + sn, ex, "side ip"); }
        if (lookup(flux, var, sn, ex)) { var_bcs[j][s] = MRH_BC_FLUX; ec.addFunction("Flux " + var + " " + sn, ex, "side ip"); }
      }
    for (auto& p : pending) ec.addFunction(p[0], p[1], p[2]);
    // fixed DOFs
    isFixed.assign((size_t)dofs.nlocal, 0); dbc_dofs.clear();
    if (cfg.use_strong_dbc)
      for (int j = 0; j < nvar; ++j)
        for (int s = 0; s < nsides; ++s) {
          if (var_bcs[j][s] != MRH_BC_DIRICHLET) continue;
          const auto sd = lay.side_dofs(mesh.sides[s].local_side);
          for (int e : mesh.sides[s].elems)
            for (int o : sd) dbc_dofs.push_back(dofs.lids[(size_t)e * N + dofs.offsets[(size_t)j * nb + o]]);
        }
    std::sort(dbc_dofs.begin(), dbc_dofs.end());
    dbc_dofs.erase(std::unique(dbc_dofs.begin(), dbc_dofs.end()), dbc_dofs.end());
    for (int l : dbc_dofs) isFixed[l] = 1;
    // compile programs
    ec.code.clear(); ec.cpool.clear();
    Prog zero{0, 0, 0, 0, 0.0};
    for (auto& p : vol_progs) p = zero;
    for (auto& kv : volNames()) vol_progs[kv.first] = ec.compile(kv.second, "ip");
    side_launches.clear();
    for (int s = 0; s < nsides; ++s) {
      SideLaunch L; L.side = s; L.local_side = mesh.sides[s].local_side;
      for (auto& p : L.progs) p = zero;
      for (int j = 0; j < nvar; ++j) {
        L.bc[j] = var_bcs[j][s];
        if (L.bc[j] == MRH_BC_WEAK_DIRICHLET || L.bc[j] == MRH_BC_NEUMANN || L.bc[j] == MRH_BC_FLUX) L.active = true;
      }
      if (phys == PHYS_CDR) {                                  // cdr::boundaryResidual is empty; only Flux acts
        L.active = false;
        for (int j = 0; j < nvar; ++j) if (L.bc[j] == MRH_BC_FLUX) L.active = true;
      }
      if (L.active && !mesh.sides[s].elems.empty()) {
        for (auto& kv : sideNames()) L.progs[kv.first] = ec.compile(kv.second, "side ip");
        const std::string& sn = mesh.sides[s].name;
        for (int j = 0; j < nvar; ++j) {
          if (L.bc[j] == MRH_BC_WEAK_DIRICHLET) L.progs[SIDE_BCSRC + j] = ec.compile("Dirichlet " + varlist[j] + " " + sn, "side ip");
          if (L.bc[j] == MRH_BC_NEUMANN) L.progs[SIDE_BCSRC + j] = ec.compile("Neumann " + varlist[j] + " " + sn, "side ip");
          if (L.bc[j] == MRH_BC_FLUX) L.progs[SIDE_FLUX + j] = ec.compile("Flux " + varlist[j] + " " + sn, "side ip");
        }
      } else L.active = false;
      side_launches.push_back(L);
    }
    // L2-projection launches: strong Dirichlet data per side, initial data over the volume.  The functions are registered by
    // PhysicsInterface under the names "Dirichlet <var> <side>" / "initial <var>" (physicsInterface.cpp:156-478)
    dirichlet_launches.clear(); proj_unsupported.clear();
    for (int s = 0; s < nsides && cfg.use_strong_dbc; ++s) {
      SideLaunch L; L.side = s; L.local_side = mesh.sides[s].local_side;
      for (auto& p : L.progs) p = zero;
      for (int j = 0; j < nvar; ++j) {
        L.bc[j] = var_bcs[j][s];
        if (L.bc[j] != MRH_BC_DIRICHLET) continue;
        L.active = true;
        L.progs[SIDE_BCSRC + j] = ec.compile("Dirichlet " + varlist[j] + " " + mesh.sides[s].name, "side ip");
        if (L.progs[SIDE_BCSRC + j].kind == 2) proj_unsupported = "Dirichlet data that depends on the solution is not supported";
      }
      if (L.active && !mesh.sides[s].elems.empty()) dirichlet_launches.push_back(L);
    }
    for (auto& p : init_progs) p = zero;
    for (int j = 0; j < nvar; ++j) {
      if (!ec.has("initial " + varlist[j], "ip")) continue;       // no initial condition given: zero
      init_progs[VOL_INIT + j] = ec.compile("initial " + varlist[j], "ip");
      if (init_progs[VOL_INIT + j].kind == 2) proj_unsupported = "initial data that depends on the solution is not supported";
    }
    if (ec.code.empty()) ec.code.push_back({(uint8_t)OP_CONST, 0, 0});
    if (ec.cpool.empty()) ec.cpool.push_back(0.0);
    host_final = true;
  }

  // ---- launch parameter blocks ----------------------------------------------------------------------
  void timeParams(TimeParams& tp, const mrh_state& st) const {
    memset(&tp, 0, sizeof(tp));
    tp.transient = st.is_transient ? 1 : 0; tp.time = st.time; tp.dt = (st.deltat != 0.0) ? st.deltat : 1.0;
    tp.alpha_u = 1.0; tp.alpha_t = 0.0; tp.timewt = 0.0;
    if (!tp.transient) return;
    if (!st.butcher_A || !st.butcher_b || !st.bdf_wts || st.nstages < 1 || st.nsteps < 1 || st.nstages > 8 || st.nsteps > 7)
      throw std::runtime_error("transient state needs butcher_A, butcher_b, bdf_wts with nstages<=8, nsteps<=7");
    tp.nstages = st.nstages; tp.nsteps = st.nsteps; tp.stage = st.stage;
    const int s = st.stage, ns = st.nstages;
    tp.alpha_u = st.butcher_A[s * ns + s] / st.butcher_b[s];
    tp.timewt = 1.0 / tp.dt / st.butcher_b[s];
    tp.alpha_t = st.bdf_wts[0] * tp.timewt;
    for (int k = 0; k < s; ++k) tp.a_over_b[k] = st.butcher_A[s * ns + k] / st.butcher_b[k];
    for (int k = 0; k <= st.nsteps; ++k) tp.bdf[k] = st.bdf_wts[k];
  }

  // side < 0: volume launch over all elements
  void fillParams(KParams& kp, const BasePtrs& bp, const mrh_state& st, double* res, double* jac, int side,
                  const int* side_elems_ptr) const {
    memset(&kp, 0, sizeof(kp));
    kp.nodes = bp.nodes; kp.e2n = bp.e2n; kp.lids = bp.lids; kp.fixed = bp.fixed; kp.rowptr = bp.rowptr;
    kp.pat = bp.pat; kp.pat_id = bp.pat_id; kp.code = bp.code; kp.cpool = bp.cpool;
    kp.u = st.u; kp.u_prev = st.u_prev; kp.u_stage = st.u_stage; kp.res = res; kp.jac = jac;
    kp.compute_jac = jac ? 1 : 0; kp.lump_mass = cfg.lump_mass;
    timeParams(kp.tp, st);
    kp.ph.have_advection = cfg.include_advection; kp.ph.useSUPG = cfg.use_supg; kp.ph.usePSPG = cfg.use_pspg; kp.ph.useLSIC = cfg.use_lsic; kp.ph.useEvisc = cfg.use_evisc;
    kp.ph.formparam = cfg.form_param; kp.ph.le_penalty = cfg.le_penalty;
    for (int i = 0; i < nvar * nb; ++i) kp.offsets[i] = dofs.offsets[i];
    if (side < 0) {
      kp.is_side = 0; kp.elem_list = nullptr; kp.nwork = mesh.nelem;
      for (int i = 0; i < MRH_MAXPROG; ++i) kp.progs[i] = vol_progs[i];
      kp.tab = bp.vol_tab;
    } else {
      const SideLaunch& L = side_launches[side];
      kp.is_side = 1; kp.elem_list = side_elems_ptr; kp.nwork = (int64_t)mesh.sides[side].elems.size();
      for (int i = 0; i < MRH_MAXPROG; ++i) kp.progs[i] = L.progs[i];
      for (int j = 0; j < 4; ++j) kp.ph.bc[j] = L.bc[j];
      kp.tab = bp.side_tab[L.local_side];
    }
  }

  // projection launch: which < 0 the volume (initial data), otherwise index into dirichlet_launches
  void fillProjectionParams(KParams& kp, const BasePtrs& bp, double time, double* rhs, double* mass, int which, const int* side_elems_ptr) const {
    memset(&kp, 0, sizeof(kp));
    kp.nodes = bp.nodes; kp.e2n = bp.e2n; kp.lids = bp.lids; kp.fixed = bp.fixed; kp.rowptr = bp.rowptr;
    kp.pat = bp.pat; kp.pat_id = bp.pat_id; kp.code = bp.code; kp.cpool = bp.cpool;
    kp.res = rhs; kp.jac = mass; kp.compute_jac = mass ? 1 : 0; kp.lump_mass = cfg.lump_mass;
    kp.tp.time = time; kp.tp.dt = 1.0; kp.tp.alpha_u = 1.0;
    for (int i = 0; i < nvar * nb; ++i) kp.offsets[i] = dofs.offsets[i];
    if (which < 0) {
      kp.ph.proj = PROJ_INITIAL; kp.is_side = 0; kp.elem_list = nullptr; kp.nwork = mesh.nelem;
      for (int i = 0; i < MRH_MAXPROG; ++i) kp.progs[i] = init_progs[i];
      kp.tab = bp.vol_tab;
    } else {
      const SideLaunch& L = dirichlet_launches[which];
      kp.ph.proj = PROJ_DIRICHLET; kp.is_side = 1; kp.elem_list = side_elems_ptr; kp.nwork = (int64_t)mesh.sides[L.side].elems.size();
      for (int i = 0; i < MRH_MAXPROG; ++i) kp.progs[i] = L.progs[i];
      for (int j = 0; j < 4; ++j) kp.ph.bc[j] = L.bc[j];
      kp.tab = bp.side_tab[L.local_side];
    }
  }

  // element order of the tile-staged scatter: 8x4x4-element bricks (16x8 blocks in 2-D) on the structured mesh, identity otherwise
  std::vector<int> tileOrder(int64_t* iface_tiles = nullptr) const {
    std::vector<int> perm; perm.reserve((size_t)mesh.nelem);
    const int nx = mesh.NXYZ[0], ny = mesh.NXYZ[1], nz = mesh.NXYZ[2];
    if (dim == 3 && nx > 0 && (int64_t)nx * ny * nz == mesh.nelem) {
      const int bx = 8, by = 4, bz = 4;
      for (int K = 0; K < nz; K += bz) for (int J = 0; J < ny; J += by) for (int I = 0; I < nx; I += bx)
        for (int k = K; k < std::min(K + bz, nz); ++k) for (int j = J; j < std::min(J + by, ny); ++j) for (int i = I; i < std::min(I + bx, nx); ++i)
          perm.push_back((int)(i + (int64_t)nx * (j + (int64_t)ny * k)));
    } else if (dim == 2 && nx > 0 && (int64_t)nx * ny == mesh.nelem) {          // 16x8-element blocks
      const int bx = 16, by = 8;
      for (int J = 0; J < ny; J += by) for (int I = 0; I < nx; I += bx)
        for (int j = J; j < std::min(J + by, ny); ++j) for (int i = I; i < std::min(I + bx, nx); ++i) perm.push_back((int)(i + (int64_t)nx * j));
    } else for (int64_t e = 0; e < mesh.nelem; ++e) perm.push_back((int)e);
    int64_t niface = 0;
    if (part.active()) {
      // tiles (128 consecutive entries) holding an element with a DOF owned elsewhere go first: their rows are complete - and
      // can be sent - while the interior is still being assembled.  A ragged last tile stays last; when it is an
      // interface tile itself no split is possible (niface = -1: the exchange then follows the whole launch).
      const int64_t nt = ((int64_t)perm.size() + 127) / 128;
      const bool ragged = (perm.size() % 128) != 0;
      std::vector<char> flag((size_t)nt, 0);
      for (int64_t k = 0; k < (int64_t)perm.size(); ++k) {
        const int* l = &dofs.lids[(size_t)perm[k] * N];
        for (int i = 0; i < N; ++i) if (part.owner[l[i]] != part.rank) { flag[k / 128] = 1; break; }
      }
      std::vector<int> out; out.reserve(perm.size());
      auto put = [&](int64_t t) { out.insert(out.end(), perm.begin() + t * 128, perm.begin() + std::min<int64_t>((t + 1) * 128, (int64_t)perm.size())); };
      const int64_t nfull = ragged ? nt - 1 : nt;
      for (int64_t t = 0; t < nfull; ++t) if (flag[t]) { put(t); ++niface; }
      for (int64_t t = 0; t < nfull; ++t) if (!flag[t]) put(t);
      if (ragged) { put(nt - 1); if (flag[nt - 1]) niface = -1; }
      perm.swap(out);
    }
    if (iface_tiles) *iface_tiles = niface;
    return perm;
  }

  // true when no coefficient of the volume launch depends on the solution (enables the specialised kernels)
  bool volumeStateIndependent() const { for (auto& p : vol_progs) if (p.kind == 2) return false; return true; }

 private:
  std::vector<std::array<std::string, 3>> pending;

  static bool lookup(const std::map<std::string, std::map<std::string, std::string>>& m, const std::string& var,
                     const std::string& side, std::string& expr) {
    auto it = m.find(var); if (it == m.end()) return false;
    auto a = it->second.find(side); if (a != it->second.end()) { expr = a->second; return true; }
    a = it->second.find("all boundaries"); if (a != it->second.end()) { expr = a->second; return true; }
    return false;
  }
  std::string userOr(const std::string& name, const std::string& dflt) const {
    for (auto& f : user_fns) if (f.first == name) return f.second;
    return dflt;
  }
  void reg(const std::string& fname, const std::string& key, const std::string& dflt, const std::string& loc) {
    ec.addFunction(fname, userOr(key, dflt), loc);
  }
  // PhysicsBase::defineFunctions of the four modules (thermal.cpp:48-65, cdr.cpp:33-57,
  // linearelasticity.cpp:67-86, navierstokes.cpp:64-76)
  void defineModuleFunctions() {
    if (phys == PHYS_THERMAL) {
      reg("thermal source", "thermal source", "0.0", "ip"); reg("thermal diffusion", "thermal diffusion", "1.0", "ip");
      reg("specific heat", "specific heat", "1.0", "ip"); reg("density", "density", "1.0", "ip");
      reg("bx", "advection x", "0.0", "ip"); reg("by", "advection y", "0.0", "ip"); reg("bz", "advection z", "0.0", "ip");
      reg("thermal diffusion", "thermal diffusion", "1.0", "side ip"); reg("robin alpha", "robin alpha", "0.0", "side ip");
    } else if (phys == PHYS_CDR) {
      reg("source", "source", "0.0", "ip"); reg("diffusion", "diffusion", "1.0", "ip"); reg("specific heat", "specific heat", "1.0", "ip");
      reg("density", "density", "1.0", "ip"); reg("reaction", "reaction", "1.0", "ip");
      reg("xvel", "xvel", "1.0", "ip"); reg("yvel", "yvel", "1.0", "ip"); reg("zvel", "zvel", "1.0", "ip");
      reg("SUPG tau", "SUPG tau", "0.0", "ip"); reg("diffusion", "diffusion", "1.0", "side ip"); reg("robin alpha", "robin alpha", "0.0", "side ip");
    } else if (phys == PHYS_LE) {
      reg("lambda", "lambda", "1.0", "ip"); reg("mu", "mu", "0.5", "ip");
      reg("source dx", "source dx", "0.0", "ip"); reg("source dy", "source dy", "0.0", "ip"); reg("source dz", "source dz", "0.0", "ip");
      reg("lambda", "lambda", "1.0", "side ip"); reg("mu", "mu", "0.5", "side ip");
    } else if (phys == PHYS_BURGERS) {         // burgers.cpp:31-45; "C1", "C2", "supg C", "supg C1", "supg C2" have no module defaults (0.0 here when absent and unused)
      reg("Burgers source", "Burgers source", "0.0", "ip"); reg("diffusion", "diffusion", "0.0", "ip");
      reg("xvel", "xvel", "1.0", "ip"); reg("yvel", "yvel", "1.0", "ip"); reg("zvel", "zvel", "1.0", "ip");
      reg("xvel", "xvel", "1.0", "side ip"); reg("yvel", "yvel", "1.0", "side ip"); reg("zvel", "zvel", "1.0", "side ip");
    } else if (phys == PHYS_HELMHOLTZ) {       // helmholtz.cpp:33-66 (functions of the non-fractional branch)
      for (const char* n : {"c2r_x", "c2i_x", "c2r_y", "c2i_y", "c2r_z", "c2i_z", "omega2r", "omega2i", "source_r", "source_i"}) reg(n, n, "0.0", "ip");
      for (const char* n : {"c2r_x", "c2i_x", "c2r_y", "c2i_y", "c2r_z", "c2i_z", "robin_alpha_r", "robin_alpha_i", "source_r_side", "source_i_side"}) reg(n, n, "0.0", "side ip");
    } else if (phys == PHYS_STOKES) {          // stokes.cpp:52-63
      reg("source ux", "source ux", "0.0", "ip"); reg("source pr", "source pr", "0.0", "ip"); reg("source uy", "source uy", "0.0", "ip");
      reg("source uz", "source uz", "0.0", "ip"); reg("viscosity", "viscosity", "1.0", "ip");
    } else {
      reg("source ux", "source ux", "0.0", "ip"); reg("source pr", "source pr", "0.0", "ip"); reg("source uy", "source uy", "0.0", "ip");
      reg("source uz", "source uz", "0.0", "ip"); reg("density", "density", "1.0", "ip"); reg("viscosity", "viscosity", "1.0", "ip");
    }
  }
  std::vector<std::pair<int, std::string>> volNames() const {
    if (phys == PHYS_THERMAL) return {{TH_SOURCE, "thermal source"}, {TH_DIFF, "thermal diffusion"}, {TH_CP, "specific heat"}, {TH_RHO, "density"},
                                      {TH_BX, "bx"}, {TH_BY, "by"}, {TH_BZ, "bz"}};
    if (phys == PHYS_CDR) return {{CDR_SOURCE, "source"}, {CDR_DIFF, "diffusion"}, {CDR_CP, "specific heat"}, {CDR_RHO, "density"},
                                  {CDR_REACTION, "reaction"}, {CDR_XVEL, "xvel"}, {CDR_YVEL, "yvel"}, {CDR_ZVEL, "zvel"}};
    if (phys == PHYS_LE) return {{LE_LAMBDA, "lambda"}, {LE_MU, "mu"}, {LE_SRC0, "source dx"}, {LE_SRC1, "source dy"}, {LE_SRC2, "source dz"}};
    if (phys == PHYS_BURGERS) {
      std::vector<std::pair<int, std::string>> v = {{BG_SRC, "Burgers source"}, {BG_DIFF, "diffusion"}, {BG_VEL, "xvel"}, {BG_VEL + 1, "yvel"}, {BG_VEL + 2, "zvel"}};
      if (dim == 2 && cfg.use_evisc) { v.push_back({BG_C1, "C1"}); v.push_back({BG_C2, "C2"}); }
      if (dim == 2 && cfg.use_supg) { v.push_back({BG_SUPG_C, "supg C"}); v.push_back({BG_SUPG_C1, "supg C1"}); v.push_back({BG_SUPG_C2, "supg C2"}); }
      return v;
    }
    if (phys == PHYS_HELMHOLTZ) return {{HZ_C2 + 0, "c2r_x"}, {HZ_C2 + 1, "c2i_x"}, {HZ_C2 + 2, "c2r_y"}, {HZ_C2 + 3, "c2i_y"}, {HZ_C2 + 4, "c2r_z"}, {HZ_C2 + 5, "c2i_z"},
                                        {HZ_SRC_R, "source_r"}, {HZ_SRC_I, "source_i"}, {HZ_OMEGA2R, "omega2r"}, {HZ_OMEGA2I, "omega2i"}};
    if (phys == PHYS_STOKES) return {{NS_SRC_UX, "source ux"}, {NS_SRC_PR, "source pr"}, {NS_SRC_UY, "source uy"}, {NS_SRC_UZ, "source uz"}, {NS_VISCOSITY, "viscosity"}};
    return {{NS_SRC_UX, "source ux"}, {NS_SRC_PR, "source pr"}, {NS_SRC_UY, "source uy"}, {NS_SRC_UZ, "source uz"},
            {NS_DENSITY, "density"}, {NS_VISCOSITY, "viscosity"}};
  }
  std::vector<std::pair<int, std::string>> sideNames() const {
    if (phys == PHYS_THERMAL) return {{SIDE_DIFF, "thermal diffusion"}};
    if (phys == PHYS_LE) return {{SIDE_LAMBDA, "lambda"}, {SIDE_MU, "mu"}};
    if (phys == PHYS_BURGERS) return {{BG_SIDE_VEL, "xvel"}, {BG_SIDE_VEL + 1, "yvel"}};
    if (phys == PHYS_HELMHOLTZ) return {{HZ_C2 + 0, "c2r_x"}, {HZ_C2 + 1, "c2i_x"}, {HZ_C2 + 2, "c2r_y"}, {HZ_C2 + 3, "c2i_y"}, {HZ_C2 + 4, "c2r_z"}, {HZ_C2 + 5, "c2i_z"},
                                        {HZ_SIDE_ALPHA_R, "robin_alpha_r"}, {HZ_SIDE_ALPHA_I, "robin_alpha_i"}, {HZ_SIDE_SRC_R, "source_r_side"}, {HZ_SIDE_SRC_I, "source_i_side"}};
    return {};
  }
};

}  // namespace mrh
