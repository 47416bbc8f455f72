// mrhyde_b200.hpp - header-only C++ facade over the C ABI (mrhyde_b200.h) that keeps the API surface of the reference's
// hot path, so that code written against MrHyDE's classes reads the same against this engine:
//   PhysicsBase<EvalT>   /root/reference/src/physics/physicsBase.hpp:30-205   (ctor(settings, dimension), defineFunctions,
//                        volumeResidual, boundaryResidual, faceResidual, computeFlux, setWorkset; label, myvars, mybasistypes)
//   thermal / cdr / linearelasticity / navierstokes / stokes / helmholtz   /root/reference/src/physics/{thermal,cdr,linearelasticity,navierstokes,stokes,helmholtz}.hpp
//   FunctionManager      /root/reference/src/managers/functionManager.hpp     (addFunction(name, expr, location), evaluate)
//   Workset              /root/reference/src/tools/workset.hpp:21-584         (time data, varlist, var_bcs, offsets)
//   AssemblyManager      /root/reference/src/managers/assemblyManager.hpp:173-245, 383-494
//                        (createGroups, allocateGroupStorage, createWorkset, finalizeFunctions, createFixedDOFs,
//                         setWorksetButcher, setWorksetBDF, updateStage, updateTimeStep, assembleJacRes, assembleRes, dofConstraints)
// Differences that follow from the fused design: a module's volumeResidual() / boundaryResidual() do not run a
// kernel of their own - they REQUEST their terms on the workset and AssemblyManager::assemble* issues ONE fused launch chain
// per call through mrh_assemble_jacres / mrh_assemble_res (gather, seeding, physics, scatter, dofConstraints).  Teuchos /
// Tpetra / Kokkos types are replaced by plain stand-ins (ParameterList, Vector, CrsMatrix, std::shared_ptr for RCP):
// the reference's third-party stack is absent from this image.  One handle per `set`; one physics module per set.
#ifndef MRHYDE_B200_HPP
#define MRHYDE_B200_HPP

#include <cstdint>
#include <cstdio>
#include <fstream>
#include <istream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "mrhyde_b200.h"

namespace mrh {

using ScalarT = double;
using std::string;
using std::vector;
template <class T> using RCP = std::shared_ptr<T>;

// ---- Teuchos::ParameterList stand-in: name -> string value, with typed access and sublists --------------------------
class ParameterList {
 public:
  template <class T> T get(const string& name, const T& dflt) const {
    auto it = vals_.find(name);
    if (it == vals_.end()) return dflt;
    return convert<T>(it->second);
  }
  template <class T> void set(const string& name, const T& v) { std::ostringstream o; o.precision(17); o << v; vals_[name] = o.str(); }
  void set(const string& name, const char* v) { vals_[name] = v; }
  void set(const string& name, const bool& v) { vals_[name] = v ? "true" : "false"; }
  bool isParameter(const string& name) const { return vals_.count(name) != 0; }
  ParameterList& sublist(const string& name) { return subs_[name]; }
  bool isSublist(const string& name) const { return subs_.count(name) != 0; }
  const std::map<string, string>& entries() const { return vals_; }
  const std::map<string, ParameterList>& sublists() const { return subs_; }

 private:
  template <class T> static T convert(const string& s) { std::istringstream i(s); T v{}; i >> v; return v; }
  std::map<string, string> vals_;
  std::map<string, ParameterList> subs_;
};
template <> inline string ParameterList::convert<string>(const string& s) { return s; }
template <> inline bool ParameterList::convert<bool>(const string& s) { return s == "true" || s == "1"; }

// ---- Tpetra stand-ins: device arrays owned by the caller ------------------------------------------------------------
struct Vector {            // Tpetra::MultiVector local view: (nlocal, ncols) row-major on the device
  double* data = nullptr; int64_t nlocal = 0; int ncols = 1;
};
struct CrsMatrix {         // Tpetra::CrsMatrix local matrix: CRS graph on the host (set-up), values on the device
  const int64_t* rowptr = nullptr; const int32_t* colind = nullptr; int64_t nrows = 0;
  double* values = nullptr;
};
using vector_RCP = RCP<Vector>;
using matrix_RCP = RCP<CrsMatrix>;

inline void check(int rc, mrh_handle h, const char* what) {      // the reference throws (TEUCHOS_TEST_FOR_EXCEPTION)
  if (rc < 0) throw std::runtime_error(string(what) + " failed (" + std::to_string(rc) + "): " + mrh_last_error(h));
}

// ---- FunctionManager: registry of named expressions at "ip" / "side ip" / "point" (functionManager.cpp:49-69) ------
template <class EvalT = ScalarT>
class FunctionManager {
 public:
  struct Entry { string name, expr, location; };
  int addFunction(const string& name, const string& expr, const string& location) {
    if (location != "ip" && location != "side ip" && location != "point") throw std::runtime_error("unknown function location: " + location);
    for (size_t k = 0; k < fns_.size(); ++k) if (fns_[k].name == name && fns_[k].location == location) return (int)k;   // first registration wins
    fns_.push_back({name, expr, location});
    if (h_) check(mrh_add_function(h_, name.c_str(), expr.c_str(), location.c_str()), h_, "mrh_add_function");
    return (int)fns_.size() - 1;
  }
  // values of a state-independent function at given points (the reference returns a Vista over the workset's ips)
  vector<ScalarT> evaluate(const string& name, const string& location, const vector<ScalarT>& xyz, int dim, ScalarT time = 0.0) const {
    if (!h_) throw std::runtime_error("FunctionManager::evaluate before the manager was attached to an engine");
    vector<ScalarT> out(xyz.size() / dim);
    check(mrh_eval_function_host(h_, name.c_str(), location.c_str(), (int64_t)out.size(), xyz.data(), time, out.data()), h_, "mrh_eval_function_host");
    return out;
  }
  const vector<Entry>& functions() const { return fns_; }
  void attach(mrh_handle h) {      // AssemblyManager::finalizeFunctions
    h_ = h;
    for (auto& f : fns_) check(mrh_add_function(h_, f.name.c_str(), f.expr.c_str(), f.location.c_str()), h_, "mrh_add_function");
  }

 private:
  vector<Entry> fns_;
  mrh_handle h_ = nullptr;
};

// ---- Workset: what a module sees of the current assembly pass (workset.hpp:21-584, hot-path members only) ----------
template <class EvalT = ScalarT>
class Workset {
 public:
  vector<string> varlist;                   // variables of the set in module order
  vector<vector<string>> var_bcs;           // [var][side]: "Dirichlet" | "weak Dirichlet" | "Neumann" | "Flux" | "none"
  vector<string> sidenames;
  int dimension = 3;
  ScalarT time = 0.0, deltat = 1.0;
  bool isTransient = false;
  int current_stage = 0;
  vector<ScalarT> butcher_A, butcher_b, butcher_c, BDF_wts;     // setWorksetButcher / setWorksetBDF
  // terms requested by the modules for the pass being assembled (fused into one launch chain by AssemblyManager)
  bool volume_requested = false, boundary_requested = false;
  void resetResidual() { volume_requested = boundary_requested = false; }     // workset.cpp:425-471: nothing to zero here
};

// ---- PhysicsBase and the four hot-path modules ----------------------------------------------------------------------
template <class EvalT = ScalarT>
class PhysicsBase {
 public:
  PhysicsBase() {}
  PhysicsBase(ParameterList& settings, const int& dimension_) : spaceDim(dimension_) { (void)settings; }
  virtual ~PhysicsBase() {}

  // registers the module's named functions ("thermal source", "lambda", ...) with their defaults, user entries of `fs` winning
  // (thermal.cpp:48-65, cdr.cpp:33-57, linearelasticity.cpp:67-86, navierstokes.cpp:64-76).  The engine owns the defaults
  // (engine_host.hpp defineModuleFunctions); the facade forwards the user's list at both locations (physicsInterface.cpp:466-483).
  virtual void defineFunctions(ParameterList& fs, RCP<FunctionManager<EvalT>>& functionManager_) {
    functionManager = functionManager_;
    for (auto& kv : fs.entries()) { functionManager->addFunction(kv.first, kv.second, "ip"); functionManager->addFunction(kv.first, kv.second, "side ip"); }
  }
  virtual void volumeResidual() { need_wkset(); wkset->volume_requested = true; }
  virtual void boundaryResidual() { need_wkset(); wkset->boundary_requested = true; }
  virtual void faceResidual() {}            // no face terms on the hot path (physicsBase.hpp:79)
  virtual void computeFlux() {}
  virtual void setWorkset(RCP<Workset<EvalT>>& wkset_) { wkset = wkset_; }
  virtual vector<string> getDerivedNames() { return {}; }
  // engine-side configuration this module contributes (replaces the module's private members read inside its kernels)
  virtual void configure(mrh_config& cfg) const { (void)cfg; }

  string label;
  vector<string> myvars, mybasistypes;
  RCP<Workset<EvalT>> wkset;
  RCP<FunctionManager<EvalT>> functionManager;
  int spaceDim = 3;

 protected:
  void need_wkset() const { if (!wkset) throw std::runtime_error(label + ": setWorkset has not been called"); }
};

template <class EvalT = ScalarT>
class thermal : public PhysicsBase<EvalT> {          // thermal.cpp:20-41
 public:
  thermal(ParameterList& settings, const int& dimension_) : PhysicsBase<EvalT>(settings, dimension_) {
    this->label = "thermal"; this->myvars = {"e"}; this->mybasistypes = {"HGRAD"};
    formparam = settings.template get<ScalarT>("form_param", 1.0);
    have_advection = settings.template get<bool>("include advection", false);
  }
  void configure(mrh_config& cfg) const override { cfg.form_param = formparam; cfg.include_advection = have_advection; }
  ScalarT formparam; bool have_advection;
};

template <class EvalT = ScalarT>
class cdr : public PhysicsBase<EvalT> {              // cdr.cpp:18-28
 public:
  cdr(ParameterList& settings, const int& dimension_) : PhysicsBase<EvalT>(settings, dimension_) {
    this->label = "cdr"; this->myvars = {"c"}; this->mybasistypes = {"HGRAD"};
  }
};

template <class EvalT = ScalarT>
class linearelasticity : public PhysicsBase<EvalT> { // linearelasticity.cpp:21-61
 public:
  linearelasticity(ParameterList& settings, const int& dimension_) : PhysicsBase<EvalT>(settings, dimension_) {
    this->label = "linearelasticity";
    this->myvars = {"dx", "dy", "dz"}; this->myvars.resize(dimension_);
    this->mybasistypes.assign(dimension_, "HGRAD");
    if (settings.template get<bool>("use crystal elasticity", false) || settings.template get<bool>("Biot", false) ||
        !settings.template get<bool>("use Lame parameters", true))
      throw std::runtime_error("linearelasticity: only the isotropic Lame branch is on the hot path (not supported)");
    formparam = settings.template get<ScalarT>("form_param", 1.0);
    penalty = settings.template get<ScalarT>("penalty", 10.0);
  }
  void configure(mrh_config& cfg) const override { cfg.form_param = formparam; cfg.le_penalty = penalty; }
  ScalarT formparam, penalty;
};

template <class EvalT = ScalarT>
class navierstokes : public PhysicsBase<EvalT> {     // navierstokes.cpp:20-62; variable order ux, pr, uy, uz
 public:
  navierstokes(ParameterList& settings, const int& dimension_) : PhysicsBase<EvalT>(settings, dimension_) {
    this->label = "navierstokes";
    this->myvars = {"ux", "pr", "uy", "uz"}; this->myvars.resize(dimension_ + 1);
    this->mybasistypes.assign(dimension_ + 1, "HGRAD");
    useSUPG = settings.template get<bool>("useSUPG", false);
    usePSPG = settings.template get<bool>("usePSPG", false);
  }
  void configure(mrh_config& cfg) const override { cfg.use_supg = useSUPG; cfg.use_pspg = usePSPG; }
  bool useSUPG, usePSPG;
};

template <class EvalT = ScalarT>
class stokes : public PhysicsBase<EvalT> {           // stokes.cpp:17-46; variable order ux, pr, uy, uz
 public:
  stokes(ParameterList& settings, const int& dimension_) : PhysicsBase<EvalT>(settings, dimension_) {
    this->label = "stokes";
    this->myvars = {"ux", "pr", "uy", "uz"}; this->myvars.resize(dimension_ + 1);
    this->mybasistypes.assign(dimension_ + 1, "HGRAD");
    useLSIC = settings.template get<bool>("useLSIC", false);
    usePSPG = settings.template get<bool>("usePSPG", false);
  }
  void configure(mrh_config& cfg) const override { cfg.use_lsic = useLSIC; cfg.use_pspg = usePSPG; }
  bool useLSIC, usePSPG;
};

template <class EvalT = ScalarT>
class helmholtz : public PhysicsBase<EvalT> {        // helmholtz.cpp:17-28; variables ureal, uimag (non-fractional branch only)
 public:
  helmholtz(ParameterList& settings, const int& dimension_) : PhysicsBase<EvalT>(settings, dimension_) {
    this->label = "helmholtz";
    this->myvars = {"ureal", "uimag"};
    this->mybasistypes.assign(2, "HGRAD");
    if (settings.template get<bool>("fractional", false)) throw std::runtime_error("helmholtz: the 'fractional' branch is not on the hot path");
  }
};

template <class EvalT = ScalarT>
class Burgers : public PhysicsBase<EvalT> {          // burgers.cpp:17-26; variable u
 public:
  Burgers(ParameterList& settings, const int& dimension_) : PhysicsBase<EvalT>(settings, dimension_) {
    this->label = "Burgers";
    this->myvars = {"u"};
    this->mybasistypes.assign(1, "HGRAD");
    use_evisc = settings.template get<bool>("use entropy viscosity", false);
    use_SUPG = settings.template get<bool>("use SUPG", false);
  }
  void configure(mrh_config& cfg) const override { cfg.use_evisc = use_evisc; cfg.use_supg = use_SUPG; }
  bool use_evisc, use_SUPG;
};

// PhysicsImporter::import (physicsImporter.cpp:60-190), hot-path modules only
template <class EvalT = ScalarT>
inline RCP<PhysicsBase<EvalT>> importPhysics(const string& name, ParameterList& settings, const int& dimension) {
  if (name == "thermal") return std::make_shared<thermal<EvalT>>(settings, dimension);
  if (name == "cdr") return std::make_shared<cdr<EvalT>>(settings, dimension);
  if (name == "linearelasticity") return std::make_shared<linearelasticity<EvalT>>(settings, dimension);
  if (name == "navierstokes") return std::make_shared<navierstokes<EvalT>>(settings, dimension);
  if (name == "stokes" || name == "Stokes") return std::make_shared<stokes<EvalT>>(settings, dimension);
  if (name == "helmholtz") return std::make_shared<helmholtz<EvalT>>(settings, dimension);
  if (name == "burgers" || name == "Burgers") return std::make_shared<Burgers<EvalT>>(settings, dimension);
  throw std::runtime_error("physics module not on the hot path: '" + name + "'");
}

// ---- what MeshInterface / DiscretizationInterface / LinearAlgebraInterface hand over for one set -------------------
struct SideSetData { string name; int local_side = 0; vector<int32_t> elems; };
struct BlockData {
  // either a structured brick built by the engine (panzer_stk Square/CubeHex factory conventions) ...
  bool brick = true;
  // ... or the caller's own tables (host pointers, copied at createGroups)
  const double* nodes = nullptr; int64_t nnode = 0;                 // (nnode, dim)
  const int32_t* elem2node = nullptr; int64_t nelem = 0;            // (nelem, 4|8) shards order
  const int32_t* lids = nullptr; const int32_t* offsets = nullptr; int64_t nlocal = 0;     // disc->dof_lids, wkset->offsets
  const int64_t* rowptr = nullptr; const int32_t* colind = nullptr;                         // overlapped CRS graph
  vector<SideSetData> sides;
};

// ---- AssemblyManager: the entry points SolverManager calls (assemblyManager.hpp:173-245) ---------------------------
// `settings` mirrors the input.yaml layout: sublists "Mesh", "Physics" (modules, "Dirichlet conditions", "Neumann conditions",
// "Flux conditions", "use weak Dirichlet", module settings), "Discretization" (sublist "order", "quadrature"), "Functions",
// "Solver" ("use strong DBCs", "fix zero rows", "lump mass").  A per-block level (eblock-0_0_0) may sit in between.
template <class Node = void>
class AssemblyManager {
 public:
  typedef RCP<Vector> vector_RCP;
  typedef RCP<CrsMatrix> matrix_RCP;

  AssemblyManager(RCP<ParameterList>& settings_, const vector<BlockData>& blocks_by_set, int device = 0)
      : settings(settings_), blocks(blocks_by_set), device_(device) {
    ParameterList& solver = settings->sublist("Solver");
    usestrongDBCs = solver.get<bool>("use strong DBCs", true);      // assemblyManager.cpp:34-40
    fix_zero_rows = solver.get<bool>("fix zero rows", false);
    lump_mass = solver.get<bool>("lump mass", false);
    ParameterList& phys = blockList("Physics");
    ParameterList& mesh = settings->sublist("Mesh");
    dimension = mesh.get<int>("dimension", 3);
    // one module per set; "modules: a, b" lists the sets' modules in order (physicsInterface.cpp:492-700 allows several per
    // block - outside the hot path's five configurations, rejected here)
    std::istringstream mods(phys.get<string>("modules", "thermal")); string m;
    while (std::getline(mods, m, ',')) {
      const size_t a = m.find_first_not_of(" \t"), b = m.find_last_not_of(" \t");
      if (a == string::npos) continue;
      modules.push_back(importPhysics<ScalarT>(m.substr(a, b - a + 1), phys, dimension));
    }
    if (modules.size() != blocks.size()) throw std::runtime_error("AssemblyManager: one BlockData per physics set expected (not supported otherwise)");
    handles.assign(modules.size(), nullptr);
    for (size_t set = 0; set < modules.size(); ++set) {
      function_managers.push_back(std::make_shared<FunctionManager<ScalarT>>());
      wkset.push_back(std::make_shared<Workset<ScalarT>>());
    }
  }
  ~AssemblyManager() { for (mrh_handle h : handles) if (h) mrh_destroy(h); }
  AssemblyManager(const AssemblyManager&) = delete;
  AssemblyManager& operator=(const AssemblyManager&) = delete;

  // createGroups (assemblyManager.cpp:272-712): creates the engine of every set and hands over mesh, DOFs, graph, sides.
  // No element groups exist afterwards - one launch covers the block.
  void createGroups() {
    ParameterList& disc = blockList("Discretization");
    ParameterList& mesh = settings->sublist("Mesh");
    for (size_t set = 0; set < modules.size(); ++set) {
      mrh_config cfg; mrh_default_config(&cfg);
      cfg.dim = dimension; cfg.physics = modules[set]->label.c_str();
      cfg.order = disc.sublist("order").get<int>(modules[set]->myvars[0], 1);
      cfg.quadrature = disc.get<int>("quadrature", 2 * cfg.order);                     // discretizationInterface.cpp:166
      cfg.side_quadrature = disc.get<int>("side quadrature", 2 * cfg.order);           // :203
      cfg.use_strong_dbc = usestrongDBCs; cfg.fix_zero_rows = fix_zero_rows; cfg.lump_mass = lump_mass;
      cfg.device = device_;
      modules[set]->configure(cfg);
      if (mrh_create(&cfg, &handles[set]) != MRH_OK) throw std::runtime_error(string("mrh_create failed: ") + mrh_last_error(nullptr));
      mrh_handle h = handles[set];
      const BlockData& B = blocks[set];
      if (B.brick) {
        const int32_t n[3] = {mesh.get<int>("NX", 1), mesh.get<int>("NY", 1), mesh.get<int>("NZ", 1)};
        const double lo[3] = {mesh.get<double>("xmin", 0.0), mesh.get<double>("ymin", 0.0), mesh.get<double>("zmin", 0.0)};
        const double hi[3] = {mesh.get<double>("xmax", 1.0), mesh.get<double>("ymax", 1.0), mesh.get<double>("zmax", 1.0)};
        check(mrh_build_brick(h, n, lo, hi), h, "mrh_build_brick");
      } else {
        check(mrh_set_mesh(h, B.nodes, B.nnode, B.elem2node, B.nelem), h, "mrh_set_mesh");
        check(mrh_set_dofmap(h, B.lids, B.offsets, B.nlocal), h, "mrh_set_dofmap");
        check(mrh_set_crs(h, B.rowptr, B.colind, B.nlocal), h, "mrh_set_crs");
        for (auto& S : B.sides) check(mrh_add_side(h, S.name.c_str(), S.local_side, S.elems.data(), (int64_t)S.elems.size()), h, "mrh_add_side");
      }
    }
  }
  void allocateGroupStorage() {}      // nothing geometric is stored: bases are rebuilt in registers

  // createWorkset (assemblyManager.cpp:760-1100): variables, boundary-condition table and functions of every set
  void createWorkset() {
    ParameterList& phys = blockList("Physics");
    const bool weak = phys.sublist("Dirichlet conditions").get<bool>("use weak Dirichlet", phys.get<bool>("use weak Dirichlet", false));
    for (size_t set = 0; set < modules.size(); ++set) {
      mrh_handle h = need(set);
      wkset[set]->varlist = modules[set]->myvars; wkset[set]->dimension = dimension;
      modules[set]->setWorkset(wkset[set]);
      // "Parameters" block (src/managers/parameterManager.cpp:85-118): type scalar -> `value`, type vector -> the numbers of the
      // file named by `source`; leaves of the coefficient expressions, constants of the assembly pass
      if (settings->isSublist("Parameters"))
        for (auto& pr : settings->sublist("Parameters").sublists()) {
          const string type = pr.second.template get<string>("type", "scalar");
          vector<double> vals;
          if (type == "scalar") vals.push_back(pr.second.template get<double>("value", 0.0));
          else if (type == "vector") {
            std::ifstream fin(pr.second.template get<string>("source", ""));
            double x; while (fin >> x) vals.push_back(x);
            if (vals.empty()) throw std::runtime_error("parameter '" + pr.first + "': no values in its source file");
          } else continue;                                         // (discretized parameters have no counterpart here)
          check(mrh_set_parameter(h, pr.first.c_str(), vals.data(), (int32_t)vals.size()), h, "mrh_set_parameter");
        }
      modules[set]->defineFunctions(settings->sublist("Functions"), function_managers[set]);
      static const struct { const char* list; int code; } kinds[3] = {{"Dirichlet conditions", MRH_BC_DIRICHLET}, {"Neumann conditions", MRH_BC_NEUMANN}, {"Flux conditions", MRH_BC_FLUX}};
      for (auto& k : kinds) {
        if (!phys.isSublist(k.list)) continue;
        for (auto& var : phys.sublist(k.list).sublists())
          for (auto& se : var.second.entries()) {
            const int code = (k.code == MRH_BC_DIRICHLET && weak) ? MRH_BC_WEAK_DIRICHLET : k.code;
            bool mine = false; for (auto& v : modules[set]->myvars) mine = mine || v == var.first;
            if (mine) check(mrh_set_bc(h, var.first.c_str(), se.first.c_str(), code, se.second.c_str()), h, "mrh_set_bc");
          }
      }
    }
  }
  // finalizeFunctions (assemblyManager.cpp:1110-1130) + createFixedDOFs (:185-265): compiles the expressions, builds the
  // fixed-DOF table, the slot-pattern dictionary and the tile plan, JIT-specialises the kernels
  void finalizeFunctions() {
    for (size_t set = 0; set < modules.size(); ++set) {
      mrh_handle h = need(set);
      function_managers[set]->attach(h);
      check(mrh_finalize(h), h, "mrh_finalize");
      int64_t sz[16]; check(mrh_get_sizes(h, sz), h, "mrh_get_sizes");
      const int nvar = (int)sz[4], nsides = (int)sz[11];
      static const char* names[5] = {"none", "Dirichlet", "weak Dirichlet", "Neumann", "Flux"};
      wkset[set]->var_bcs.assign(nvar, vector<string>(nsides));
      wkset[set]->sidenames.resize(nsides);
      for (int s = 0; s < nsides; ++s) {
        char buf[64]; mrh_get_side(h, s, buf, 64, nullptr, nullptr); wkset[set]->sidenames[s] = buf;
        for (int v = 0; v < nvar; ++v) wkset[set]->var_bcs[v][s] = names[mrh_get_bc(h, v, s)];
      }
    }
  }
  void createFixedDOFs() {}           // done by finalizeFunctions (needs the compiled boundary-condition table)

  // time-integration data of the workset (assemblyManager.cpp:3195-3436)
  void setWorksetButcher(const size_t& set, const size_t& /*block*/, const vector<ScalarT>& butcher_A, const vector<ScalarT>& butcher_b,
                         const vector<ScalarT>& butcher_c) {
    wkset[set]->butcher_A = butcher_A; wkset[set]->butcher_b = butcher_b; wkset[set]->butcher_c = butcher_c;
  }
  void setWorksetBDF(const size_t& set, const size_t& /*block*/, const vector<ScalarT>& BDF_wts) { wkset[set]->BDF_wts = BDF_wts; }
  void updateStage(const int& stage, const ScalarT& current_time, const ScalarT& deltat) {
    for (auto& w : wkset) { w->current_stage = stage; w->time = current_time; w->deltat = deltat; }
  }
  void updateTimeStep() {}

  // assembleJacRes (assemblyManager.cpp:2045-2140, 2150-2665): res += -r(u), J += dr/du, fixed rows skipped and given a unit
  // diagonal (dofConstraints / updateJacDBC are part of the fused call).  sol / sol_stage / sol_prev are indexed by set;
  // phi*, Psol, sensitivities and adjoints have no counterpart (compute_sens / compute_disc_sens / useadjoint must be false).
  void assembleJacRes(const size_t& set, const size_t& stage, vector<vector_RCP>& sol, vector<vector_RCP>& sol_stage,
                      vector<vector_RCP>& sol_prev, vector<vector_RCP>& /*phi*/, vector<vector_RCP>& /*phi_stage*/,
                      vector<vector_RCP>& /*phi_prev*/, const bool& compute_jacobian, const bool& compute_sens,
                      const bool& compute_disc_sens, vector_RCP& res, matrix_RCP& J, const bool& isTransient,
                      const ScalarT& current_time, const bool& useadjoint, const bool& /*store_adjPrev*/,
                      const int& num_active_params, vector_RCP& /*Psol*/, const bool& /*is_final_time*/, const ScalarT& deltat,
                      void* stream = nullptr) {
    if (compute_sens || compute_disc_sens || useadjoint || num_active_params > 0)
      throw std::runtime_error("assembleJacRes: sensitivities / adjoints are not on the hot path (not supported)");
    mrh_handle h = need(set);
    Workset<ScalarT>& w = *wkset[set];
    w.resetResidual();
    w.isTransient = isTransient; w.deltat = deltat; w.current_stage = (int)stage;
    // updateWorksetTime (assemblyManager.cpp:2717-2732): the workset time of a transient pass is the stage time
    w.time = (isTransient && stage < w.butcher_c.size()) ? current_time + w.butcher_c[stage] * deltat : current_time;
    // the reference's element loop calls the module through PhysicsInterface::volumeResidual / boundaryResidual
    // (physicsInterface.cpp:1231-1365); here the module requests its terms and ONE fused call assembles them
    modules[set]->volumeResidual();
    modules[set]->boundaryResidual();
    mrh_state st = state_of(w, set, sol, sol_stage, sol_prev, stream);
    const int rc = compute_jacobian ? mrh_assemble_jacres(h, &st, res->data, J->values) : mrh_assemble_res(h, &st, res->data);
    check(rc, h, compute_jacobian ? "mrh_assemble_jacres" : "mrh_assemble_res");
  }
  // assembleRes (assemblyManager.cpp:2916-3150): residual only
  void assembleRes(const size_t& set, const size_t& stage, vector<vector_RCP>& sol, vector<vector_RCP>& sol_stage,
                   vector<vector_RCP>& sol_prev, vector<vector_RCP>& phi, vector<vector_RCP>& phi_stage, vector<vector_RCP>& phi_prev,
                   vector_RCP& Psol, vector_RCP& res, matrix_RCP& J, const bool& isTransient, const ScalarT& current_time,
                   const ScalarT& deltat, void* stream = nullptr) {
    assembleJacRes(set, stage, sol, sol_stage, sol_prev, phi, phi_stage, phi_prev, false, false, false, res, J, isTransient,
                   current_time, false, false, 0, Psol, false, deltat, stream);
  }
  // dofConstraints (assemblyManager.cpp:3158-3188): already applied inside assembleJacRes; kept for call-site compatibility
  void dofConstraints(const size_t& /*set*/, matrix_RCP& /*J*/, vector_RCP& /*res*/, const ScalarT& /*current_time*/,
                      const bool& /*compute_jacobian*/, const bool& /*compute_disc_sens*/) {}

  // ---- what the caller needs to size its vectors and matrix (disc->dof_lids, linalg graph) ---------------------------
  struct Sizes { int64_t dim, nelem, nnode, nvert, nvar, nb, ndof_elem, nlocal, nnz, nq, nqs, nsides, npatterns, nfixed; };
  Sizes sizes(const size_t& set) const {
    int64_t v[16]; check(mrh_get_sizes(need(set), v), handles[set], "mrh_get_sizes");
    return Sizes{v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8], v[9], v[10], v[11], v[12], v[13]};
  }
  void getGraph(const size_t& set, vector<int64_t>& rowptr, vector<int32_t>& colind) const {
    const Sizes s = sizes(set); rowptr.resize(s.nlocal + 1); colind.resize(s.nnz);
    check(mrh_get_crs(need(set), rowptr.data(), colind.data()), handles[set], "mrh_get_crs");
  }
  void getDOFs(const size_t& set, vector<int32_t>& lids, vector<int32_t>& offsets, vector<uint8_t>& isFixedDOF, vector<double>& dofxyz,
               vector<int32_t>& dofvar) const {
    const Sizes s = sizes(set);
    lids.resize(s.nelem * s.ndof_elem); offsets.resize(s.nvar * s.nb); isFixedDOF.resize(s.nlocal); dofxyz.resize(s.nlocal * s.dim); dofvar.resize(s.nlocal);
    check(mrh_get_dofmap(need(set), lids.data(), offsets.data(), isFixedDOF.data(), blocks[set].brick ? dofxyz.data() : nullptr,
                         blocks[set].brick ? dofvar.data() : nullptr), handles[set], "mrh_get_dofmap");
  }
  const char* kernelPath(const size_t& set) const { return mrh_last_kernel_path(need(set)); }
  mrh_handle handle(const size_t& set) const { return need(set); }

  RCP<ParameterList> settings;
  vector<RCP<PhysicsBase<ScalarT>>> modules;                 // one per set
  vector<RCP<FunctionManager<ScalarT>>> function_managers;   // one per set
  vector<RCP<Workset<ScalarT>>> wkset;                       // one per set
  bool usestrongDBCs = true, fix_zero_rows = false, lump_mass = false;
  int dimension = 3;

 private:
  mrh_handle need(const size_t& set) const {
    if (set >= handles.size() || !handles[set]) throw std::runtime_error("AssemblyManager: createGroups has not been called for this set");
    return handles[set];
  }
  // a per-block level ("eblock-0_0_0") may sit between the top-level list and its entries (input.yaml)
  ParameterList& blockList(const string& name) {
    ParameterList& top = settings->sublist(name);
    for (auto& kv : top.sublists()) if (kv.first.compare(0, 6, "eblock") == 0) return top.sublist(kv.first);
    return top;
  }
  mrh_state state_of(const Workset<ScalarT>& w, const size_t& set, vector<vector_RCP>& sol, vector<vector_RCP>& sol_stage,
                     vector<vector_RCP>& sol_prev, void* stream) const {
    mrh_state st{};
    st.u = sol.at(set)->data;
    st.is_transient = w.isTransient ? 1 : 0; st.time = w.time; st.deltat = w.deltat; st.stream = stream;
    if (w.isTransient) {
      if (set >= sol_prev.size() || !sol_prev[set]) throw std::runtime_error("transient assembly needs sol_prev");
      st.u_prev = sol_prev[set]->data; st.nsteps = sol_prev[set]->ncols;
      if (set < sol_stage.size() && sol_stage[set]) st.u_stage = sol_stage[set]->data;
      st.nstages = (int)w.butcher_b.size(); st.stage = w.current_stage;
      st.butcher_A = w.butcher_A.data(); st.butcher_b = w.butcher_b.data(); st.bdf_wts = w.BDF_wts.data();
      if ((int)w.BDF_wts.size() != st.nsteps + 1 || (int)w.butcher_A.size() != st.nstages * st.nstages)
        throw std::runtime_error("setWorksetButcher / setWorksetBDF sizes do not match sol_prev / the stage count");
    }
    return st;
  }
  vector<BlockData> blocks;
  vector<mrh_handle> handles;
  int device_ = 0;
};

// ---- on-disk formats ---------------------------------------------------------------------------------------
// readInput: the subset of YAML that MrHyDE's input decks use (UserInterface, /root/reference/src/interfaces/userInterface.hpp:30-270,
// reads them with Teuchos' YAML reader): nested block mappings by indentation, scalar values (plain, 'single' or "double" quoted),
// '#' comments, the '%YAML' / '---' / '...' markers and the 'ANONYMOUS:' root.  Sequences, anchors and flow collections are not
// part of those decks and are rejected.
inline RCP<ParameterList> readInput(std::istream& in) {
  auto root = std::make_shared<ParameterList>();
  struct Level { int indent; ParameterList* list; };
  vector<Level> stack{{-1, root.get()}};
  string line; int lineno = 0;
  auto fail = [&](const string& why) { throw std::runtime_error("input line " + std::to_string(lineno) + ": " + why); };
  while (std::getline(in, line)) {
    ++lineno;
    if (!line.empty() && line.back() == '\r') line.pop_back();
    // strip comments outside quotes
    { bool q1 = false, q2 = false; for (size_t k = 0; k < line.size(); ++k) { if (line[k] == '\'' && !q2) q1 = !q1; else if (line[k] == '"' && !q1) q2 = !q2;
        else if (line[k] == '#' && !q1 && !q2 && (k == 0 || line[k - 1] == ' ' || line[k - 1] == '\t')) { line.erase(k); break; } } }
    const size_t first = line.find_first_not_of(" \t");
    if (first == string::npos) continue;
    if (line.compare(first, 5, "%YAML") == 0 || line.compare(first, 3, "---") == 0 || line.compare(first, 3, "...") == 0) continue;
    if (line.find('\t') != string::npos && line.find('\t') < first) fail("tabs in the indentation");
    if (line[first] == '-' || line[first] == '[' || line[first] == '{' || line[first] == '&' || line[first] == '*') fail("YAML sequences / flow collections / anchors are not supported");
    // key: up to the first ':' that is followed by a space or ends the line (keys of the decks contain spaces, commas, brackets)
    size_t colon = string::npos;
    { bool q1 = false, q2 = false; for (size_t k = first; k < line.size(); ++k) { if (line[k] == '\'' && !q2) q1 = !q1; else if (line[k] == '"' && !q1) q2 = !q2;
        else if (line[k] == ':' && !q1 && !q2 && (k + 1 == line.size() || line[k + 1] == ' ')) { colon = k; break; } } }
    if (colon == string::npos) fail("expected 'key: value'");
    auto trim = [](string t) { const size_t a = t.find_first_not_of(" \t"), b = t.find_last_not_of(" \t"); return a == string::npos ? string() : t.substr(a, b - a + 1); };
    auto unquote = [](string t) { if (t.size() >= 2 && ((t.front() == '\'' && t.back() == '\'') || (t.front() == '"' && t.back() == '"'))) return t.substr(1, t.size() - 2); return t; };
    const string key = unquote(trim(line.substr(first, colon - first)));
    const string val = trim(line.substr(colon + 1));
    const int indent = (int)first;
    while (stack.size() > 1 && indent <= stack.back().indent) stack.pop_back();
    ParameterList* cur = stack.back().list;
    if (val.empty()) {
      if (key == "ANONYMOUS" && stack.size() == 1) { stack.push_back({indent, root.get()}); continue; }      // the decks' root mapping
      stack.push_back({indent, &cur->sublist(key)});
    } else cur->set(key, unquote(val).c_str());
  }
  return root;
}
inline RCP<ParameterList> readInput(const string& path) {
  std::ifstream f(path);
  if (!f) throw std::runtime_error("cannot open input file " + path);
  return readInput(f);
}

// Matrix Market (coordinate real general, array real general) dumps of a CRS matrix / a vector on the HOST: what
// LinearAlgebraInterface::writeToFile produces for J, res, du (/root/reference/src/managers/solverManager.cpp:1653), so that an
// assembly of this engine can be compared offline with a real MrHyDE build.  1-based indices, 17 significant digits.
inline void writeMatrixMarket(const string& path, const int64_t* rowptr, const int32_t* colind, const double* values, int64_t nrows, int64_t ncols) {
  FILE* f = fopen(path.c_str(), "w");
  if (!f) throw std::runtime_error("cannot write " + path);
  fprintf(f, "%%%%MatrixMarket matrix coordinate real general\n%lld %lld %lld\n", (long long)nrows, (long long)ncols, (long long)rowptr[nrows]);
  for (int64_t r = 0; r < nrows; ++r)
    for (int64_t p = rowptr[r]; p < rowptr[r + 1]; ++p) fprintf(f, "%lld %d %.17g\n", (long long)r + 1, colind[p] + 1, values[p]);
  fclose(f);
}
inline void writeMatrixMarket(const string& path, const double* vec, int64_t n) {
  FILE* f = fopen(path.c_str(), "w");
  if (!f) throw std::runtime_error("cannot write " + path);
  fprintf(f, "%%%%MatrixMarket matrix array real general\n%lld 1\n", (long long)n);
  for (int64_t i = 0; i < n; ++i) fprintf(f, "%.17g\n", vec[i]);
  fclose(f);
}

}  // namespace mrh
#endif  // MRHYDE_B200_HPP
