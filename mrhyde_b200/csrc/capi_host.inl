// C-ABI entry points that touch only host state (include/mrhyde_b200.h).  Included by engine.cu (product)
// and by tests/emu/emu.cpp (CPU stepping harness for the kernels); the backend supplies
//   backend_create / backend_destroy  and the finalize / assemble entry points.
#include <cstdio>
#include <cstring>
#include "engine_host.hpp"

struct mrh_engine {
  mrh::EngineHost host;
  std::string err;
  void* impl = nullptr;
  int64_t npatterns = 0;
};

static thread_local std::string g_create_err;
static int backend_create(mrh_engine* h);
static void backend_destroy(mrh_engine* h);
static void backend_nodes_changed(mrh_engine* h);      // vertex coordinates replaced after finalize

#define MRH_TRY(h) try {
#define MRH_CATCH(h)                                                                                  \
  } catch (const std::bad_alloc&) { (h)->err = "out of host memory"; return MRH_ENOMEM; }             \
  catch (const std::exception& e) {                                                                   \
    (h)->err = e.what();                                                                              \
    if ((h)->err.find("not able to decompose") != std::string::npos || (h)->err.find("could not evaluate") != std::string::npos) return MRH_EPARSE; \
    if ((h)->err.find("not supported") != std::string::npos) return MRH_EUNSUPPORTED;                 \
    return MRH_EINVAL; }

extern "C" {

void mrh_default_config(mrh_config* c) {
  memset(c, 0, sizeof(*c));
  c->dim = 3; c->physics = "thermal"; c->order = 1; c->quadrature = -1; c->side_quadrature = -1;
  c->use_strong_dbc = 1; c->form_param = 1.0; c->le_penalty = 10.0;
}

int mrh_create(const mrh_config* cfg, mrh_handle* out) {
  if (!cfg || !out) { g_create_err = "null argument"; return MRH_EINVAL; }
  mrh_engine* h = nullptr;
  try {
    h = new mrh_engine();
    h->host.configure(*cfg);
    const int rc = backend_create(h);
    if (rc != MRH_OK) { g_create_err = h->err; delete h; return rc; }
  } catch (const std::exception& e) { g_create_err = e.what(); delete h; return MRH_EINVAL; }
  *out = h; return MRH_OK;
}

int mrh_destroy(mrh_handle h) { if (!h) return MRH_EINVAL; backend_destroy(h); delete h; return MRH_OK; }

const char* mrh_last_error(mrh_handle h) { return h ? h->err.c_str() : g_create_err.c_str(); }

int mrh_build_brick(mrh_handle h, const int32_t* nxyz, const double* lo, const double* hi) {
  if (!h || !nxyz || !lo || !hi) return MRH_EINVAL;
  MRH_TRY(h) h->host.buildBrick(nxyz, lo, hi); return MRH_OK; MRH_CATCH(h)
}

int mrh_build_brick_part(mrh_handle h, const int32_t* nxyz, const double* lo, const double* hi, const int32_t* procs, int32_t rank) {
  if (!h || !nxyz || !lo || !hi || !procs) return MRH_EINVAL;
  MRH_TRY(h) h->host.buildBrickPart(nxyz, lo, hi, procs, rank); return MRH_OK; MRH_CATCH(h)
}
int mrh_set_partition(mrh_handle h, const int64_t* gids, const int32_t* owner, int32_t rank, int32_t world) {
  if (!h || !gids || !owner || world < 1 || rank < 0 || rank >= world) return MRH_EINVAL;
  MRH_TRY(h) h->host.setPartition(gids, owner, rank, world); return MRH_OK; MRH_CATCH(h)
}
int mrh_get_partition(mrh_handle h, int64_t* gids, int32_t* owner) {
  if (!h || h->host.part.gids.empty()) return MRH_EINVAL;
  if (gids) memcpy(gids, h->host.part.gids.data(), h->host.part.gids.size() * sizeof(int64_t));
  if (owner) memcpy(owner, h->host.part.owner.data(), h->host.part.owner.size() * sizeof(int));
  return MRH_OK;
}
// exchange-plan metadata for a transport of the caller's choice (tests: gloo; mrh_comm_init moves it over NCCL itself)
int64_t mrh_exchange_meta_size(mrh_handle h, int32_t peer) {
  if (!h || !h->host.part.active()) return 0;
  return (int64_t)h->host.plan.meta_for(peer, h->host.part, h->host.crs).size();
}
int mrh_exchange_meta_get(mrh_handle h, int32_t peer, int64_t* out) {
  if (!h || !out || !h->host.part.active()) return MRH_EINVAL;
  MRH_TRY(h)
  const auto m = h->host.plan.meta_for(peer, h->host.part, h->host.crs);
  memcpy(out, m.data(), m.size() * sizeof(int64_t));
  return MRH_OK;
  MRH_CATCH(h)
}
int mrh_exchange_meta_put(mrh_handle h, int32_t peer, const int64_t* in, int64_t n) {
  if (!h || (n > 0 && !in) || !h->host.part.active()) return MRH_EINVAL;
  MRH_TRY(h) h->host.plan.ingest(peer, in, n, h->host.part, h->host.crs); return MRH_OK; MRH_CATCH(h)
}
// which: 0 send rows, 1 send slots, 2 receive rows, 3 receive slots (-(k+1): ghost entry k) of `peer`; 4 / 5: ghost rows / ghost
// column gids (peer ignored).  Returns the entry count; out may be NULL to query it.
int64_t mrh_get_exchange_table(mrh_handle h, int32_t peer, int32_t which, int64_t* out) {
  if (!h) return MRH_EINVAL;
  const auto& P = h->host.plan;
  std::vector<int64_t> v;
  if (which == 4) v.assign(P.ghost_row.begin(), P.ghost_row.end());
  else if (which == 5) v = P.ghost_colgid;
  else {
    const mrh::ExchangePlan::Peer* p = P.find(peer);
    if (!p) return 0;
    if (which == 0) v.assign(p->send_rows.begin(), p->send_rows.end());
    else if (which == 1) v = p->send_slots;
    else if (which == 2) v.assign(p->recv_rows.begin(), p->recv_rows.end());
    else if (which == 3) v = p->recv_slots;
    else return MRH_EINVAL;
  }
  if (out && !v.empty()) memcpy(out, v.data(), v.size() * sizeof(int64_t));
  return (int64_t)v.size();
}

int mrh_set_mesh(mrh_handle h, const double* nodes, int64_t nnode, const int32_t* elem2node, int64_t nelem) {
  if (!h || !nodes || !elem2node || nnode <= 0 || nelem <= 0) return MRH_EINVAL;
  MRH_TRY(h)
  auto& M = h->host.mesh;
  M.dim = h->host.dim; M.nvert = h->host.nv; M.nnode = nnode; M.nelem = nelem;
  for (int64_t k = 0; k < nelem * M.nvert; ++k) if (elem2node[k] < 0 || elem2node[k] >= nnode) throw std::runtime_error("element vertex id out of range");
  M.nodes.assign(nodes, nodes + nnode * M.dim);
  M.elem2node.assign(elem2node, elem2node + nelem * M.nvert);
  for (int d = 0; d < 3; ++d) M.NXYZ[d] = 0;
  M.sides.clear(); h->host.have_mesh = true; h->host.host_final = false;
  return MRH_OK;
  MRH_CATCH(h)
}

int mrh_set_nodes(mrh_handle h, const double* nodes, int64_t nnode) {
  if (!h || !nodes || !h->host.have_mesh || nnode != h->host.mesh.nnode) return MRH_EINVAL;
  MRH_TRY(h)
  auto& M = h->host.mesh;
  M.nodes.assign(nodes, nodes + nnode * M.dim);
  backend_nodes_changed(h);
  return MRH_OK;
  MRH_CATCH(h)
}

int mrh_set_dofmap(mrh_handle h, const int32_t* lids, const int32_t* offsets, int64_t nlocal) {
  if (!h || !lids || !offsets || nlocal <= 0 || !h->host.have_mesh) return MRH_EINVAL;
  MRH_TRY(h)
  auto& D = h->host.dofs; auto& H = h->host;
  D.nvar = H.nvar; D.nb = H.nb; D.N = H.N; D.nlocal = nlocal;
  D.lids.assign(lids, lids + H.mesh.nelem * H.N);
  D.offsets.assign(offsets, offsets + H.nvar * H.nb);
  for (int v : D.lids) if (v < 0 || v >= nlocal) throw std::runtime_error("LID out of range");
  for (int v : D.offsets) if (v < 0 || v >= H.N) throw std::runtime_error("offset out of range");
  D.dofxyz.clear(); D.dofvar.clear(); H.have_dofs = true; H.host_final = false;
  return MRH_OK;
  MRH_CATCH(h)
}

int mrh_set_crs(mrh_handle h, const int64_t* rowptr, const int32_t* colind, int64_t nrows) {
  if (!h || !rowptr || !colind || !h->host.have_dofs || nrows != h->host.dofs.nlocal) return MRH_EINVAL;
  MRH_TRY(h)
  auto& G = h->host.crs; G.nrows = nrows;
  G.rowptr.assign(rowptr, rowptr + nrows + 1);
  G.colind.assign(colind, colind + rowptr[nrows]);
  for (int64_t r = 0; r < nrows; ++r) {
    if (rowptr[r + 1] - rowptr[r] > 65535) throw std::runtime_error("CRS row longer than 65535 entries");
    for (int64_t p = rowptr[r] + 1; p < rowptr[r + 1]; ++p) if (colind[p] <= colind[p - 1]) throw std::runtime_error("CRS columns must be ascending within a row");
  }
  h->host.have_crs = true; h->host.host_final = false;
  return MRH_OK;
  MRH_CATCH(h)
}

int mrh_add_side(mrh_handle h, const char* name, int32_t local_side, const int32_t* elems, int64_t n) {
  if (!h || !name || (n > 0 && !elems) || !h->host.have_mesh) return MRH_EINVAL;
  MRH_TRY(h)
  if (local_side < 0 || local_side >= h->host.ref.nsides) throw std::runtime_error("local side id out of range");
  for (int64_t k = 0; k < n; ++k) if (elems[k] < 0 || elems[k] >= h->host.mesh.nelem) throw std::runtime_error("side-set element id out of range");
  mrh::SideSet S; S.name = name; S.local_side = local_side; S.elems.assign(elems, elems + n);
  h->host.mesh.sides.push_back(S); h->host.host_final = false;
  return MRH_OK;
  MRH_CATCH(h)
}

int mrh_set_bc(mrh_handle h, const char* var, const char* side, int32_t bc_code, const char* expr) {
  if (!h || !var || !side || !expr) return MRH_EINVAL;
  MRH_TRY(h) h->host.setBC(var, side, bc_code, expr); return MRH_OK; MRH_CATCH(h)
}

int mrh_add_function(mrh_handle h, const char* name, const char* expr, const char* location) {
  if (!h || !name || !expr || !location) return MRH_EINVAL;
  MRH_TRY(h)
  const std::string loc(location);
  if (loc != "ip" && loc != "side ip" && loc != "point") throw std::runtime_error("unknown function location: " + loc);
  h->host.addUserFunction(name, expr, loc); h->host.host_final = false;
  return MRH_OK;
  MRH_CATCH(h)
}

int mrh_set_parameter(mrh_handle h, const char* name, const double* values, int32_t n) {
  if (!h || !name || !values || n < 1 || n > 100) return MRH_EINVAL;
  MRH_TRY(h) h->host.setParameter(name, values, n); h->host.host_final = false; return MRH_OK; MRH_CATCH(h)
}

int mrh_finalize_host(mrh_handle h) {
  if (!h) return MRH_EINVAL;
  MRH_TRY(h) h->host.finalizeHost(); return MRH_OK; MRH_CATCH(h)
}

int mrh_get_sizes(mrh_handle h, int64_t* out) {
  if (!h || !out) return MRH_EINVAL;
  auto& H = h->host;
  int64_t nfixed = 0; for (auto f : H.isFixed) nfixed += f;
  int64_t v[16] = {H.dim, H.mesh.nelem, H.mesh.nnode, H.nv, H.nvar, H.nb, H.N, H.dofs.nlocal,
                   (int64_t)H.crs.colind.size(), H.ref.vol.nq, H.ref.side.empty() ? 0 : H.ref.side[0].nq,
                   (int64_t)H.mesh.sides.size(), h->npatterns, nfixed, 0, 0};
  memcpy(out, v, sizeof(v));
  return MRH_OK;
}
int mrh_get_mesh(mrh_handle h, double* nodes, int32_t* elem2node) {
  if (!h) return MRH_EINVAL;
  if (nodes) memcpy(nodes, h->host.mesh.nodes.data(), h->host.mesh.nodes.size() * sizeof(double));
  if (elem2node) memcpy(elem2node, h->host.mesh.elem2node.data(), h->host.mesh.elem2node.size() * sizeof(int));
  return MRH_OK;
}
int mrh_get_dofmap(mrh_handle h, int32_t* lids, int32_t* offsets, uint8_t* is_fixed, double* dofxyz, int32_t* dofvar) {
  if (!h) return MRH_EINVAL;
  auto& D = h->host.dofs;
  if (lids) memcpy(lids, D.lids.data(), D.lids.size() * sizeof(int));
  if (offsets) memcpy(offsets, D.offsets.data(), D.offsets.size() * sizeof(int));
  if (is_fixed) { if (h->host.isFixed.size() != (size_t)D.nlocal) return MRH_EINVAL; memcpy(is_fixed, h->host.isFixed.data(), h->host.isFixed.size()); }
  if (dofxyz) { if (D.dofxyz.empty()) return MRH_EINVAL; memcpy(dofxyz, D.dofxyz.data(), D.dofxyz.size() * sizeof(double)); }
  if (dofvar) { if (D.dofvar.empty()) return MRH_EINVAL; memcpy(dofvar, D.dofvar.data(), D.dofvar.size() * sizeof(int)); }
  return MRH_OK;
}
int mrh_get_crs(mrh_handle h, int64_t* rowptr, int32_t* colind) {
  if (!h) return MRH_EINVAL;
  if (rowptr) memcpy(rowptr, h->host.crs.rowptr.data(), h->host.crs.rowptr.size() * sizeof(int64_t));
  if (colind) memcpy(colind, h->host.crs.colind.data(), h->host.crs.colind.size() * sizeof(int));
  return MRH_OK;
}
int mrh_get_side(mrh_handle h, int32_t s, char* name, int32_t namelen, int32_t* local_side, int32_t* elems) {
  if (!h || s < 0 || s >= (int)h->host.mesh.sides.size()) return MRH_EINVAL;
  const auto& S = h->host.mesh.sides[s];
  if (name && namelen > 0) { strncpy(name, S.name.c_str(), namelen - 1); name[namelen - 1] = 0; }
  if (local_side) *local_side = S.local_side;
  if (elems) memcpy(elems, S.elems.data(), S.elems.size() * sizeof(int));
  return (int)S.elems.size();
}
int mrh_get_bc(mrh_handle h, int32_t var, int32_t side) {
  if (!h || var < 0 || var >= (int)h->host.var_bcs.size() || side < 0 || side >= (int)h->host.var_bcs[var].size()) return MRH_EINVAL;
  return h->host.var_bcs[var][side];
}
const char* mrh_get_varname(mrh_handle h, int32_t var) {
  if (!h || var < 0 || var >= h->host.nvar) return "";
  return h->host.varlist[var].c_str();
}

int mrh_eval_function_host(mrh_handle h, const char* name, const char* location, int64_t npts, const double* xyz,
                           double time, double* out) {
  if (!h || !name || !location || !xyz || !out) return MRH_EINVAL;
  MRH_TRY(h)
  auto& H = h->host;
  if (!H.host_final) H.finalizeHost();
  mrh::ExprCompiler ec = H.ec;                      // private copy: compile() appends code
  const mrh::Prog pr = ec.compile(name, location);
  if (pr.kind == 2) throw std::runtime_error("function depends on the solution; host evaluation not supported");
  for (int64_t i = 0; i < npts; ++i) {
    mrh::PointState<double> ps{};
    for (int c = 0; c < H.dim; ++c) ps.x[c] = xyz[i * H.dim + c];
    ps.time = time; ps.fields = nullptr;
    out[i] = mrh::eval_prog<double>(pr, ec.code.data(), ec.cpool.data(), ps);
  }
  return MRH_OK;
  MRH_CATCH(h)
}

}  // extern "C"
