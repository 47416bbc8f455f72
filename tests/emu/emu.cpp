// TEST INFRASTRUCTURE (not product code): steps the product's __host__ __device__ kernel phases
// (mrhyde_b200/csrc/assemble.h) on the CPU so the kernel arithmetic can be checked against the oracle in the
// GPU-less build container.  Exports the same C ABI as libmrhyde_b200.so so tests drive both with one wrapper.
// The product library never links or calls this file; without a GPU the product fails with MRH_ECUDA.
#include <memory>
#include "../../mrhyde_b200/csrc/capi_host.inl"
#include "../../mrhyde_b200/csrc/dispatch.h"
#include "../../mrhyde_b200/csrc/assemble_fast.h"
#include "../../mrhyde_b200/csrc/assemble_fast2d.h"
#include "../../mrhyde_b200/csrc/assemble_batched.h"

using namespace mrh;

struct EmuImpl {
  std::vector<uint16_t> pat; std::vector<int> pat_id;
  int64_t launches = 0, fast_launches = 0, batched_launches = 0;
};

static int backend_create(mrh_engine* h) { h->impl = new EmuImpl(); return MRH_OK; }
static void backend_destroy(mrh_engine* h) { delete (EmuImpl*)h->impl; }
static void backend_nodes_changed(mrh_engine*) {}      // the stepping harness reads the host arrays directly

template <class C>
static void run_launch(const KParams& kp) {
  using G = GenPhases<C>;
  const int EB = 3;
  const bool jac = kp.compute_jac != 0;
  GenSmem<C> L(EB, kp.tab.nq, jac, kp.tp.transient != 0);
  std::vector<double> sm(L.total_doubles());
  for (int64_t blk0 = 0; blk0 < kp.nwork; blk0 += EB) {
    for (int it = 0; it < EB * C::N; ++it) G::gather_dof(kp, L, sm.data(), blk0, it);
    for (int it = 0; it < EB * C::NV; ++it) G::gather_vertex(kp, L, sm.data(), blk0, it);
    for (int it = 0; it < EB * L.nq; ++it) G::geometry(kp, L, sm.data(), blk0, it);
    for (int it = 0; it < EB * L.nq; ++it) {
      if (jac) G::template point<PAD<C::P>>(kp, L, sm.data(), blk0, it);
      else G::template point<double>(kp, L, sm.data(), blk0, it);
    }
    for (int it = 0; it < EB * C::NB * C::NVAR; ++it) G::contract(kp, L, sm.data(), blk0, it);
  }
}

// the batched volume kernel (batched_kernels.cu), stepped phase by phase: one "warp" of 32 explicit lanes per element
template <class C>
static bool run_batched(const KParams& kp) {
  using G = GenPhases<C>;
  using B = BatPhases<C>;
  WarpSmem<C> L(kp.tab.nq, kp.tp.transient != 0);
  std::vector<double> sm(L.total_doubles());
  std::vector<int> offs(kp.offsets, kp.offsets + C::N);
  const int nq = L.nq, nparts = B::geom_parts(nq), n1 = B::num_kind1(kp);
  auto acc = std::make_unique<typename B::Acc>();
  for (int64_t w = 0; w < kp.nwork; ++w) {
    for (int it = 0; it < C::N; ++it) G::gather_dof(kp, L, sm.data(), w, it);
    for (int it = 0; it < C::NV; ++it) G::gather_vertex(kp, L, sm.data(), w, it);
    for (int it = 0; it < nq * nparts; ++it) B::geometry(kp, L, sm.data(), it, nparts);
    for (int it = 0; it < nq * n1; ++it) B::prep(kp, L, sm.data(), it, n1);
    if (NeedsElementSize<C::PHYS>::value || kp.is_side) B::element_size(L, sm.data(), kp.is_side != 0);
    for (int it = 0; it < nq * B::NCH; ++it) B::point_chunk(kp, L, sm.data(), it);
    B::contract(kp, L, sm.data(), *acc);
    B::store_acc(L, sm.data(), *acc, kp.compute_jac != 0);
    for (int lane = 0; lane < 32; ++lane) B::row_setup(kp, L, sm.data(), offs.data(), lane);
    for (int lane = 0; lane < 32; ++lane) B::scatter(kp, L, sm.data(), offs.data(), G::elem_of(kp, w), lane);
  }
  return true;
}

static int emu_assemble(mrh_engine* h, const mrh_state* st, double* res, double* jac) {
  auto& H = h->host; auto* E = (EmuImpl*)h->impl;
  if (!H.host_final || E->pat_id.empty()) throw std::runtime_error("mrh_finalize has not been called");
  BasePtrs bp;
  bp.nodes = H.mesh.nodes.data(); bp.e2n = H.mesh.elem2node.data(); bp.lids = H.dofs.lids.data(); bp.fixed = H.isFixed.data();
  bp.rowptr = H.crs.rowptr.data(); bp.pat = E->pat.data(); bp.pat_id = E->pat_id.data();
  bp.code = H.ec.code.data(); bp.cpool = H.ec.cpool.data();
  auto tab = [](const PointTables& T) { RefTab r{}; r.B = T.B.data(); r.G = T.G.data(); r.VB = T.VB.data(); r.VG = T.VG.data(); r.W = T.W.data(); r.FR = T.FR.data(); r.nq = T.nq;
    for (int c = 0; c < 3; ++c) { r.tanU[c] = T.tanU[c]; r.tanV[c] = T.tanV[c]; } return r; };
  bp.vol_tab = tab(H.ref.vol);
  for (int s = 0; s < H.ref.nsides; ++s) bp.side_tab[s] = tab(H.ref.side[s]);
  E->launches = 0; E->fast_launches = 0; E->batched_launches = 0;
  for (int side = -1; side < (int)H.side_launches.size(); ++side) {
    if (side >= 0 && !H.side_launches[side].active) continue;
    KParams kp;
    H.fillParams(kp, bp, *st, res, jac, side, side >= 0 ? H.mesh.sides[side].elems.data() : nullptr);
    if (side < 0 && !(H.cfg.flags & 1) &&
        thermal_hex_q1_eligible(H.phys, H.dim, H.order, H.ref.vol.nq, H.volumeStateIndependent(), H.cfg.lump_mass)) {
      const bool tr = kp.tp.transient != 0, adv = H.cfg.include_advection != 0, jc = jac != nullptr;
      for (int64_t e = 0; e < kp.nwork; ++e) {
#define MRH_FAST(T, A, Jc) if (tr == T && adv == A && jc == Jc) { double stage[48]; ThermalHexQ1<T, A, Jc>::element(kp, e, stage, 1); }
        MRH_FAST(false, false, false) MRH_FAST(false, false, true) MRH_FAST(false, true, false) MRH_FAST(false, true, true)
        MRH_FAST(true, false, false) MRH_FAST(true, false, true) MRH_FAST(true, true, false) MRH_FAST(true, true, true)
#undef MRH_FAST
      }
      E->launches++; E->fast_launches++;
      continue;
    }
    if (side < 0 && !(H.cfg.flags & 1) &&
        thermal_quad_q1_eligible(H.phys, H.dim, H.order, H.ref.vol.nq, H.volumeStateIndependent(), H.cfg.lump_mass)) {
      const bool tr = kp.tp.transient != 0, adv = H.cfg.include_advection != 0, jc = jac != nullptr;
      for (int64_t e = 0; e < kp.nwork; ++e) {
#define MRH_FAST2D(T, A, Jc) if (tr == T && adv == A && jc == Jc) ThermalQuadQ1<T, A, Jc>::element(kp, e);
        MRH_FAST2D(false, false, false) MRH_FAST2D(false, false, true) MRH_FAST2D(false, true, false) MRH_FAST2D(false, true, true)
        MRH_FAST2D(true, false, false) MRH_FAST2D(true, false, true) MRH_FAST2D(true, true, false) MRH_FAST2D(true, true, true)
#undef MRH_FAST2D
      }
      E->launches++; E->fast_launches++;
      continue;
    }
    if (!(H.cfg.flags & 16)) {          // volume and side launches through the batched phases
      if (!dispatch_cfg(H.phys, H.dim, H.order, [&](auto c) { run_batched<decltype(c)>(kp); })) throw std::runtime_error("configuration not supported");
      E->launches++; if (side < 0) E->batched_launches++;
      continue;
    }
    if (!dispatch_cfg(H.phys, H.dim, H.order, [&](auto c) { run_launch<decltype(c)>(kp); }))
      throw std::runtime_error("configuration not supported");
    E->launches++;
  }
  if (jac) {
    auto diag = [&](int r) { for (int64_t p = H.crs.rowptr[r]; p < H.crs.rowptr[r + 1]; ++p) if (H.crs.colind[p] == r) return p; return (int64_t)-1; };
    for (int d : H.dbc_dofs) { const int64_t p = diag(d); if (p >= 0) jac[p] = 1.0; }
    if (H.cfg.fix_zero_rows)
      for (int64_t r = 0; r < H.crs.nrows; ++r) {
        double s = 0.0; for (int64_t p = H.crs.rowptr[r]; p < H.crs.rowptr[r + 1]; ++p) s += std::fabs(jac[p]);
        if (s < 1.0e-14) { const int64_t p = diag((int)r); if (p >= 0) jac[p] = 1.0; }
      }
  }
  return MRH_OK;
}

// L2-projection systems stepped on the CPU (GenPhases::projection)
template <class C>
static void run_projection(const KParams& kp) {
  using G = GenPhases<C>;
  const int EB = 2;
  GenSmem<C> L(EB, kp.tab.nq, false, false);
  std::vector<double> sm(L.total_doubles());
  for (int64_t blk0 = 0; blk0 < kp.nwork; blk0 += EB) {
    for (int it = 0; it < EB * C::N; ++it) {
      const int le = it / C::N, k = it % C::N;
      if (blk0 + le < kp.nwork) reinterpret_cast<int*>(sm.data() + L.oLid())[le * C::N + k] = kp.lids[G::elem_of(kp, blk0 + le) * C::N + k];
    }
    for (int it = 0; it < EB * C::NV; ++it) G::gather_vertex(kp, L, sm.data(), blk0, it);
    for (int it = 0; it < EB * L.nq; ++it) G::geometry(kp, L, sm.data(), blk0, it);
    for (int it = 0; it < EB * C::NB * C::NVAR; ++it) G::projection(kp, L, sm.data(), blk0, it);
  }
}
extern "C" {

int mrh_finalize(mrh_handle h) {
  if (!h) return MRH_EINVAL;
  MRH_TRY(h)
  auto& H = h->host; auto* E = (EmuImpl*)h->impl;
  H.finalizeHost();
  const int N = H.N;
  E->pat.assign((size_t)H.mesh.nelem * N * N, 0); E->pat_id.resize((size_t)H.mesh.nelem);
  for (int64_t e = 0; e < H.mesh.nelem; ++e) {
    E->pat_id[e] = (int)e;
    const int* l = &H.dofs.lids[(size_t)e * N];
    for (int r = 0; r < N; ++r) {
      const int64_t b = H.crs.rowptr[l[r]], en = H.crs.rowptr[l[r] + 1];
      for (int c = 0; c < N; ++c) {
        const int* it = std::lower_bound(&H.crs.colind[b], &H.crs.colind[0] + en, l[c]);
        if (it == &H.crs.colind[0] + en || *it != l[c]) throw std::runtime_error("CRS graph misses an element coupling");
        E->pat[((size_t)e * N + r) * N + c] = (uint16_t)(it - &H.crs.colind[b]);
      }
    }
  }
  return MRH_OK;
  MRH_CATCH(h)
}
int mrh_assemble_res(mrh_handle h, const mrh_state* st, double* res) {
  if (!h || !st || !res) return MRH_EINVAL;
  MRH_TRY(h) return emu_assemble(h, st, res, nullptr); MRH_CATCH(h)
}
int mrh_assemble_jacres(mrh_handle h, const mrh_state* st, double* res, double* jac) {
  if (!h || !st || !res || !jac) return MRH_EINVAL;
  MRH_TRY(h) return emu_assemble(h, st, res, jac); MRH_CATCH(h)
}
int mrh_assemble_jacres_host(mrh_handle h, const mrh_state* st, double* res, double* jac) {
  if (!h || !st || !res || !jac) return MRH_EINVAL;
  MRH_TRY(h)
  memset(res, 0, sizeof(double) * h->host.dofs.nlocal); memset(jac, 0, sizeof(double) * h->host.crs.colind.size());
  return emu_assemble(h, st, res, jac);
  MRH_CATCH(h)
}
int mrh_assemble_projection(mrh_handle h, int32_t kind, double time, double* rhs, double* mass, void*) {
  if (!h || !rhs || (kind != 0 && kind != 1)) return MRH_EINVAL;
  MRH_TRY(h)
  auto& H = h->host; auto* E = (EmuImpl*)h->impl;
  if (!H.host_final || E->pat_id.empty()) throw std::runtime_error("mrh_finalize has not been called");
  if (!H.proj_unsupported.empty()) throw std::runtime_error(H.proj_unsupported + " (not supported)");
  BasePtrs bp;
  bp.nodes = H.mesh.nodes.data(); bp.e2n = H.mesh.elem2node.data(); bp.lids = H.dofs.lids.data(); bp.fixed = H.isFixed.data();
  bp.rowptr = H.crs.rowptr.data(); bp.pat = E->pat.data(); bp.pat_id = E->pat_id.data();
  bp.code = H.ec.code.data(); bp.cpool = H.ec.cpool.data();
  auto tab = [](const PointTables& T) { RefTab r{}; r.B = T.B.data(); r.G = T.G.data(); r.VB = T.VB.data(); r.VG = T.VG.data(); r.W = T.W.data(); r.FR = T.FR.data(); r.nq = T.nq;
    for (int c = 0; c < 3; ++c) { r.tanU[c] = T.tanU[c]; r.tanV[c] = T.tanV[c]; } return r; };
  bp.vol_tab = tab(H.ref.vol);
  for (int s = 0; s < H.ref.nsides; ++s) bp.side_tab[s] = tab(H.ref.side[s]);
  auto diag = [&](int r) { for (int64_t p = H.crs.rowptr[r]; p < H.crs.rowptr[r + 1]; ++p) if (H.crs.colind[p] == r) return p; return (int64_t)-1; };
  if (kind == 0) {
    for (size_t k = 0; k < H.dirichlet_launches.size(); ++k) {
      KParams kp;
      H.fillProjectionParams(kp, bp, time, rhs, mass, (int)k, H.mesh.sides[H.dirichlet_launches[k].side].elems.data());
      if (!dispatch_cfg(H.phys, H.dim, H.order, [&](auto c) { run_projection<decltype(c)>(kp); })) throw std::runtime_error("configuration not supported");
    }
    if (mass) for (int64_t r = 0; r < H.crs.nrows; ++r) if (!H.isFixed[r]) { const int64_t p = diag((int)r); if (p >= 0) mass[p] = 1.0; }
  } else {
    KParams kp;
    H.fillProjectionParams(kp, bp, time, rhs, mass, -1, nullptr);
    if (!dispatch_cfg(H.phys, H.dim, H.order, [&](auto c) { run_projection<decltype(c)>(kp); })) throw std::runtime_error("configuration not supported");
    if (mass) for (int64_t r = 0; r < H.crs.nrows; ++r) {
      double sum = 0.0; for (int64_t p = H.crs.rowptr[r]; p < H.crs.rowptr[r + 1]; ++p) sum += std::fabs(mass[p]);
      if (sum < 1.0e-14) { const int64_t p = diag((int)r); if (p >= 0) mass[p] = 1.0; }
    }
  }
  return MRH_OK;
  MRH_CATCH(h)
}
int mrh_set_dirichlet(mrh_handle h, double*, double, double*, double*, const mrh_solver_opts*, mrh_solver_info*, void*) { if (h) h->err = "not supported by the CPU stepping harness"; return MRH_EUNSUPPORTED; }
int mrh_set_initial(mrh_handle h, double*, double*, double*, const mrh_solver_opts*, mrh_solver_info*, void*) { if (h) h->err = "not supported by the CPU stepping harness"; return MRH_EUNSUPPORTED; }
int mrh_comm_unique_id(void*) { return MRH_EUNSUPPORTED; }
int mrh_comm_init(mrh_handle h, const void*, int32_t, int32_t) { if (h) h->err = "not supported by the CPU stepping harness"; return MRH_EUNSUPPORTED; }
int mrh_exchange(mrh_handle h, double*, double*, void*) { if (h) h->err = "not supported by the CPU stepping harness"; return MRH_EUNSUPPORTED; }
int64_t mrh_get_ghost_block(mrh_handle h, int32_t* rows, int64_t* col_gids, double*) {
  if (!h) return MRH_EINVAL;
  const auto& P = h->host.plan; const int64_t n = (int64_t)P.ghost_row.size();
  if (rows) memcpy(rows, P.ghost_row.data(), n * sizeof(int));
  if (col_gids) memcpy(col_gids, P.ghost_colgid.data(), n * sizeof(int64_t));
  return n;
}
int mrh_assemble_res_host_static(mrh_handle h, const mrh_state*, double*) {      // device-only entry point
  if (!h) return MRH_EINVAL;
  h->err = "not supported by the CPU stepping harness"; return MRH_EUNSUPPORTED;
}
const char* mrh_last_kernel_path(mrh_handle h) {
  if (!h) return "generic";
  auto* E = (EmuImpl*)h->impl;
  return E->fast_launches ? "specialised" : E->batched_launches ? "batched" : "generic";
}
int64_t mrh_last_launch_count(mrh_handle h) { return h ? ((EmuImpl*)h->impl)->launches : 0; }
double mrh_last_volume_kernel_ms(mrh_handle h) { return h ? -(double)((EmuImpl*)h->impl)->fast_launches : 0.0; }  // emu: -#specialised launches
int mrh_microbench(int32_t, double*, double*, double*) { return MRH_EUNSUPPORTED; }

}  // extern "C"
