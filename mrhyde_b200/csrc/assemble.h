// Fused element assembly: gather -> geometry/basis at quadrature points -> point physics with derivative
// lanes -> basis-by-quadrature contraction -> scatter-add into the CRS matrix and the residual vector.
// Replaces, for all elements of a block (or of one side set) in ONE launch, the reference's per-group chain
//   performGather            /root/reference/src/managers/assemblyManager.cpp:3598-3643
//   updateWorkset            :6512-6596  (resetResidual, computeSoln{Steady,Transient}Seeded workset.cpp:479-859)
//   getPhysicalVolumetricBasis / getPhysicalBoundary*  /root/reference/src/interfaces/discretizationInterface.cpp:898-981, :1608-1887
//   evaluateSolutionField    /root/reference/src/tools/workset.cpp:937-1185
//   physics volumeResidual / boundaryResidual / fluxConditions
//   scatter / scatterRes     assemblyManager.cpp:4031-4145, :4154-4202
// The phase functions are __host__ __device__ so that tests/emu can step the very same code on the CPU;
// the product only ever launches them from kernels.cu.
#pragma once
#include "physics.h"

namespace mrh {

MRH_HD void atomic_add(double* p, double v) {
#if defined(__CUDA_ARCH__)
  atomicAdd(p, v);          // result unused: compiles to RED.E.ADD.F64
#else
  *p += v;
#endif
}

template <int DIM> MRH_HD double det_inv(const double (&J)[DIM][DIM], double (&Ji)[DIM][DIM]);
template <> MRH_HD double det_inv<2>(const double (&J)[2][2], double (&Ji)[2][2]) {
  const double det = J[0][0] * J[1][1] - J[0][1] * J[1][0];
  const double id = 1.0 / det;
  Ji[0][0] = J[1][1] * id; Ji[0][1] = -J[0][1] * id; Ji[1][0] = -J[1][0] * id; Ji[1][1] = J[0][0] * id;
  return det;
}
template <> MRH_HD double det_inv<3>(const double (&J)[3][3], double (&Ji)[3][3]) {
  const double c0 = J[1][1] * J[2][2] - J[1][2] * J[2][1], c1 = J[1][2] * J[2][0] - J[1][0] * J[2][2], c2 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
  const double det = J[0][0] * c0 + J[0][1] * c1 + J[0][2] * c2;
  const double id = 1.0 / det;
  Ji[0][0] = c0 * id; Ji[0][1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) * id; Ji[0][2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) * id;
  Ji[1][0] = c1 * id; Ji[1][1] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) * id; Ji[1][2] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) * id;
  Ji[2][0] = c2 * id; Ji[2][1] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) * id; Ji[2][2] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) * id;
  return det;
}

// ================================================================================================
// Generic path: any of the four physics, any state-dependent coefficient, volume or side launch.
// A block of EB elements is staged in shared memory; work items are spread over the CTA's threads.
// ================================================================================================
template <int PHYS_, int DIM_, int ORDER_>
struct GenCfg {
  static constexpr int PHYS = PHYS_, DIM = DIM_, ORDER = ORDER_;
  static constexpr int NVAR = PhysTraits<PHYS_, DIM_>::NVAR;
  static constexpr int NB = (DIM_ == 3) ? (ORDER_ + 1) * (ORDER_ + 1) * (ORDER_ + 1) : (ORDER_ + 1) * (ORDER_ + 1);
  static constexpr int NV = (DIM_ == 3) ? 8 : 4;
  static constexpr int N = NVAR * NB;
  static constexpr int P = NVAR * (1 + DIM_);
  static constexpr int TA = 1 + DIM_;
};

// shared-memory carve-up for EB elements and nq points (offsets in doubles)
template <class C>
struct GenSmem {
  int EB, nq;
  bool jac, transient;
  MRH_HD GenSmem(int eb, int nq_, bool j, bool tr) : EB(eb), nq(nq_), jac(j), transient(tr) {}
  MRH_HD size_t oX() const { return 0; }
  MRH_HD size_t oU() const { return oX() + (size_t)EB * C::NV * C::DIM; }
  MRH_HD size_t oUd() const { return oU() + (size_t)EB * C::N; }
  MRH_HD size_t oW() const { return oUd() + (transient ? (size_t)EB * C::N : 0); }
  MRH_HD size_t oXq() const { return oW() + (size_t)EB * nq; }
  MRH_HD size_t oNrm() const { return oXq() + (size_t)EB * nq * 3; }
  MRH_HD size_t oT() const { return oNrm() + (size_t)EB * nq * 6; }
  MRH_HD size_t oF() const { return oT() + (size_t)EB * nq * C::NB * C::TA; }
  MRH_HD size_t oD() const { return oF() + (size_t)EB * nq * C::P; }
  MRH_HD size_t oLid() const { return oD() + (jac ? (size_t)EB * nq * C::P * C::P : 0); }
  MRH_HD size_t total_doubles() const { return oLid() + ((size_t)EB * C::N + 1) / 2 + 1; }
};

template <class C>
struct GenPhases {
  static constexpr int DIM = C::DIM, NB = C::NB, NV = C::NV, N = C::N, NVAR = C::NVAR, P = C::P, TA = C::TA;

  MRH_HD static int64_t elem_of(const KParams& kp, int64_t w) { return kp.elem_list ? (int64_t)kp.elem_list[w] : w; }

  // ---- phase 0: gather (items: EB*N dof slots, then EB*NV vertices) --------------------------------
  template <class LAY> MRH_HDN static void gather_dof(const KParams& kp, const LAY& L, double* sm, int64_t blk0, int item) {
    const int le = item / N, k = item % N;
    if (blk0 + le >= kp.nwork) return;
    const int64_t e = elem_of(kp, blk0 + le);
    const int lid = kp.lids[e * N + k];
    int* sLid = reinterpret_cast<int*>(sm + L.oLid());
    sLid[le * N + k] = lid;
    const double uval = kp.u[lid];
    if (!kp.tp.transient) { sm[L.oU() + (size_t)le * N + k] = uval; return; }
    const TimeParams& tp = kp.tp;
    const double prev0 = kp.u_prev[(size_t)lid * tp.nsteps];
    double beta_u = (1.0 - tp.alpha_u) * prev0;
    for (int s = 0; s < tp.stage; ++s) beta_u += tp.a_over_b[s] * (kp.u_stage[(size_t)lid * tp.nstages + s] - prev0);
    double beta_t = 0.0;
    for (int s = 1; s <= tp.nsteps; ++s) beta_t += tp.bdf[s] * kp.u_prev[(size_t)lid * tp.nsteps + (s - 1)];
    beta_t *= tp.timewt;
    sm[L.oU() + (size_t)le * N + k] = tp.alpha_u * uval + beta_u;
    sm[L.oUd() + (size_t)le * N + k] = tp.alpha_t * uval + beta_t;
  }
  template <class LAY> MRH_HDN static void gather_vertex(const KParams& kp, const LAY& L, double* sm, int64_t blk0, int item) {
    const int le = item / NV, a = item % NV;
    if (blk0 + le >= kp.nwork) return;
    const int64_t e = elem_of(kp, blk0 + le);
    const int64_t nd = kp.e2n[e * NV + a];
    for (int c = 0; c < DIM; ++c) sm[L.oX() + ((size_t)le * NV + a) * DIM + c] = kp.nodes[nd * DIM + c];
  }

  // ---- phase 1a: geometry + physical basis at one (element, point) ---------------------------------
  template <class LAY> MRH_HDN static void geometry(const KParams& kp, const LAY& L, double* sm, int64_t blk0, int item) {
    const int nq = L.nq, le = item / nq, q = item % nq;
    if (blk0 + le >= kp.nwork) return;
    const double* X = sm + L.oX() + (size_t)le * NV * DIM;
    const RefTab& tb = kp.tab;
    double J[DIM][DIM], Ji[DIM][DIM], xq[3] = {0.0, 0.0, 0.0};
    for (int i = 0; i < DIM; ++i) for (int j = 0; j < DIM; ++j) J[i][j] = 0.0;
    for (int a = 0; a < NV; ++a) {
      const double vb = tb.VB[q * NV + a];
      for (int i = 0; i < DIM; ++i) {
        const double xa = X[a * DIM + i];
        xq[i] += xa * vb;
        for (int j = 0; j < DIM; ++j) J[i][j] += xa * tb.VG[(q * NV + a) * DIM + j];
      }
    }
    const double det = det_inv<DIM>(J, Ji);
    double w, nrm[3] = {0.0, 0.0, 0.0}, tng[3] = {0.0, 0.0, 0.0};
    if (!kp.is_side) w = fabs(det) * tb.W[q];
    else {
      double measure;
      if (DIM == 2) {
        for (int i = 0; i < 2; ++i) { tng[i] = 0.0; for (int j = 0; j < 2; ++j) tng[i] += J[i][j] * tb.tanU[j]; }
        nrm[0] = tng[1]; nrm[1] = -tng[0];
        measure = sqrt(tng[0] * tng[0] + tng[1] * tng[1]);
      } else {
        double tu[3], tv[3];
        for (int i = 0; i < DIM; ++i) { tu[i] = tv[i] = 0.0; for (int j = 0; j < DIM; ++j) { tu[i] += J[i][j] * tb.tanU[j]; tv[i] += J[i][j] * tb.tanV[j]; } }
        nrm[0] = tu[1] * tv[2] - tu[2] * tv[1]; nrm[1] = tu[2] * tv[0] - tu[0] * tv[2]; nrm[2] = tu[0] * tv[1] - tu[1] * tv[0];
        measure = sqrt(nrm[0] * nrm[0] + nrm[1] * nrm[1] + nrm[2] * nrm[2]);
      }
      w = measure * tb.W[q];
      double nl = 0.0; for (int c = 0; c < DIM; ++c) nl += nrm[c] * nrm[c];
      nl = 1.0 / sqrt(nl);
      for (int c = 0; c < DIM; ++c) nrm[c] *= nl;
    }
    sm[L.oW() + (size_t)le * nq + q] = w;
    for (int c = 0; c < 3; ++c) {
      sm[L.oXq() + ((size_t)le * nq + q) * 3 + c] = xq[c];
      sm[L.oNrm() + ((size_t)le * nq + q) * 6 + c] = nrm[c];
      sm[L.oNrm() + ((size_t)le * nq + q) * 6 + 3 + c] = tng[c];
    }
    double* T = sm + L.oT() + ((size_t)le * nq + q) * NB * TA;
    for (int j = 0; j < NB; ++j) {
      T[j * TA] = tb.B[q * NB + j];
      for (int c = 0; c < DIM; ++c) {
        double s = 0.0;
        for (int m = 0; m < DIM; ++m) s += Ji[m][c] * tb.G[(q * NB + j) * DIM + m];
        T[j * TA + 1 + c] = s;
      }
    }
  }

  // ---- phase 1b: fields + point physics at one (element, point) ------------------------------------
  template <class S>
  MRH_HDN static void point(const KParams& kp, const GenSmem<C>& L, double* sm, int64_t blk0, int item) {
    const int nq = L.nq, le = item / nq, q = item % nq;
    if (blk0 + le >= kp.nwork) return;
    const double* T = sm + L.oT() + ((size_t)le * nq + q) * NB * TA;
    const double* U = sm + L.oU() + (size_t)le * N;
    const double* Ud = sm + L.oUd() + (size_t)le * N;
    S fields[NVAR * FIELD_KINDS];
    constexpr int PL = lanes_of<S>::value;
    for (int m = 0; m < NVAR; ++m) {
      double f[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
      for (int j = 0; j < NB; ++j) {
        const int slot = kp.offsets[m * NB + j];
        const double uj = U[slot];
        f[0] += uj * T[j * TA];
        for (int c = 0; c < DIM; ++c) f[1 + c] += uj * T[j * TA + 1 + c];
        if (kp.tp.transient && !kp.is_side) f[4] += Ud[slot] * T[j * TA];
      }
      for (int k = 0; k < FIELD_KINDS; ++k) fields[m * FIELD_KINDS + k] = S(k < 5 ? f[k] : 0.0);
      if (PL > 0 && kp.compute_jac) {
        const double au = kp.tp.transient ? kp.tp.alpha_u : 1.0;
        set_lane(fields[m * FIELD_KINDS + 0], m * TA, au);
        for (int c = 0; c < DIM; ++c) set_lane(fields[m * FIELD_KINDS + 1 + c], m * TA + 1 + c, au);
        if (kp.tp.transient && !kp.is_side) set_lane(fields[m * FIELD_KINDS + 4], m * TA, kp.tp.alpha_t);
      }
    }
    PointIn<S> in;
    in.kp = &kp; in.ps.fields = fields; in.ps.time = kp.tp.time;
    for (int c = 0; c < 3; ++c) {
      in.ps.x[c] = sm[L.oXq() + ((size_t)le * nq + q) * 3 + c];
      in.ps.nrm[c] = sm[L.oNrm() + ((size_t)le * nq + q) * 6 + c];
      in.ps.tng[c] = sm[L.oNrm() + ((size_t)le * nq + q) * 6 + 3 + c];
    }
    in.w = sm[L.oW() + (size_t)le * nq + q];
    double vol = 0.0;
    for (int k = 0; k < nq; ++k) vol += sm[L.oW() + (size_t)le * nq + k];
    in.h = pow(vol, 1.0 / (kp.is_side ? (double)(DIM - 1) : (double)DIM));
    S F[NVAR][TA];
    for (int v = 0; v < NVAR; ++v) for (int a = 0; a < TA; ++a) F[v][a] = S(0.0);
    if (!kp.is_side) Physics<C::PHYS, DIM, S>::volume(in, F);
    else {
      Physics<C::PHYS, DIM, S>::boundary(in, F);
      for (int v = 0; v < NVAR; ++v)        // generic "Flux" condition, physicsInterface.cpp:1698-1766
        if (kp.ph.bc[v] == BC_FLUX) F[v][0] = F[v][0] + (-in.fn(SIDE_FLUX + v)) * in.w;
    }
    double* sF = sm + L.oF() + ((size_t)le * nq + q) * P;
    for (int v = 0; v < NVAR; ++v) for (int a = 0; a < TA; ++a) sF[v * TA + a] = val(F[v][a]);
    if (PL > 0 && L.jac) {
      double* sD = sm + L.oD() + ((size_t)le * nq + q) * P * P;
      for (int v = 0; v < NVAR; ++v) for (int a = 0; a < TA; ++a)
        for (int l = 0; l < P; ++l) sD[(v * TA + a) * P + l] = lane(F[v][a], l);
    }
  }
  // ---- L2-projection systems (no AD): item = (element, variable v, basis function i) -----------------------------
  // PROJ_DIRICHLET (side launch): rows of FIXED DOFs collect the side mass matrix of their variable and the side moments of the
  //   Dirichlet function  - AssemblyManager::setDirichlet / getDirichletBoundary / getMassBoundary
  //   (/root/reference/src/managers/assemblyManager.cpp:1855-1948, 6288-6420);
  // PROJ_INITIAL (volume launch): every row collects the element mass matrix of its variable and the moments of "initial <var>"
  //   - AssemblyManager::setInitial (:1185-1310).  rhs is ADDED to kp.res, the mass entries to kp.jac (lumped when lump_mass).
  MRH_HDN static void projection(const KParams& kp, const GenSmem<C>& L, double* sm, int64_t blk0, int item) {
    const int nq = L.nq;
    const int v = item % NVAR, i = (item / NVAR) % NB, le = item / (NVAR * NB);
    if (blk0 + le >= kp.nwork) return;
    const bool dir = kp.ph.proj == PROJ_DIRICHLET;
    if (dir && kp.ph.bc[v] != BC_DIRICHLET) return;
    const int* sLid = reinterpret_cast<const int*>(sm + L.oLid()) + le * N;
    const int rowslot = kp.offsets[v * NB + i];
    const int row = sLid[rowslot];
    if (dir && !kp.fixed[row]) return;
    const Prog& pr = kp.progs[(dir ? SIDE_BCSRC : VOL_INIT) + v];
    double dval = 0.0, m[NB];
    for (int k = 0; k < NB; ++k) m[k] = 0.0;
    for (int q = 0; q < nq; ++q) {
      const double* T = sm + L.oT() + ((size_t)le * nq + q) * NB * TA;
      const double w = sm[L.oW() + (size_t)le * nq + q];
      PointState<double> ps;
      for (int c = 0; c < 3; ++c) {
        ps.x[c] = sm[L.oXq() + ((size_t)le * nq + q) * 3 + c];
        ps.nrm[c] = sm[L.oNrm() + ((size_t)le * nq + q) * 6 + c];
        ps.tng[c] = sm[L.oNrm() + ((size_t)le * nq + q) * 6 + 3 + c];
      }
      ps.time = kp.tp.time; ps.fields = nullptr;
      const double g = (pr.kind == 0) ? pr.cval : eval_code<double>(kp.code, kp.cpool, pr.start, pr.len, ps);
      const double Tiw = T[i * TA] * w;
      dval += g * Tiw;
      for (int k = 0; k < NB; ++k) m[k] += Tiw * T[k * TA];
    }
    atomic_add(kp.res + row, dval);
    if (!kp.jac) return;
    const int64_t e = elem_of(kp, blk0 + le);
    const uint16_t* pat = kp.pat + (size_t)kp.pat_id[e] * N * N + (size_t)rowslot * N;
    double* jrow = kp.jac + kp.rowptr[row];
    if (kp.lump_mass && !dir) {
      double s = 0.0;
      for (int k = 0; k < NB; ++k) s += m[k];
      atomic_add(jrow + pat[rowslot], s);
      return;
    }
    for (int k = 0; k < NB; ++k) atomic_add(jrow + pat[kp.offsets[v * NB + k]], m[k]);
  }
  MRH_HD static void set_lane(double&, int, double) {}
  template <int PP> MRH_HD static void set_lane(PAD<PP>& x, int l, double v) { x.d[l] = v; }

  // ---- phase 2: contraction + scatter for one (element, test node i, row variable v) ---------------
  MRH_HDN static void contract(const KParams& kp, const GenSmem<C>& L, double* sm, int64_t blk0, int item) {
    const int nq = L.nq;
    const int v = item % NVAR, i = (item / NVAR) % NB, le = item / (NVAR * NB);
    if (blk0 + le >= kp.nwork) return;
    const int* sLid = reinterpret_cast<const int*>(sm + L.oLid()) + le * N;
    const int rowslot = kp.offsets[v * NB + i];
    const int row = sLid[rowslot];
    if (kp.fixed[row]) return;                                   // isFixedDOF rows are never scattered into
    double r = 0.0;
    double acc[NVAR][NB];
    for (int m = 0; m < NVAR; ++m) for (int j = 0; j < NB; ++j) acc[m][j] = 0.0;
    for (int q = 0; q < nq; ++q) {
      const double* T = sm + L.oT() + ((size_t)le * nq + q) * NB * TA;
      const double* sF = sm + L.oF() + ((size_t)le * nq + q) * P;
      double Ti[TA];
      for (int a = 0; a < TA; ++a) { Ti[a] = T[i * TA + a]; r += Ti[a] * sF[v * TA + a]; }
      if (!L.jac) continue;
      const double* sD = sm + L.oD() + ((size_t)le * nq + q) * P * P;
      double Lv[P];
      for (int l = 0; l < P; ++l) {
        double s = 0.0;
        for (int a = 0; a < TA; ++a) s += Ti[a] * sD[(v * TA + a) * P + l];
        Lv[l] = s;
      }
      for (int j = 0; j < NB; ++j)
        for (int m = 0; m < NVAR; ++m) {
          double s = acc[m][j];
          for (int b = 0; b < TA; ++b) s += Lv[m * TA + b] * T[j * TA + b];
          acc[m][j] = s;
        }
    }
    atomic_add(kp.res + row, -r);
    if (!L.jac) return;
    const int64_t e = elem_of(kp, blk0 + le);
    const uint16_t* pat = kp.pat + (size_t)kp.pat_id[e] * N * N + (size_t)rowslot * N;
    double* jrow = kp.jac + kp.rowptr[row];
    if (kp.lump_mass) {
      double s = 0.0;
      for (int m = 0; m < NVAR; ++m) for (int j = 0; j < NB; ++j) s += acc[m][j];
      atomic_add(jrow + pat[rowslot], s);
      return;
    }
    for (int m = 0; m < NVAR; ++m)
      for (int j = 0; j < NB; ++j) atomic_add(jrow + pat[kp.offsets[m * NB + j]], acc[m][j]);
  }
};

}  // namespace mrh
