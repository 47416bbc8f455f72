// Batched fused assembly (volume terms, residual + Jacobian) for the multi-variable / higher-order configurations
// (C3 cdr Q2, C4 linear elasticity, C5 Navier-Stokes; also every state-dependent thermal case).  Same reference chain
// as assemble.h (performGather -> computeSoln*Seeded -> getPhysicalVolumetricBasis -> volumeResidual -> scatter,
// /root/reference/src/managers/assemblyManager.cpp:2150-2665, :3598-3643, :4031-4145), restructured around what the
// profiles of the generic kernel showed (profiles/r1_generic_c4_n64_*, r1_batched_v1_*): the PAD<P> interpreter spilled
// 10 KB per thread, only nq threads per element worked in the point phase, every RED lane hit its own sector, and the
// DFMA contraction was bound by shared-memory operand loads at 8 warps per SM.  Here ONE WARP owns ONE ELEMENT from
// gather to scatter (no CTA barrier; warps of an SM drift apart so that scatter, point physics and contraction overlap):
//   * state-independent coefficient programs are evaluated once per point on plain doubles (phase `prep`);
//   * the point physics runs on lane chunks: lane (point, variable m) carries PAD<1+dim> and differentiates the point
//     fluxes with respect to (u_m, grad u_m); fields and fluxes stay in registers;
//   * the element contraction K = sum_q T^T D T and r = sum_q T^T f runs on the FP64 tensor path
//     (mma.sync m8n8k4, DMMA.8x8x4): measured 37 TFLOP/s with one warp per scheduler (profiles/r1_dmma_microbench.txt)
//     against 8x the issue slots and one shared-memory operand per FMA for the DFMA form.  D is written by the point
//     phase directly in B-fragment order with the columns of a tile interleaved (m even | m odd) so that the C
//     fragment of the first product IS the A fragment of the second - no shuffle, no shared-memory round trip;
//   * the scatter walks a row of K with lanes along (node, variable), i.e. along consecutive CRS slots.
// Phase functions are __host__ __device__ and written against WV (one value per lane): tests/emu steps the same code
// on the CPU with WV = 32 explicit lanes.
#pragma once
#include "assemble.h"

namespace mrh {

#if !defined(MRH_NS_HALF_CHUNKS)
#define MRH_NS_HALF_CHUNKS 0     // measured on B200, C5 128^3: PAD<2> half-variable chunks at 10 warps / SM 29.9 ms, PAD<4> at 8 warps 23.1 ms
#endif
constexpr int MRH_NCOEF = 8;      // volume program slots of one physics module (params.h: at most CDR's 8)

// ---- one value per lane of a warp --------------------------------------------------------------------------
#if defined(__CUDA_ARCH__)
struct WV { double x; };
MRH_HD double wv_get(const WV& v, int) { return v.x; }
MRH_HD WV wv_zero() { return WV{0.0}; }
template <class F> MRH_HD WV wv_make(F f) { return WV{f((int)(threadIdx.x & 31))}; }
template <class F> MRH_HD void wv_each(F f) { f((int)(threadIdx.x & 31)); }
// c(8x8) += a(8x4) b(4x8): lane holds a[lane/4][lane%4], b[lane%4][lane/4], c[lane/4][2(lane%4) + {0,1}]
MRH_HD void wv_mma(WV& c0, WV& c1, const WV& a, const WV& b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0.x), "+d"(c1.x) : "d"(a.x), "d"(b.x));
}
MRH_HD void warp_sync() { __syncwarp(); }
#else
struct WV { double x[32]; };
MRH_HD double wv_get(const WV& v, int lane) { return v.x[lane]; }
MRH_HD WV wv_zero() { WV r; for (int l = 0; l < 32; ++l) r.x[l] = 0.0; return r; }
template <class F> MRH_HD WV wv_make(F f) { WV r; for (int l = 0; l < 32; ++l) r.x[l] = f(l); return r; }
template <class F> MRH_HD void wv_each(F f) { for (int l = 0; l < 32; ++l) f(l); }
MRH_HD void wv_mma(WV& c0, WV& c1, const WV& a, const WV& b) {
  for (int l = 0; l < 32; ++l) {
    const int row = l / 4, p = l % 4;
    for (int k = 0; k < 4; ++k) {
      c0.x[l] += a.x[row * 4 + k] * b.x[(2 * p) * 4 + k];
      c1.x[l] += a.x[row * 4 + k] * b.x[(2 * p + 1) * 4 + k];
    }
  }
}
MRH_HD void warp_sync() {}
#endif

MRH_HD void st2(double* p, double a, double b) {
#if defined(__CUDA_ARCH__)
  *reinterpret_cast<double2*>(p) = make_double2(a, b);
#else
  p[0] = a; p[1] = b;
#endif
}

MRH_HD size_t even_up(size_t x) { return (x + 1) & ~(size_t)1; }

// shared-memory carve-up of ONE element / warp (offsets in doubles, every region 16-byte aligned)
template <class C>
struct WarpSmem {
  int nq;
  bool transient;
  static constexpr int N = C::N, NVAR = C::NVAR, NB = C::NB, TA = C::TA;
  static constexpr int NT = (NVAR + 1) / 2;                    // column tiles of D: variables (2t, 2t+1)
  static constexpr int NRT = (NB + 7) / 8;                     // row tiles of 8 basis functions
  static constexpr int QS = NVAR * NT * 32 + 1;                // D fragments of one point; with dq() the lanes (q, m) of one store spread evenly over the banks
  static constexpr int KS = ((N + 7) / 16) * 16 + 8;           // row stride of K, == 8 (mod 16): conflict-free 16-byte stores
  static_assert(KS >= N, "row stride of the element matrix");
  MRH_HD WarpSmem(int nq_, bool tr) : nq(nq_), transient(tr) {}
  MRH_HD static size_t dq(int q) { return (size_t)q * QS + (size_t)((q >> 2) << 2); }
  MRH_HD size_t dsize() const { return even_up(dq(nq)); }
  MRH_HD size_t oX() const { return 0; }
  MRH_HD size_t oU() const { return even_up(oX() + (size_t)C::NV * C::DIM); }
  MRH_HD size_t oUd() const { return even_up(oU() + N); }
  MRH_HD size_t oW() const { return even_up(oUd() + (transient ? N : 0)); }
  MRH_HD size_t oXq() const { return even_up(oW() + nq); }
  MRH_HD size_t oNrm() const { return even_up(oXq() + (size_t)nq * 3); }                 // side launches: unit normal | tangent per point
  MRH_HD size_t oJi() const { return even_up(oNrm() + (size_t)nq * 6); }                 // J^-1 per point, [m][c]
  static constexpr bool STORE_T = NB <= 8;                     // physical basis rows kept in shared memory (Q1); rebuilt from J^-1 otherwise
  MRH_HD size_t oT() const { return even_up(oJi() + (size_t)nq * C::DIM * C::DIM); }     // T[q][j][0..dim], STORE_T only
  MRH_HD size_t oF() const { return even_up(oT() + (STORE_T ? (size_t)nq * NB * TA : 0)); }
  MRH_HD size_t oD() const { return even_up(oF() + (size_t)nq * NVAR * TA); }
  // F, D and the pre-evaluated coefficients are dead once the contraction has read them: the element matrix K takes
  // their place (a separate block only when even the three together are too small)
  MRH_HD size_t oCoef() const { return even_up(oD() + dsize()); }
  MRH_HD size_t dead_end() const { return even_up(oCoef() + (size_t)nq * MRH_NCOEF); }
  MRH_HD bool k_aliases_d() const { return (size_t)N * KS <= dead_end() - oF(); }
  MRH_HD size_t oKx() const { return dead_end(); }
  MRH_HD size_t oH() const { return even_up(oKx() + (k_aliases_d() ? 0 : (size_t)N * KS)); }
  MRH_HD size_t oR() const { return even_up(oH() + 1); }
  MRH_HD size_t oRowBase() const { return even_up(oR() + N); }                         // int64 per natural row
  MRH_HD size_t oPatOff() const { return even_up(oRowBase() + N); }                    // int per natural row: row offset in the slot pattern
  MRH_HD size_t oLid() const { return even_up(oPatOff() + (N + 1) / 2); }
  MRH_HD size_t total_doubles() const { return even_up(oLid() + (N + 1) / 2 + 1); }
  MRH_HD size_t oK() const { return k_aliases_d() ? oF() : oKx(); }
  // what GenPhases::gather_* expect of a layout
  int EB = 1;
};

// What a physics routine sees at one point in the batched kernel: same interface as PointIn<S> (physics.h), but the
// interpolated fields are a by-value array with compile-time indices (registers) and state-independent programs
// come pre-evaluated from shared memory.
template <class S, int NVAR, int DIM>
struct PointBat {
  const KParams* kp;
  S f[NVAR][5];            // value, d/dx, d/dy, d/dz, d/dt
  const double* coef;      // [MRH_NCOEF] pre-evaluated programs of kind 1 at this point
  PointState<S> ps;        // x, time; fields -> FIELD_KINDS-strided copy, written only when a state-dependent program exists
  double w, h;
  MRH_HD S fn(int k) const {
    const Prog& pr = kp->progs[k];
    if (pr.kind == 0) return S(pr.cval);
    if (pr.kind == 1) {
      if (k < MRH_NCOEF) return S(coef[k]);
      PointState<double> pd;                                 // side launches: boundary-condition functions (slots SIDE_BCSRC..)
      for (int c = 0; c < 3; ++c) { pd.x[c] = ps.x[c]; pd.nrm[c] = ps.nrm[c]; pd.tng[c] = ps.tng[c]; }
      pd.time = ps.time; pd.fields = nullptr;
      return S(eval_code<double>(kp->code, kp->cpool, pr.start, pr.len, pd));
    }
    return eval_code<S>(kp->code, kp->cpool, pr.start, pr.len, ps);
  }
  MRH_HD const S& U(int m) const { return f[m][0]; }
  MRH_HD const S& dU(int m, int c) const { return f[m][1 + c]; }
  MRH_HD const S& Ut(int m) const { return f[m][4]; }
};

// SIDES = false compiles the side-launch branches away (volume-only kernels: the NVRTC build and the ahead-of-time volume
// instantiation; measured on B200, C4 192^3: 86 ms with the branches present in the volume kernel against 61 ms without)
template <class C, bool SIDES = true>
struct BatPhases {
  MRH_HD static bool is_side(const KParams& kp) { if constexpr (SIDES) return kp.is_side != 0; else return false; }
  static constexpr int DIM = C::DIM, NB = C::NB, NV = C::NV, N = C::N, NVAR = C::NVAR, P = C::P, TA = C::TA;
  // lane chunk of the point phase: LCH of the P = NVAR (1 + dim) point-state lanes per work item.  Default one chunk per
  // variable (u_m, grad u_m).  MRH_NS_HALF_CHUNKS: 3-D Navier-Stokes with half a variable per item (PAD<2>, 168 instead of
  // 250 registers, 10 instead of 8 warps per SM) - slower in practice (the point values are recomputed twice as often)
  static constexpr int LCH = (C::PHYS == PHYS_NS && DIM == 3 && MRH_NS_HALF_CHUNKS) ? 2 : TA, NCH = P / LCH;
  static_assert(TA % LCH == 0, "a lane chunk must not straddle two variables");
  using S = PAD<LCH>;
  using LAY = WarpSmem<C>;
  static constexpr int NT = LAY::NT, NRT = LAY::NRT, QS = LAY::QS, KS = LAY::KS;
  static_assert(N <= 32, "one lane per element row");

  // ---- geometry, item = (point, part): every part rebuilds J, det, J^-1; part 0 stores them.  Small bases (STORE_T):
  //      part p also stores the physical basis rows j == p (mod nparts), T[j][1+c] = sum_m J^-1[m][c] G[j][m].
  //      Larger bases never store T: the point phase transforms the reference gradient of u and the contraction
  //      builds its fragments from J^-1 and the reference tables (basis_entry). ----
  static constexpr bool STORE_T = LAY::STORE_T;
  MRH_HD static int geom_parts(int nq) { return !STORE_T ? 1 : nq * 4 <= 32 ? 4 : (nq * 2 <= 32 ? 2 : 1); }
  MRH_HDN static void geometry(const KParams& kp, const LAY& L, double* sm, int item, int nparts) {
    const int q = item / nparts, part = item - q * nparts;
    const double* X = sm + L.oX();
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
    if (part == 0 && is_side(kp)) {
      // side launch: weight = |J t_U x J t_V| w (3-D) or |J t| w (2-D), unit outward normal (discretizationInterface.cpp:1681-1787)
      double nrm[3] = {0.0, 0.0, 0.0}, tng[3] = {0.0, 0.0, 0.0}, measure;
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
      double nl = 0.0; for (int c = 0; c < DIM; ++c) nl += nrm[c] * nrm[c];
      nl = 1.0 / sqrt(nl);
      sm[L.oW() + q] = measure * tb.W[q];
      for (int c = 0; c < 3; ++c) { sm[L.oXq() + (size_t)q * 3 + c] = xq[c]; sm[L.oNrm() + (size_t)q * 6 + c] = nrm[c] * nl; sm[L.oNrm() + (size_t)q * 6 + 3 + c] = tng[c]; }
      for (int m = 0; m < DIM; ++m) for (int c = 0; c < DIM; ++c) sm[L.oJi() + (size_t)q * DIM * DIM + m * DIM + c] = Ji[m][c];
    } else if (part == 0) {
      sm[L.oW() + q] = fabs(det) * tb.W[q];
      for (int c = 0; c < 3; ++c) sm[L.oXq() + (size_t)q * 3 + c] = xq[c];
      for (int m = 0; m < DIM; ++m) for (int c = 0; c < DIM; ++c) sm[L.oJi() + (size_t)q * DIM * DIM + m * DIM + c] = Ji[m][c];
    }
    if (STORE_T) {
      double* T = sm + L.oT() + (size_t)q * NB * TA;
      for (int j = part; j < NB; j += nparts) {
        T[j * TA] = tb.B[q * NB + j];
        for (int c = 0; c < DIM; ++c) {
          double s = 0.0;
          for (int m = 0; m < DIM; ++m) s += Ji[m][c] * tb.G[(q * NB + j) * DIM + m];
          T[j * TA + 1 + c] = s;
        }
      }
    }
  }
  // physical basis entry T[j][a] at point q (a = 0: value, a = 1 + c: d/dx_c)
  MRH_HD static double basis_entry(const RefTab& tb, const double* Ji, int q, int j, int a) {
    if (a == 0) return tb.B[q * NB + j];
    double s = 0.0;
#pragma unroll
    for (int m = 0; m < DIM; ++m) s += Ji[m * DIM + (a - 1)] * tb.G[(q * NB + j) * DIM + m];
    return s;
  }

  // ---- prep: state-independent programs, item = (point, i-th program of kind 1); lane 0 adds the element size ----
  MRH_HD static int num_kind1(const KParams& kp) { int n = 0; for (int k = 0; k < MRH_NCOEF; ++k) n += (kp.progs[k].kind == 1); return n; }
  MRH_HDN static void prep(const KParams& kp, const LAY& L, double* sm, int item, int n1) {
    const int q = item / n1; int idx = item - q * n1, k = 0;
    for (; k < MRH_NCOEF; ++k) if (kp.progs[k].kind == 1 && idx-- == 0) break;
    PointState<double> pd;
    for (int c = 0; c < 3; ++c) {
      pd.x[c] = sm[L.oXq() + (size_t)q * 3 + c];
      pd.nrm[c] = is_side(kp) ? sm[L.oNrm() + (size_t)q * 6 + c] : 0.0; pd.tng[c] = is_side(kp) ? sm[L.oNrm() + (size_t)q * 6 + 3 + c] : 0.0;
    }
    pd.time = kp.tp.time; pd.fields = nullptr;
    sm[L.oCoef() + (size_t)q * MRH_NCOEF + k] = eval_code<double>(kp.code, kp.cpool, kp.progs[k].start, kp.progs[k].len, pd);
  }
  MRH_HDN static void element_size(const LAY& L, double* sm, bool side = false) {
    double vol = 0.0;
    for (int k = 0; k < L.nq; ++k) vol += sm[L.oW() + k];
    if (side) { sm[L.oH()] = (DIM == 3) ? sqrt(vol) : vol; return; }          // side element size (sum w_side)^(1/(d-1)), workset.cpp:2681-2697
    sm[L.oH()] = (DIM == 3) ? cbrt(vol) : (DIM == 2) ? sqrt(vol) : vol;      // vol^(1/dim): within an ulp of the reference's pow(), a fifth of its instructions
  }

  // ---- point physics on one lane chunk, item = (point, variable m) ----------------------------------------
  MRH_HDN static void point_chunk(const KParams& kp, const LAY& L, double* sm, int item) {
    const int q = item / NCH, ch = item - q * NCH;
    if (!kp.compute_jac && ch != 0) return;                     // residual-only pass: the values come from chunk 0 alone
    const int lane0 = ch * LCH;
    const double* Ji = sm + L.oJi() + (size_t)q * DIM * DIM;
    const RefTab& tb = kp.tab;
    const double* U = sm + L.oU();
    const double* Ud = sm + L.oUd();
    const bool tr = kp.tp.transient != 0;
    const double au = tr ? kp.tp.alpha_u : 1.0;
    PointBat<S, NVAR, DIM> in;
#pragma unroll
    for (int m = 0; m < NVAR; ++m) {
      double fv[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
      if (STORE_T) {
        const double* T = sm + L.oT() + (size_t)q * NB * TA;
        for (int j = 0; j < NB; ++j) {
          const int slot = kp.offsets[m * NB + j];
          const double uj = U[slot], b = T[j * TA];
          fv[0] += uj * b;
#pragma unroll
          for (int c = 0; c < DIM; ++c) fv[1 + c] += uj * T[j * TA + 1 + c];
          if (tr && !is_side(kp)) fv[4] += Ud[slot] * b;
        }
      } else {
        double gr[DIM];
#pragma unroll
        for (int c = 0; c < DIM; ++c) gr[c] = 0.0;
        for (int j = 0; j < NB; ++j) {
          const int slot = kp.offsets[m * NB + j];
          const double uj = U[slot], b = tb.B[q * NB + j];
          fv[0] += uj * b;
#pragma unroll
          for (int c = 0; c < DIM; ++c) gr[c] += uj * tb.G[(q * NB + j) * DIM + c];
          if (tr && !is_side(kp)) fv[4] += Ud[slot] * b;
        }
#pragma unroll
        for (int c = 0; c < DIM; ++c) {
          double g = 0.0;
#pragma unroll
          for (int mm = 0; mm < DIM; ++mm) g += Ji[mm * DIM + c] * gr[mm];
          fv[1 + c] = g;
        }
      }
#pragma unroll
      for (int k = 0; k < 5; ++k) in.f[m][k] = S(fv[k]);
#pragma unroll
      for (int k = 0; k <= DIM; ++k)
#pragma unroll
        for (int l = 0; l < LCH; ++l) if (m * TA + k == lane0 + l) in.f[m][k].d[l] = au;
      if (tr && !is_side(kp)) {
#pragma unroll
        for (int l = 0; l < LCH; ++l) if (m * TA == lane0 + l) in.f[m][4].d[l] = kp.tp.alpha_t;
      }
    }
    bool sd = false;
    for (int k = 0; k < MRH_NCOEF; ++k) sd = sd || (kp.progs[k].kind == 2);
    S fm[NVAR * FIELD_KINDS];
    if (sd) {
#pragma unroll
      for (int m = 0; m < NVAR; ++m)
#pragma unroll
        for (int k = 0; k < FIELD_KINDS; ++k) fm[m * FIELD_KINDS + k] = (k < 5) ? in.f[m][k] : S(0.0);
    }
    in.kp = &kp; in.coef = sm + L.oCoef() + (size_t)q * MRH_NCOEF;
    in.ps.fields = fm; in.ps.time = kp.tp.time;
    for (int c = 0; c < 3; ++c) {
      in.ps.x[c] = sm[L.oXq() + (size_t)q * 3 + c];
      in.ps.nrm[c] = is_side(kp) ? sm[L.oNrm() + (size_t)q * 6 + c] : 0.0; in.ps.tng[c] = is_side(kp) ? sm[L.oNrm() + (size_t)q * 6 + 3 + c] : 0.0;
    }
    in.w = sm[L.oW() + q];
    in.h = (NeedsElementSize<C::PHYS>::value || is_side(kp)) ? sm[L.oH()] : 0.0;
    S F[NVAR][TA];
#pragma unroll
    for (int v = 0; v < NVAR; ++v)
#pragma unroll
      for (int a = 0; a < TA; ++a) F[v][a] = S(0.0);
    if (!is_side(kp)) Physics<C::PHYS, DIM, S>::volume(in, F);
    else {
      Physics<C::PHYS, DIM, S>::boundary(in, F);
#pragma unroll
      for (int v = 0; v < NVAR; ++v)          // generic "Flux" condition, physicsInterface.cpp:1698-1766
        if (kp.ph.bc[v] == BC_FLUX) F[v][0] = F[v][0] + (-in.fn(SIDE_FLUX + v)) * in.w;
    }
    // f -> F[q][v][a];  dF_(v,a)/d(m,b) -> B-fragment order: tile (v, t = m/2), lane = a + 4 ((m & 1) + 2 b)
    double* sF = sm + L.oF() + (size_t)q * NVAR * TA;
    double* Dq = sm + L.oD() + LAY::dq(q);
    const int mv = lane0 / TA, b0 = lane0 - mv * TA;              // variable and first slot this chunk differentiates by
    const int t = mv >> 1, half = mv & 1;
    const bool phantom = (NVAR & 1) && mv == NVAR - 1;           // odd variable count: this item also zeroes column m + 1
    const bool last = b0 + LCH == TA;                            // 2-D (3 slots): the last chunk also zeroes slot 3
#pragma unroll
    for (int v = 0; v < NVAR; ++v) {
      double* Dt = Dq + (size_t)(v * NT + t) * 32;
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        if (a < TA && ch == 0) sF[v * TA + a] = F[v][a < TA ? a : 0].v;
        if (!kp.compute_jac) continue;
#pragma unroll
        for (int l = 0; l < LCH; ++l) {
          const int b = b0 + l;
          Dt[a + 4 * (half + 2 * b)] = (a < TA) ? F[v][a < TA ? a : 0].d[l] : 0.0;
          if (phantom) Dt[a + 4 * (1 + 2 * b)] = 0.0;
        }
        if (TA < 4 && last) {
#pragma unroll
          for (int b = TA; b < 4; ++b) { Dt[a + 4 * (half + 2 * b)] = 0.0; if (phantom) Dt[a + 4 * (1 + 2 * b)] = 0.0; }
        }
      }
    }
  }

  // ---- contraction on the FP64 tensor path (warp-collective) ------------------------------------------------
  // step 1 (per v, column tile t):  L = T_rt (8 x 4: basis i, slot a) x D_(v,t) (4 x 8: slot a, columns (b, m & 1))
  // step 2 (per m of the tile)   :  K_(v,rt),(m,ct) += L_m (8 x 4: basis i, slot b) x T_ct^T (4 x 8: slot b, basis j)
  // The A fragment of step 2 is the C fragment of step 1; the B fragment of step 2 is the A fragment of step 1.
  struct Acc { WV k[NVAR][NRT][NVAR][NRT][2]; WV r[NRT][2]; };
  MRH_HD static void contract(const KParams& kp, const LAY& L, const double* sm, Acc& A) {
    const RefTab& tb = kp.tab;
#pragma unroll
    for (int v = 0; v < NVAR; ++v)
#pragma unroll
      for (int rt = 0; rt < NRT; ++rt)
#pragma unroll
        for (int m = 0; m < NVAR; ++m)
#pragma unroll
          for (int ct = 0; ct < NRT; ++ct) { A.k[v][rt][m][ct][0] = wv_zero(); A.k[v][rt][m][ct][1] = wv_zero(); }
#pragma unroll
    for (int rt = 0; rt < NRT; ++rt) { A.r[rt][0] = wv_zero(); A.r[rt][1] = wv_zero(); }
    // basis fragments of point q: loaded (Q1) or rebuilt from J^-1 and the reference tables (larger bases) one point
    // ahead, so that their loads overlap the DMMAs of the current point
    // Bases that do not store T (Q2): the fragment is itself a tensor product, T_q (8 x 4) = (B | G)_q (8 x 4, reference table
    // in fragment order, RefTab::FR) x M_q (4 x 4) with M = diag(1, J^-1).  The B operand carries column a of M in column 2 a
    // (odd columns zero), so that the FIRST accumulator register of lane 4 j + a is T[j][a]: the product's C layout is the A
    // layout the contraction needs - one coalesced load and one DMMA per fragment, no shuffle, against three gathers and three
    // predicated FMAs per lane before (profiles/r1_batched_final_C3: 2/3 of the instructions of C3 were fragment building)
    auto m_frag = [&](int q) {
      const double* Ji = sm + L.oJi() + (size_t)q * DIM * DIM;
      return wv_make([&](int lane) {                   // b[k = lane % 4][n = lane / 4] = (n even) ? M[k][n / 2] : 0
        const int k = lane & 3, n = lane >> 2, a = n >> 1;
        if (n & 1) return 0.0;
        if (a == 0) return k == 0 ? 1.0 : 0.0;
        return (k >= 1 && k <= DIM && a <= DIM) ? Ji[(k - 1) * DIM + (a - 1)] : 0.0; });
    };
    auto basis_frag = [&](int q, int rt, const WV& mf) {
      if constexpr (STORE_T) {
        return wv_make([&](int lane) { const int j = rt * 8 + (lane >> 2), a = lane & 3; return (j < NB && a < TA) ? sm[L.oT() + (size_t)q * NB * TA + j * TA + a] : 0.0; });
#if defined(MRH_BAT_FRAG_GATHER)                // previous form, kept for A/B timing (MRH_BAT_DEFINE=MRH_BAT_FRAG_GATHER)
      } else if constexpr (true) {
        const double* Ji = sm + L.oJi() + (size_t)q * DIM * DIM;
        return wv_make([&](int lane) { const int j = rt * 8 + (lane >> 2), a = lane & 3; return (j < NB && a < TA) ? basis_entry(tb, Ji, q, j, a) : 0.0; });
#endif
      } else {
        const double* fr = tb.FR + ((size_t)q * NRT + rt) * 32;
        const WV ref = wv_make([&](int lane) { return fr[lane]; });
        WV c0 = wv_zero(), c1 = wv_zero();
        wv_mma(c0, c1, ref, mf);
        return c0;
      }
    };
    WV mfn = STORE_T ? wv_zero() : m_frag(0);
    WV tfn[NRT];
#pragma unroll
    for (int rt = 0; rt < NRT; ++rt) tfn[rt] = basis_frag(0, rt, mfn);
#if defined(MRH_BAT_QUNROLL)                   // tuning experiment: unrolled point loop (the one-ahead fragments then need no register copies)
#pragma unroll MRH_BAT_QUNROLL
#endif
    for (int q = 0; q < L.nq; ++q) {
      const double* sF = sm + L.oF() + (size_t)q * NVAR * TA;
      const double* Dq = sm + L.oD() + LAY::dq(q);
      WV tf[NRT];
#pragma unroll
      for (int rt = 0; rt < NRT; ++rt) tf[rt] = tfn[rt];
      if (q + 1 < L.nq) {
        if (!STORE_T) mfn = m_frag(q + 1);
#pragma unroll
        for (int rt = 0; rt < NRT; ++rt) tfn[rt] = basis_frag(q + 1, rt, mfn);
      }
      const WV ff = wv_make([&](int lane) { const int v = lane >> 2, a = lane & 3; return (v < NVAR && a < TA) ? sF[v * TA + a] : 0.0; });
#pragma unroll
      for (int rt = 0; rt < NRT; ++rt) wv_mma(A.r[rt][0], A.r[rt][1], tf[rt], ff);
      if (!kp.compute_jac) continue;
#pragma unroll
      for (int v = 0; v < NVAR; ++v) {
        // all first products of this row variable are issued before the second products that consume them, so a
        // dependent DMMA never follows its producer back to back
        WV l0[NT][NRT], l1[NT][NRT];
#pragma unroll
        for (int t = 0; t < NT; ++t) {
          const double* Dt = Dq + (size_t)(v * NT + t) * 32;
          const WV df = wv_make([&](int lane) { return Dt[lane]; });
#pragma unroll
          for (int rt = 0; rt < NRT; ++rt) { l0[t][rt] = wv_zero(); l1[t][rt] = wv_zero(); wv_mma(l0[t][rt], l1[t][rt], tf[rt], df); }
        }
#pragma unroll
        for (int t = 0; t < NT; ++t)
#pragma unroll
          for (int rt = 0; rt < NRT; ++rt)
#pragma unroll
            for (int ct = 0; ct < NRT; ++ct) {
              wv_mma(A.k[v][rt][2 * t][ct][0], A.k[v][rt][2 * t][ct][1], l0[t][rt], tf[ct]);
              if (2 * t + 1 < NVAR) wv_mma(A.k[v][rt][(2 * t + 1 < NVAR) ? 2 * t + 1 : 0][ct][0], A.k[v][rt][(2 * t + 1 < NVAR) ? 2 * t + 1 : 0][ct][1], l1[t][rt], tf[ct]);
            }
      }
    }
  }

  // ---- park K (natural order: row (v, i), column (m, j)) and r in shared memory; lane = fragment lane ------
  MRH_HD static void store_acc(const LAY& L, double* sm, const Acc& A, bool jac) {
    double* K = sm + L.oK();
    double* R = sm + L.oR();
    wv_each([&](int lane) {
      const int i0 = lane >> 2, p = lane & 3;
#pragma unroll
      for (int rt = 0; rt < NRT; ++rt) {
        const int i = rt * 8 + i0;
        if (i >= NB) continue;
#pragma unroll
        for (int e = 0; e < 2; ++e) { const int v = 2 * p + e; if (v < NVAR) R[v * NB + i] = wv_get(A.r[rt][e], lane); }
        if (!jac) continue;
#pragma unroll
        for (int v = 0; v < NVAR; ++v)
#pragma unroll
          for (int m = 0; m < NVAR; ++m)
#pragma unroll
            for (int ct = 0; ct < NRT; ++ct) {
              const int j = ct * 8 + 2 * p;
              double* dst = K + (size_t)(v * NB + i) * KS + m * NB + j;
              const double k0 = wv_get(A.k[v][rt][m][ct][0], lane), k1 = wv_get(A.k[v][rt][m][ct][1], lane);
              if ((NB & 1) == 0) { if (j < NB) st2(dst, k0, k1); }
              else { if (j < NB) dst[0] = k0; if (j + 1 < NB) dst[1] = k1; }
            }
      }
    });
  }

  // ---- per-row data: CRS row base (or -1 for isFixedDOF rows), lumping; lane = natural row ----------------
  MRH_HDN static void row_setup(const KParams& kp, const LAY& L, double* sm, const int* offs, int lane) {
    if (lane >= N) return;
    const int* sLid = reinterpret_cast<const int*>(sm + L.oLid());
    const int rowslot = offs[lane], row = sLid[rowslot];
    reinterpret_cast<int64_t*>(sm + L.oRowBase())[lane] = kp.fixed[row] ? (int64_t)-1 : kp.rowptr[row];
    reinterpret_cast<int*>(sm + L.oPatOff())[lane] = rowslot * N;
    if (kp.lump_mass && kp.compute_jac) {                            // assemblyManager.cpp:4130-4132
      double* K = sm + L.oK() + (size_t)lane * KS;
      double s = 0.0;
      for (int c = 0; c < N; ++c) { s += K[c]; K[c] = 0.0; }
      K[lane] = s;
    }
  }

  // ---- scatter: rows one after the other, lanes along (node j, variable m) = consecutive element slots -----
  MRH_HDN static void scatter(const KParams& kp, const LAY& L, const double* sm, const int* offs, int64_t e, int lane) {
    const int64_t* rb = reinterpret_cast<const int64_t*>(sm + L.oRowBase());
    const int* sLid = reinterpret_cast<const int*>(sm + L.oLid());
    const double* K = sm + L.oK();
    if (lane < N && rb[lane] >= 0) atomic_add(kp.res + sLid[offs[lane]], -sm[L.oR() + lane]);
    if (lane >= N || !kp.compute_jac) return;
    const int col = (lane % NVAR) * NB + lane / NVAR;                // natural column this lane serves
    const int colslot = offs[col];
    const uint16_t* pat = kp.pat + (size_t)kp.pat_id[e] * N * N + colslot;
    const int* po = reinterpret_cast<const int*>(sm + L.oPatOff());
    const double* Kc = K + col;
    constexpr int UR = 4;                                            // rows in flight: the slot loads of UR rows overlap
#pragma unroll
    for (int r0 = 0; r0 < N; r0 += UR) {
      int64_t base[UR]; uint16_t slot[UR];
#pragma unroll
      for (int k = 0; k < UR; ++k) {
        const int r = r0 + k;
        base[k] = r < N ? rb[r] : (int64_t)-1;
        slot[k] = base[k] >= 0 ? pat[po[r < N ? r : 0]] : (uint16_t)0;
      }
#pragma unroll
      for (int k = 0; k < UR; ++k)
        if (base[k] >= 0) atomic_add(kp.jac + base[k] + slot[k], Kc[(r0 + k) * KS]);
    }
  }
};

#if defined(__CUDACC__)
// Coefficient-program policy of the kernel body: the bytecode interpreter (ahead-of-time build, batched_kernels.cu)
// or straight-line functions printed from the bytecode and compiled with NVRTC at mrh_finalize (jit.hpp: JitProgs).
struct InterpProgs {
  static constexpr bool jit = false;
  static constexpr int NA = 0;
  static __device__ __forceinline__ double atom_arg(int, double, double, double, double) { return 0.0; }
  static __device__ __forceinline__ int atom_shift(int) { return 0; }
  static __device__ __forceinline__ double eval(int, double, double, double, double, const double*) { return 0.0; }
};

// One warp per element, grid-stride over the elements of the launch.  smem_all: [offsets | one WarpSmem block per warp].
template <class C, class PROGS, bool SIDES = true>
__device__ __forceinline__ void batched_body(const KParams& kp, double* smem_all) {
  using G = GenPhases<C>;
  using B = BatPhases<C, SIDES>;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const WarpSmem<C> L(kp.tab.nq, kp.tp.transient != 0);
  int* offs = reinterpret_cast<int*>(smem_all);                        // offsets(var, basis) -> element slot, CTA-shared
  for (int k = threadIdx.x; k < C::N; k += blockDim.x) offs[k] = kp.offsets[k];
  __syncthreads();
  double* sm = smem_all + even_up((C::N + 1) / 2) + (size_t)warp * L.total_doubles();
  const int nq = L.nq, nparts = B::geom_parts(nq), n1 = B::num_kind1(kp);
  // Steady-state gather, software-pipelined two elements deep: the element -> DOF / vertex indices of element w + 2 stride
  // and the solution / coordinate values of element w + stride are in flight while element w is processed
  // (lane k < N owns DOF slot k, lane a < NV owns vertex a).  Transient launches use the plain gather (stage / BDF sums).
  const bool pipelined = !kp.tp.transient;
  const int64_t w_first = (int64_t)blockIdx.x * nwarp + warp, w_stride = (int64_t)gridDim.x * nwarp;
  // (plain statements, no lambdas: variables captured by reference were placed in local memory, and storing a value
  //  that is still in flight there stalls the warp on the load - exactly what the pipeline is meant to avoid)
  int lidA = 0, ndA = 0, lidB = 0, ndB = 0;
  double uA = 0.0, xA0 = 0.0, xA1 = 0.0, xA2 = 0.0;
#define MRH_LOAD_IDX(W, LID, ND) { const int64_t e_ = G::elem_of(kp, (W)); \
    if (lane < C::N) LID = kp.lids[e_ * C::N + lane]; if (lane < C::NV) ND = kp.e2n[e_ * C::NV + lane]; }
#define MRH_LOAD_VAL(LID, ND) { if (lane < C::N) uA = kp.u[LID]; \
    if (lane < C::NV) { const double* xp_ = kp.nodes + (int64_t)(ND) * C::DIM; xA0 = xp_[0]; xA1 = xp_[1]; if (C::DIM == 3) xA2 = xp_[C::DIM - 1]; } }
  if (pipelined && w_first < kp.nwork) {
    MRH_LOAD_IDX(w_first, lidA, ndA)
    if (w_first + w_stride < kp.nwork) MRH_LOAD_IDX(w_first + w_stride, lidB, ndB)
    MRH_LOAD_VAL(lidA, ndA)
  }
  for (int64_t w = w_first; w < kp.nwork; w += w_stride) {
    if (pipelined) {
      if (lane < C::N) { reinterpret_cast<int*>(sm + L.oLid())[lane] = lidA; sm[L.oU() + lane] = uA; }
      if (lane < C::NV) {
        double* xs_ = sm + L.oX() + lane * C::DIM;
        xs_[0] = xA0; xs_[1] = xA1; if (C::DIM == 3) xs_[C::DIM - 1] = xA2;
      }
      if (w + w_stride < kp.nwork) { lidA = lidB; ndA = ndB; MRH_LOAD_VAL(lidA, ndA) }
      if (w + 2 * w_stride < kp.nwork) MRH_LOAD_IDX(w + 2 * w_stride, lidB, ndB)
    } else {
      for (int it = lane; it < C::N; it += 32) G::gather_dof(kp, L, sm, w, it);
      for (int it = lane; it < C::NV; it += 32) G::gather_vertex(kp, L, sm, w, it);
    }
    __syncwarp();
    for (int it = lane; it < nq * nparts; it += 32) B::geometry(kp, L, sm, it, nparts);
    __syncwarp();
    if (PROGS::jit) {
      // sin / cos atoms of all points first: every lane busy, uniform code.  They live at the start of the D region, which the
      // point phase only writes after the programs have read them (jit.hpp checks NA against the region's size).
      double* at = sm + L.oD();
      if (PROGS::NA > 0) {
        for (int it = lane; it < nq * PROGS::NA; it += 32) {
          const int q = it / PROGS::NA, a = it - q * PROGS::NA;
          const double arg = PROGS::atom_arg(a, sm[L.oXq() + (size_t)q * 3], sm[L.oXq() + (size_t)q * 3 + 1], sm[L.oXq() + (size_t)q * 3 + 2], kp.tp.time);
          const int sh = PROGS::atom_shift(a);
          bool ok = true;
          double v = trig_fast(arg, sh, ok);
          if (!ok) v = sh ? cos(arg) : sin(arg);
          at[it] = v;
        }
        __syncwarp();
      }
      for (int q = lane; q < nq; q += 32) {
        const double x = sm[L.oXq() + (size_t)q * 3], y = sm[L.oXq() + (size_t)q * 3 + 1], z = sm[L.oXq() + (size_t)q * 3 + 2];
        // the point's atoms go to registers once and every program is evaluated before the first store: with the stores in between
        // the compiler had to assume they alias the atom table and re-read the atoms for every term of every program
        if constexpr (PROGS::NA > 0) {
          double areg[PROGS::NA > 0 ? PROGS::NA : 1];
#pragma unroll
          for (int i = 0; i < PROGS::NA; ++i) areg[i] = at[q * PROGS::NA + i];
          double val[MRH_NCOEF];
#pragma unroll
          for (int k = 0; k < MRH_NCOEF; ++k) val[k] = PROGS::eval(k, x, y, z, kp.tp.time, areg);
#pragma unroll
          for (int k = 0; k < MRH_NCOEF; ++k)
            if (kp.progs[k].kind == 1) sm[L.oCoef() + (size_t)q * MRH_NCOEF + k] = val[k];
        } else {
#pragma unroll
          for (int k = 0; k < MRH_NCOEF; ++k)
            if (kp.progs[k].kind == 1) sm[L.oCoef() + (size_t)q * MRH_NCOEF + k] = PROGS::eval(k, x, y, z, kp.tp.time, nullptr);
        }
      }
    } else {
      for (int it = lane; it < nq * n1; it += 32) B::prep(kp, L, sm, it, n1);
    }
    if (lane == 31 && (NeedsElementSize<C::PHYS>::value || B::is_side(kp))) B::element_size(L, sm, B::is_side(kp));       // h enters volume terms through the stabilisation only
    __syncwarp();
    for (int it = lane; it < nq * B::NCH; it += 32) B::point_chunk(kp, L, sm, it);
    __syncwarp();
    typename B::Acc acc;
    B::contract(kp, L, sm, acc);
    __syncwarp();                                          // every lane is done with D: K may take its place
    B::store_acc(L, sm, acc, kp.compute_jac != 0);
    __syncwarp();
    B::row_setup(kp, L, sm, offs, lane);
    __syncwarp();
    B::scatter(kp, L, sm, offs, G::elem_of(kp, w), lane);
    __syncwarp();                                          // the next element reuses the block
  }
#undef MRH_LOAD_IDX
#undef MRH_LOAD_VAL
}
#endif

#if !defined(__CUDACC_RTC__)
constexpr int MRH_BATCHED_MAXW = 6;       // most warps per CTA the launchers try
// dynamic shared memory of a CTA of nw warps (one element block per warp + the CTA-shared offsets table)
template <class C>
inline size_t batched_smem(int nq, bool transient, int nw) {
  return even_up((C::N + 1) / 2) * sizeof(double) + (size_t)nw * WarpSmem<C>(nq, transient).total_doubles() * sizeof(double);
}
// warps per CTA that put the most warps on one SM; occ(threads, smem) = resident CTAs per SM of the kernel at hand
// (registers included, which is why it is asked of the driver instead of computed from the layout).  0: nothing fits.
template <class C, class OCC>
inline int batched_pick_warps(int nq, bool transient, OCC occ, int& blocks_per_sm) {
  int best = 0, nw = 0; blocks_per_sm = 0;
  for (int c = 1; c <= MRH_BATCHED_MAXW; ++c) {
    const size_t smem = batched_smem<C>(nq, transient, c);
    if (smem > (size_t)227 * 1024) break;
    const int b = occ(32 * c, smem);
    if (b * c > best) { best = b * c; nw = c; blocks_per_sm = b; }
  }
  return nw;
}
#endif

}  // namespace mrh
