// Specialised volume kernel for the headline configuration: thermal, HGRAD Q1 hexahedra, 2x2x2 Gauss rule,
// coefficients that do not depend on the solution (BASELINE.json configs[1]).  One thread owns one element and
// keeps everything — vertex coordinates, gathered DOFs, J / J^-1, the 8 physical gradients, the 8 residual
// entries and the 8x8 element matrix — in registers; nothing goes through shared or global scratch.
// It fuses what the reference runs as separate Kokkos launches per 100-element group:
//   "assembly gather"                    /root/reference/src/managers/assemblyManager.cpp:3633-3641
//   "wkset reset res" / seeding          /root/reference/src/tools/workset.cpp:430-459, :589-623, :836-857
//   setJacobian/Inv/Det, HGRADtransformGRAD, computeCellMeasure
//                                        /root/reference/src/interfaces/discretizationInterface.cpp:771-774, :923-971
//   "wkset soln ip HGRAD" (e, grad e, e_t) workset.cpp:1044-1056
//   "funcman evaluate <op>" (one launch per DAG node)   /root/reference/src/managers/functionManager.cpp:744-1900
//   "Thermal volume resid 3D part 1"     /root/reference/src/physics/thermal.cpp:125-163
//   "assembly insert Jac"                assemblyManager.cpp:4063-4144
// Derivative lanes: with state-independent coefficients d r_i / d u_j = sum_q [rho cp alpha_t phi_i phi_j
//   + kappa alpha_u grad phi_i . grad phi_j + alpha_u (b . grad phi_j) phi_i] w, written out explicitly; without
// advection the element matrix is symmetric and only its 36 upper entries are accumulated.
// For Q1 the field basis equals the vertex (geometry) basis, so everything is indexed by element slot
// (= shards vertex) and the 2-point Gauss tables are compile-time constants of the unrolled point loop.
#pragma once
#include "assemble.h"

namespace mrh {

struct HexQ1 {
  static constexpr double g = 0.57735026918962576451;   // 1/sqrt(3); 2-point Gauss, weights 1
  // shards vertex s -> tensor bits
  MRH_HD static constexpr int bx(int s) { return (s == 1 || s == 2 || s == 5 || s == 6) ? 1 : 0; }
  MRH_HD static constexpr int by(int s) { return (s == 2 || s == 3 || s == 6 || s == 7) ? 1 : 0; }
  MRH_HD static constexpr int bz(int s) { return s >> 2; }
  // point q -> abscissa per direction (descending order: bit 0 -> +g), x fastest
  MRH_HD static constexpr double xi(int qbit) { return qbit ? -g : g; }
  MRH_HD static constexpr double l(int bit, int qbit) { return bit ? 0.5 * (1.0 + xi(qbit)) : 0.5 * (1.0 - xi(qbit)); }
  MRH_HD static constexpr double dl(int bit) { return bit ? 0.5 : -0.5; }
  MRH_HD static constexpr double N(int s, int q) { return l(bx(s), q & 1) * l(by(s), (q >> 1) & 1) * l(bz(s), q >> 2); }
  MRH_HD static constexpr double dN(int s, int q, int c) {
    return (c == 0) ? dl(bx(s)) * l(by(s), (q >> 1) & 1) * l(bz(s), q >> 2)
         : (c == 1) ? l(bx(s), q & 1) * dl(by(s)) * l(bz(s), q >> 2)
                    : l(bx(s), q & 1) * l(by(s), (q >> 1) & 1) * dl(bz(s));
  }
  MRH_HD static constexpr int tri(int i, int j) { return i * 8 - (i * (i - 1)) / 2 + (j - i); }   // i <= j, 36 entries
};

#if !defined(__CUDACC_RTC__)
// eligibility of a volume launch for this kernel (checked on the host)
inline bool thermal_hex_q1_eligible(int phys, int dim, int order, int nq, bool state_independent, int lump_mass) {
  return phys == PHYS_THERMAL && dim == 3 && order == 1 && nq == 8 && state_independent && !lump_mass;
}
#endif

// Coefficient policy.  InterpCoefs: the bytecode programs are run by the vector interpreter before the point loop
// and their values staged in shared memory.  The NVRTC build (jit.hpp) supplies a policy with staged = false whose
// source()/diff()/rhocp()/bx()/by()/bz() are the expressions themselves, compiled inline into the point loop.
struct InterpCoefs { static constexpr bool staged = true; };

template <bool TRANSIENT, bool ADV, bool JAC, class CF = InterpCoefs>
struct ThermalHexQ1 {
  static constexpr int NK = ADV ? 64 : 36;
  static constexpr int NSTAGE = 2 + (TRANSIENT ? 1 : 0) + (ADV ? 3 : 0);   // staged coefficient arrays (8 points each)
  MRH_HD static constexpr int vtx(int a, int b, int c) {                   // tensor bits -> shards vertex
    return (b ? (a ? 2 : 3) : (a ? 1 : 0)) + 4 * c;
  }

  // stage: per-thread scratch for the coefficient values at the 8 points, element k of array m at
  // stage[(m*8+k)*stride] (shared memory, stride = blockDim.x, on the device).
  MRH_HDN static void element(const KParams& kp, int64_t e, double* stage, int stride) {
#if defined(MRH_DEBUG_KSMEM_COMPUTE) && defined(__CUDA_ARCH__)     // tuning experiment: element computation only, K accumulated in shared memory
    {
      __shared__ double kbuf[NK * 128];
      int lid[8]; double r[8], X[8][3], us[8], ud[8];
      gather(kp, kp.lids + e * 8, kp.e2n + e * 8, lid, X, us, ud);
      KSmem acc{kbuf + threadIdx.x};
      work(kp, X, us, ud, stage, stride, r, acc);
      kp.res[lid[0]] = r[0] + acc.get(0) + acc.get(NK - 1);
      return;
    }
#endif
    int lid[8]; double r[8], K[NK];
    compute(kp, e, stage, stride, lid, r, K);
    scatter(kp, e, lid, r, K);
  }

  // gather + geometry + point loop of one element: r = element residual, K = element matrix (36 upper entries
  // when symmetric, else 8x8 row-major)
  MRH_HDN static void compute(const KParams& kp, int64_t e, double* stage, int stride, int (&lid)[8], double (&r)[8], double (&K)[NK]) {
    double X[8][3], us[8], ud[8];
    gather(kp, kp.lids + e * 8, kp.e2n + e * 8, lid, X, us, ud);
    work(kp, X, us, ud, stage, stride, r, K);
  }
  // ---- gather: LIDs, vertex coordinates, DOFs (time-stage combination for transient runs) ----------------
  // lrow / nrow: the element's 8 LIDs / 8 node ids (16-byte aligned; global or shared memory)
  MRH_HDN static void gather(const KParams& kp, const int* lrow, const int* nrow, int (&lid)[8], double (&X)[8][3], double (&us)[8], double (&ud)[8]) {
    {
      const int4* lp = reinterpret_cast<const int4*>(lrow);
      const int4 a = lp[0], b = lp[1];
      lid[0] = a.x; lid[1] = a.y; lid[2] = a.z; lid[3] = a.w; lid[4] = b.x; lid[5] = b.y; lid[6] = b.z; lid[7] = b.w;
    }
    {
      const int4* np = reinterpret_cast<const int4*>(nrow);
      const int4 a = np[0], b = np[1];
      const int nd[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        const double* p = kp.nodes + (int64_t)nd[s] * 3;
        X[s][0] = p[0]; X[s][1] = p[1]; X[s][2] = p[2];
      }
    }
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      const double uval = kp.u[lid[s]];
      if (!TRANSIENT) { us[s] = uval; ud[s] = 0.0; }
      else {
        const TimeParams& tp = kp.tp;
        const double prev0 = kp.u_prev[(size_t)lid[s] * tp.nsteps];
        double beta_u = (1.0 - tp.alpha_u) * prev0;
        for (int k = 0; k < tp.stage; ++k) beta_u += tp.a_over_b[k] * (kp.u_stage[(size_t)lid[s] * tp.nstages + k] - prev0);
        double beta_t = 0.0;
        for (int k = 1; k <= tp.nsteps; ++k) beta_t += tp.bdf[k] * kp.u_prev[(size_t)lid[s] * tp.nsteps + (k - 1)];
        us[s] = tp.alpha_u * uval + beta_u;
        ud[s] = tp.alpha_t * uval + beta_t * tp.timewt;
      }
    }
  }
  // ---- geometry + point loop of one element: r = element residual, K = element matrix -----------------------
  // element-matrix accumulators: registers, or one column of an entry-major shared-memory array (stride 128)
  struct KReg { static constexpr bool in_smem = false; double (&K)[NK]; MRH_HD double get(int k) const { return K[k]; } MRH_HD void set(int k, double v) { K[k] = v; } };
  struct KSmem { static constexpr bool in_smem = true; double* p; MRH_HD double get(int k) const { return p[k * 128]; } MRH_HD void set(int k, double v) { p[k * 128] = v; } };
  MRH_HDN static void work(const KParams& kp, double (&X)[8][3], double (&us)[8], const double (&ud)[8], double* stage, int stride,
                           double (&r)[8], double (&K)[NK]) {
    KReg acc{K};
    work(kp, X, us, ud, stage, stride, r, acc);
  }
  // X and us are overwritten by their monomial coefficients (see below)
  template <class KA>
  MRH_HDN static void work(const KParams& kp, double (&X)[8][3], double (&us)[8], const double (&ud)[8], double* stage, int stride,
                           double (&r)[8], KA& K) { work<KA, false>(kp, X, us, ud, stage, stride, r, K, nullptr); }
  // CSM: the 24 monomial coefficients of the trilinear map live in shared memory (cs[(m * 3 + i) * 128], one column per thread)
  // and are re-read at every point, which frees 48 registers - enough to keep the element matrix in registers instead
  template <class KA, bool CSM>
  MRH_HDN static void work(const KParams& kp, double (&X)[8][3], double (&us)[8], const double (&ud)[8], double* stage, int stride,
                           double (&r)[8], KA& K, double* cs) {
    const double au = TRANSIENT ? kp.tp.alpha_u : 1.0, at = TRANSIENT ? kp.tp.alpha_t : 0.0;

    // ---- coefficients at the 8 points (vector interpreter), staged for the rolled point loop ------------
    if constexpr (CF::staged) {
      double px[8], py[8], pz[8], val[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        double x = 0.0, y = 0.0, z = 0.0;
#pragma unroll
        for (int s = 0; s < 8; ++s) { x += X[s][0] * HexQ1::N(s, q); y += X[s][1] * HexQ1::N(s, q); z += X[s][2] * HexQ1::N(s, q); }
        px[q] = x; py[q] = y; pz[q] = z;
      }
      int m = 0;
      auto put = [&](int slot) {
        eval_prog_vec<8>(kp.progs[slot], kp.code, kp.cpool, px, py, pz, kp.tp.time, val);
#pragma unroll
        for (int q = 0; q < 8; ++q) stage[(m * 8 + q) * stride] = val[q];
        ++m;
      };
      put(TH_SOURCE); put(TH_DIFF);
      if (TRANSIENT) {
        double cp[8];
        eval_prog_vec<8>(kp.progs[TH_CP], kp.code, kp.cpool, px, py, pz, kp.tp.time, cp);
        eval_prog_vec<8>(kp.progs[TH_RHO], kp.code, kp.cpool, px, py, pz, kp.tp.time, val);
#pragma unroll
        for (int q = 0; q < 8; ++q) stage[(m * 8 + q) * stride] = val[q] * cp[q];
        ++m;
      }
      if (ADV) { put(TH_BX); put(TH_BY); put(TH_BZ); }
    }

    // ---- trilinear map and solution in monomial form --------------------------------------------------
    // x(xi, eta, zeta) = sum_m C_m xi^i eta^j zeta^k, m = i + 2 j + 4 k: a three-stage butterfly over the vertex sign
    // bits.  The Jacobian of the (general) trilinear map then costs 30 FP64 operations per point instead of 81
    // and the reference gradient of u 9 instead of 24.
    double (&C)[8][3] = X;
    double (&A)[8] = us;
    {
      double v[4][8];
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int b = 0; b < 2; ++b)
#pragma unroll
          for (int a = 0; a < 2; ++a) {
            const int sv = vtx(a, b, c), m = a + 2 * b + 4 * c;
            v[0][m] = X[sv][0]; v[1][m] = X[sv][1]; v[2][m] = X[sv][2]; v[3][m] = us[sv];
          }
#pragma unroll
      for (int d = 0; d < 4; ++d) {
#pragma unroll
        for (int bit = 1; bit < 8; bit <<= 1)
#pragma unroll
          for (int m = 0; m < 8; ++m)
            if (!(m & bit)) { const double lo = v[d][m], hi = v[d][m | bit]; v[d][m] = hi + lo; v[d][m | bit] = hi - lo; }
      }
#pragma unroll
      for (int m = 0; m < 8; ++m) { C[m][0] = 0.125 * v[0][m]; C[m][1] = 0.125 * v[1][m]; C[m][2] = 0.125 * v[2][m]; A[m] = 0.125 * v[3][m]; }
      if constexpr (CSM) {
#pragma unroll
        for (int m = 0; m < 8; ++m)
#pragma unroll
          for (int i = 0; i < 3; ++i) cs[(m * 3 + i) * 128] = C[m][i];
#pragma unroll
        for (int m = 1; m < 8; ++m) cs[(24 + m) * 128] = A[m];
      }
    }

    // ---- point loop (rolled: the body stays resident in the instruction cache) ------------------------
#pragma unroll
    for (int s = 0; s < 8; ++s) r[s] = 0.0;
#pragma unroll
    for (int k = 0; k < NK; ++k) K.set(k, 0.0);

#if defined(MRH_WS_QUNROLL)              // NVRTC build: two points per trip, so that the scheduler can interleave their dependent chains
#pragma unroll MRH_WS_QUNROLL
#else
#pragma unroll 1
#endif
    for (int q = 0; q < 8; ++q) {
      const double g = HexQ1::g;
      const double sx = (q & 1) ? -g : g, sy = (q & 2) ? -g : g, sz = (q & 4) ? -g : g;
      const double lx[2] = {0.5 - 0.5 * sx, 0.5 + 0.5 * sx}, ly[2] = {0.5 - 0.5 * sy, 0.5 + 0.5 * sy}, lz[2] = {0.5 - 0.5 * sz, 0.5 + 0.5 * sz};
      double yz[2][2], xz[2][2], xy[2][2];
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) { yz[a][b] = ly[a] * lz[b]; xz[a][b] = lx[a] * lz[b]; xy[a][b] = lx[a] * ly[b]; }
      // J_ij = d x_i / d xi_j of the trilinear map, from its monomial coefficients
      const double xe = sx * sy, ez2 = sy * sz, xz2 = sx * sz, xez = xe * sz;
      double J[3][3];
      double ptc[3] = {0.0, 0.0, 0.0};                   // CSM: the point's coordinates are formed here, while the coefficients are at hand
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        if constexpr (CSM) {
          const double c0 = cs[(0 * 3 + i) * 128], c1 = cs[(1 * 3 + i) * 128], c2 = cs[(2 * 3 + i) * 128], c3 = cs[(3 * 3 + i) * 128],
                       c4 = cs[(4 * 3 + i) * 128], c5 = cs[(5 * 3 + i) * 128], c6 = cs[(6 * 3 + i) * 128], c7 = cs[(7 * 3 + i) * 128];
          J[i][0] = c1 + c3 * sy + c5 * sz + c7 * ez2;
          J[i][1] = c2 + c3 * sx + c6 * sz + c7 * xz2;
          J[i][2] = c4 + c5 * sx + c6 * sy + c7 * xe;
          ptc[i] = c0 + c1 * sx + c2 * sy + c4 * sz + c3 * xe + c5 * xz2 + c6 * ez2 + c7 * xez;
        } else {
        J[i][0] = C[1][i] + C[3][i] * sy + C[5][i] * sz + C[7][i] * ez2;
        J[i][1] = C[2][i] + C[3][i] * sx + C[6][i] * sz + C[7][i] * xz2;
        J[i][2] = C[4][i] + C[5][i] * sx + C[6][i] * sy + C[7][i] * xe;
        }
      }
      double Ji[3][3];
      const double det = det_inv<3>(J, Ji);
      const double w = fabs(det);                       // reference weight 1
      double gx[8], gy[8], gz[8], Nq[8];
      double ex = 0.0, ey = 0.0, ez = 0.0, edot = 0.0;
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        const int a = HexQ1::bx(s), b = HexQ1::by(s), c = HexQ1::bz(s);
        const double d0 = (a ? 0.5 : -0.5) * yz[b][c], d1 = (b ? 0.5 : -0.5) * xz[a][c], d2 = (c ? 0.5 : -0.5) * xy[a][b];
        Nq[s] = xy[a][b] * lz[c];
        gx[s] = Ji[0][0] * d0 + Ji[1][0] * d1 + Ji[2][0] * d2;
        gy[s] = Ji[0][1] * d0 + Ji[1][1] * d1 + Ji[2][1] * d2;
        gz[s] = Ji[0][2] * d0 + Ji[1][2] * d1 + Ji[2][2] * d2;
        if (TRANSIENT) edot += ud[s] * Nq[s];
      }
      {
        double a1 = A[1], a2 = A[2], a3 = A[3], a4 = A[4], a5 = A[5], a6 = A[6], a7 = A[7];
        if constexpr (CSM) { a1 = cs[25 * 128]; a2 = cs[26 * 128]; a3 = cs[27 * 128]; a4 = cs[28 * 128]; a5 = cs[29 * 128]; a6 = cs[30 * 128]; a7 = cs[31 * 128]; }
        const double g0 = a1 + a3 * sy + a5 * sz + a7 * ez2;        // reference gradient of u
        const double g1 = a2 + a3 * sx + a6 * sz + a7 * xz2;
        const double g2 = a4 + a5 * sx + a6 * sy + a7 * xe;
        ex = Ji[0][0] * g0 + Ji[1][0] * g1 + Ji[2][0] * g2;
        ey = Ji[0][1] * g0 + Ji[1][1] * g1 + Ji[2][1] * g2;
        ez = Ji[0][2] * g0 + Ji[1][2] * g1 + Ji[2][2] * g2;
      }
      double srcq, kapq, rcpq = 0.0, bxq = 0.0, byq = 0.0, bzq = 0.0;
      if constexpr (CF::staged) {
        constexpr int AB = TRANSIENT ? 3 : 2;
        srcq = stage[(0 * 8 + q) * stride]; kapq = stage[(1 * 8 + q) * stride];
        if (TRANSIENT) rcpq = stage[(2 * 8 + q) * stride];
        if (ADV) { bxq = stage[((AB + 0) * 8 + q) * stride]; byq = stage[((AB + 1) * 8 + q) * stride]; bzq = stage[((AB + 2) * 8 + q) * stride]; }
      } else {
        double pt[3];
#pragma unroll
        for (int i = 0; i < 3; ++i)
          pt[i] = CSM ? ptc[i] : C[0][i] + C[1][i] * sx + C[2][i] * sy + C[4][i] * sz + C[3][i] * xe + C[5][i] * xz2 + C[6][i] * ez2 + C[7][i] * xez;
        const double x = pt[0], y = pt[1], z = pt[2];
        const double t = kp.tp.time;
        srcq = CF::source(x, y, z, t); kapq = CF::diff(x, y, z, t);
        if (TRANSIENT) rcpq = CF::rhocp(x, y, z, t);
        if (ADV) { bxq = CF::bx(x, y, z, t); byq = CF::by(x, y, z, t); bzq = CF::bz(x, y, z, t); }
      }
      const double kw = kapq * w;
      double f0 = -srcq * w;
      if (TRANSIENT) f0 += rcpq * edot * w;
      if (ADV) f0 += (bxq * ex + byq * ey + bzq * ez) * w;
      const double fx = kw * ex, fy = kw * ey, fz = kw * ez;
#pragma unroll
      for (int s = 0; s < 8; ++s) r[s] += f0 * Nq[s] + fx * gx[s] + fy * gy[s] + fz * gz[s];
      if (JAC) {
        const double kwa = kw * au;
        const double mw = TRANSIENT ? rcpq * at * w : 0.0;
        // accumulators in shared memory are walked row by row with the next row's loads issued ahead of this row's
        // arithmetic; the compiler barrier keeps it from hoisting all 36 loads (72 registers) at once
        double knext[8];
        if constexpr (KA::in_smem) {
#pragma unroll
          for (int j = 0; j < 8; ++j) knext[j] = K.get(ADV ? j : HexQ1::tri(0, j));
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const double ax = kwa * gx[i], ay = kwa * gy[i], az = kwa * gz[i];
          const double mi = TRANSIENT ? mw * Nq[i] : 0.0;
          double kcur[8];
          if constexpr (KA::in_smem) {
#pragma unroll
            for (int j = 0; j < 8; ++j) kcur[j] = knext[j];
            if (i < 7) {
#pragma unroll
              for (int j = (ADV ? 0 : i + 1); j < 8; ++j) knext[j] = K.get(ADV ? (i + 1) * 8 + j : HexQ1::tri(i + 1, j));
            }
          } else {
#pragma unroll
            for (int j = (ADV ? 0 : i); j < 8; ++j) kcur[j] = K.get(ADV ? i * 8 + j : HexQ1::tri(i, j));
          }
          if (!ADV) {
#pragma unroll
            for (int j = i; j < 8; ++j) {
              double v = kcur[j] + ax * gx[j] + ay * gy[j] + az * gz[j];
              if (TRANSIENT) v += mi * Nq[j];
              K.set(HexQ1::tri(i, j), v);
            }
          } else {
            const double ci = au * w * Nq[i];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              double v = kcur[j] + ax * gx[j] + ay * gy[j] + az * gz[j];
              v += ci * (bxq * gx[j] + byq * gy[j] + bzq * gz[j]);
              if (TRANSIENT) v += mi * Nq[j];
              K.set(i * 8 + j, v);
            }
          }
#if defined(__CUDA_ARCH__)
          if constexpr (KA::in_smem) asm volatile("" ::: "memory");
#endif
        }
      }
    }

  }

  MRH_HD static double kval(const double (&K)[NK], int i, int j) {
    return ADV ? K[i * 8 + j] : (i <= j ? K[HexQ1::tri(i, j)] : K[HexQ1::tri(j, i)]);
  }

  // ---- direct scatter: residual subtracted, Jacobian added, fixed rows skipped (one RED per entry) --------
  MRH_HDN static void scatter(const KParams& kp, int64_t e, const int (&lid)[8], const double (&r)[8], const double (&K)[NK]) {
    bool fx_[8];
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      fx_[s] = kp.fixed[lid[s]] != 0;
      if (!fx_[s]) atomic_add(kp.res + lid[s], -r[s]);
    }
    if (!JAC) return;
    uint16_t pat[64];
    {
      const uint4* pp = reinterpret_cast<const uint4*>(kp.pat + (size_t)kp.pat_id[e] * 64);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint4 v = pp[k];
        pat[k * 8 + 0] = (uint16_t)(v.x & 0xffff); pat[k * 8 + 1] = (uint16_t)(v.x >> 16);
        pat[k * 8 + 2] = (uint16_t)(v.y & 0xffff); pat[k * 8 + 3] = (uint16_t)(v.y >> 16);
        pat[k * 8 + 4] = (uint16_t)(v.z & 0xffff); pat[k * 8 + 5] = (uint16_t)(v.z >> 16);
        pat[k * 8 + 6] = (uint16_t)(v.w & 0xffff); pat[k * 8 + 7] = (uint16_t)(v.w >> 16);
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (fx_[i]) continue;
      double* jrow = kp.jac + kp.rowptr[lid[i]];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        atomic_add(jrow + pat[i * 8 + j], kval(K, i, j));
      }
    }
  }
#if defined(__CUDACC__)
  // ---- tile-staged scatter (device only) ------------------------------------------------------------------
  // One CTA owns a tile of 128 elements (kp.tile, built at finalize).  Every thread parks its element matrix and
  // residual in shared memory (entry-major, conflict free).  The tile's 128 x 72 contributions were sorted by
  // destination (CRS slot, then residual row) at plan time; thread t walks sorted positions [72 t, 72 t + 72),
  // sums each run of equal destinations in a register and stores the run total into the tile's accumulation
  // array - no shared-memory atomics except for the one run per thread that started in the previous thread's
  // chunk.  The CTA then flushes the array with REDs whose lanes walk consecutive CRS slots: against one RED
  // per element entry this halves the REDs and lets several lanes share a 32-byte sector, which is what bounds
  // RED.F64 throughput on B200 (profiles/r1_red_microbench.txt).
  // While an element is computed, LDGSTS copies bring the tile's slot / row lists into shared memory; the gather
  // loads of the NEXT tile are issued before this tile's scatter and the LID / node rows of the tile after that are
  // copied asynchronously, so that neither the flush nor the next element computation waits on DRAM.
  // dynamic shared memory: Ks[(NK + 8) * 128] | zero cell | pad | acc[max_acc + 1] | carry[128] | rowbuf[128][16] int
  //                        | off[max_e] uint32 | rows[max_r] int
  static constexpr int NE = NK + 8;
  __host__ __device__ static constexpr size_t tiled_smem_bytes(int max_acc, int max_e, int max_r) {
    return ((size_t)NE * 128 + 2 + (size_t)((max_acc + 2) & ~1) + 128) * sizeof(double) + 128 * 64 + ((size_t)max_e + (size_t)max_r) * 4 + WS_LIST_SLACK;
  }
  __device__ __forceinline__ static void cp_async4(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
  }
  __device__ __forceinline__ static void cp_async16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
  }
  __device__ static void tiled(const KParams& kp, double* stage, int stride, double* sm) {
    static_assert(JAC, "the tile-staged scatter is used for Jacobian assembly only");
    const TilePlan& tp = kp.tile;
    const int tid = threadIdx.x;
    double* Ks = sm;
    double* acc = sm + NE * 128 + 2;
    double* carry = acc + ((tp.max_acc + 2) & ~1);        // keeps the row buffer 16-byte aligned
    int* rowbuf = reinterpret_cast<int*>(carry + 128) + tid * 16;          // my element's 8 LIDs | 8 node ids
    unsigned* s_off = reinterpret_cast<unsigned*>(carry + 128) + 128 * 16;
    int* s_rows = reinterpret_cast<int*>(s_off + tp.max_e);
    if (tid == 0) Ks[NE * 128] = 0.0;                     // masked contributions read this cell
    const int64_t tstep = gridDim.x;
    auto element_of = [&](int64_t t) -> int { const int64_t e0 = t * 128 + tid; return (t < tp.tile_end && e0 < kp.nwork) ? tp.perm[e0] : -1; };
    auto fetch_rows = [&](int e) {
      if (e < 0) return;
      cp_async16(rowbuf, kp.lids + (int64_t)e * 8); cp_async16(rowbuf + 4, kp.lids + (int64_t)e * 8 + 4);
      cp_async16(rowbuf + 8, kp.e2n + (int64_t)e * 8); cp_async16(rowbuf + 12, kp.e2n + (int64_t)e * 8 + 4);
    };
    // software pipeline over the CTA's tiles: while tile t is scattered, the gather loads of tile t + 1 are in
    // flight and the LID / node rows of tile t + 2 are being copied into rowbuf
    const int64_t tfirst = tp.tile_begin + blockIdx.x;
    int e_cur = element_of(tfirst), e_next = element_of(tfirst + tstep), e_next2 = element_of(tfirst + 2 * tstep);
    double X[8][3], us[8], ud[8];
    {
      int lid[8];
      if (e_cur >= 0) gather(kp, kp.lids + (int64_t)e_cur * 8, kp.e2n + (int64_t)e_cur * 8, lid, X, us, ud);
    }
    fetch_rows(e_next);
    asm volatile("cp.async.commit_group;" ::: "memory");
    for (int64_t t = tfirst; t < tp.tile_end; t += tstep) {
      const int64_t eb = tp.ebeg[t], rb = tp.rptr[t];
      const int E = (int)(tp.eptr[t + 1] - tp.eptr[t]), R = (int)(tp.rptr[t + 1] - rb);
      for (int k = tid; k < E; k += 128) cp_async4(s_off + k, tp.off + eb + k);
      for (int k = tid; k < R; k += 128) cp_async4(s_rows + k, tp.rows + rb + k);
      asm volatile("cp.async.commit_group;" ::: "memory");
      if (e_cur >= 0) {
        double r[8], K[NK];
        work(kp, X, us, ud, stage, stride, r, K);
#pragma unroll
        for (int k = 0; k < NK; ++k) Ks[k * 128 + tid] = K[k];
#pragma unroll
        for (int s = 0; s < 8; ++s) Ks[(NK + s) * 128 + tid] = -r[s];
      }
      asm volatile("cp.async.wait_group 0;" ::: "memory");                 // slot / row lists of this tile, my rows of the next
      if (e_next >= 0) {
        int lid[8];
        gather(kp, rowbuf, rowbuf + 8, lid, X, us, ud);                    // consumed after this tile's scatter
      }
      const int e_next3 = element_of(t + 3 * tstep);
#if defined(MRH_DEBUG_SKIP_SCATTER)                         // tuning experiment: element computation only (results are NOT assembled)
      __syncthreads();
      fetch_rows(e_next2);
      asm volatile("cp.async.commit_group;" ::: "memory");
      e_cur = e_next; e_next = e_next2; e_next2 = e_next3;
      continue;
#endif
      const size_t cid = (size_t)tp.cl_id[t] * 128 + tid;
      unsigned w[36];
      {
        const uint4* cl = reinterpret_cast<const uint4*>(tp.clist + cid * 72);
#pragma unroll
        for (int i = 0; i < 9; ++i) { const uint4 v = __ldg(cl + i); w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w; }
      }
      const int loc0 = tp.cbase[cid];
      const bool cont = (w[0] & 0x8000u) == 0;            // my first run started in the previous thread's chunk
      __syncthreads();                                    // element store and list copies complete
      {
        double sum = 0.0;
        int loc = loc0;
        double* dst = cont ? carry + tid : acc + loc;
#pragma unroll
        for (int g = 0; g < 72; g += 8) {                 // loads of a group are issued ahead of its (possibly aliasing) stores
          double v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int k = g + u;
            const unsigned c = (k & 1) ? (w[k >> 1] >> 16) : (w[k >> 1] & 0xffffu);
            v[u] = Ks[c & 0x7fffu];
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int k = g + u;
            const unsigned c = (k & 1) ? (w[k >> 1] >> 16) : (w[k >> 1] & 0xffffu);
            if (k > 0 && (c & 0x8000u)) { *dst = sum; ++loc; dst = acc + loc; sum = 0.0; }
            sum += v[u];
          }
        }
        *dst = sum;
      }
      __syncthreads();                                    // run totals stored
      if (cont) atomicAdd(acc + loc0, carry[tid]);
      __syncthreads();
      {
        double* jb = kp.jac + tp.base[t];
        for (int k0 = tid; k0 < E; k0 += 512) {
          unsigned o[4]; double v[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) { const int k = k0 + 128 * u; if (k < E) { o[u] = s_off[k]; v[u] = acc[k]; } }
#pragma unroll
          for (int u = 0; u < 4; ++u) { const int k = k0 + 128 * u; if (k < E) atomicAdd(jb + o[u], v[u]); }
        }
        for (int k = tid; k < R; k += 128) atomicAdd(kp.res + s_rows[k], acc[E + k]);
      }
      __syncthreads();                                    // s_off / s_rows / acc free for the next tile's copies
      fetch_rows(e_next2);                                // rowbuf was consumed by the gather above
      asm volatile("cp.async.commit_group;" ::: "memory");
      e_cur = e_next; e_next = e_next2; e_next2 = e_next3;
    }
  }
  // ---- warp-specialised tile-staged assembly (device only; NVRTC kernel) ---------------------------------------
  // One CTA of 384 threads per SM: warpgroups 1 and 2 (MRH_WS_CREG registers per thread after setmaxnreg) only compute
  // elements - each owns a stream of tiles, parks K and r in its own shared element store and signals a named
  // barrier; warpgroup 0 (MRH_WS_SREG registers) does everything else for both streams: contribution-list run sums, carry
  // adds, RED flush, list copies.  The FP64 pipe therefore always sees two computing warps per scheduler instead of
  // alternating with the scatter phases of the same warps.
  // dynamic shared memory: Ks[WS_NBUF][NE * 128 + 2] | acc[(max_acc + 2) & ~1] | carry[128] | rowbuf[2][128][16] int
  //                        | off[max_e] uint32 | rows[max_r] int
  static constexpr int WS_NBUF = 3;               // element stores shared by the two compute groups (see tiled_ws)
  static constexpr int WS_LIST_SLACK = 96;        // bulk-copy staging of the lists: 16-byte groups around each list + one mbarrier
  __host__ __device__ static constexpr size_t ws_smem_bytes(int max_acc, int max_e, int max_r) {
    return (WS_NBUF * ((size_t)NE * 128 + 2) + (size_t)((max_acc + 2) & ~1) + 128) * sizeof(double) + 2 * 128 * 64 + ((size_t)max_e + (size_t)max_r) * 4 + WS_LIST_SLACK;
  }
  // K-in-registers kernel (tiled_wsk): two element stores, monomial coefficients of both compute groups in shared memory
  static constexpr int WSK_NBUF = 2;
  __host__ __device__ static constexpr size_t wsk_smem_bytes(int max_acc, int max_e, int max_r) {
    return (WSK_NBUF * ((size_t)NE * 128 + 2) + (size_t)((max_acc + 2) & ~1) + 128) * sizeof(double) + 2 * 128 * 64 + 2 * 32 * 128 * sizeof(double) +
           ((size_t)max_e + (size_t)max_r) * 4 + WS_LIST_SLACK;
  }
  // lane-per-point kernel (tiled_ws8): element-store stride 129, NCG compute warpgroups of 16 elements in flight each
  static constexpr int W8_STR = TRANSIENT ? 44 : 36;      // doubles of monomial scratch per element in flight (bank-spread for LDS.128)
  __host__ __device__ static constexpr size_t ws8_scratch_bytes_of(int ncg) { return (size_t)ncg * 16 * W8_STR * sizeof(double); }
  __host__ __device__ static constexpr size_t ws8_smem_bytes(int max_acc, int max_e, int max_r, int ncg) {
    return (WS_NBUF * ((size_t)NE * 129 + 2) + (size_t)((max_acc + 2) & ~1) + 128) * sizeof(double) + ws8_scratch_bytes_of(ncg) + ((size_t)max_e + (size_t)max_r) * 4 + WS_LIST_SLACK;
  }
#if defined(MRH_WS_CREG)
  __device__ __forceinline__ static void bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
  __device__ __forceinline__ static void bar_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }
  // scatter warpgroup of the warp-specialised kernels: serves the element stores of NCG compute groups in hand-over order
  // (sequence number s = NCG i + g is tile i of compute group g and lives in element store s % WS_NBUF).  KPAD: the element
  // store's entry stride is 128 + KPAD doubles (contribution lists address cells as entry * 128 + thread).
  template <int NCG, int KPAD, int NBUF = WS_NBUF>
  __device__ __forceinline__ static void ws_scatter_group(const KParams& kp, double* sm, int tid, int scratch_bytes) {
    const TilePlan& tp = kp.tile;
    constexpr int KSZ = NE * (128 + KPAD) + 2;
    double* acc = sm + NBUF * KSZ;
    double* carry = acc + ((tp.max_acc + 2) & ~1);
#if defined(MRH_WS_BULK)
    unsigned* s_off0 = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(carry + 128) + scratch_bytes);       // 16-byte aligned
    int* s_rows0 = reinterpret_cast<int*>(s_off0 + ((tp.max_e + 7) & ~3));
    const unsigned mbar = (unsigned)__cvta_generic_to_shared(s_rows0 + ((tp.max_r + 7) & ~3));
    unsigned phase = 0;
    if (tid == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    bar_sync(1 + 2 * NBUF, 128);
#else
    unsigned* s_off = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(carry + 128) + scratch_bytes);
    int* s_rows = reinterpret_cast<int*>(s_off + tp.max_e);
#endif
    const int64_t tstep = NCG * (int64_t)gridDim.x;
    constexpr int BAR_FULL = 1, BAR_EMPTY = 1 + NBUF, BAR_SCATTER = 1 + 2 * NBUF;
    constexpr int NPAIR = 256;                              // a FULL / EMPTY barrier joins one compute group (128) and the scatter group (128)
    auto tile_of = [&](int s) -> int64_t { return tp.tile_begin + NCG * (int64_t)blockIdx.x + (s % NCG) + (int64_t)(s / NCG) * tstep; };
    if (tid < NBUF) sm[tid * KSZ + NE * (128 + KPAD)] = 0.0;    // masked contributions read these cells
    struct TileMeta { int64_t eb, rb, base; int E, R, clid; };
    auto load_meta = [&](int64_t t, TileMeta& m) {
      if (t >= tp.tile_end) { m.E = m.R = 0; m.clid = 0; m.eb = m.rb = m.base = 0; return; }
      m.eb = tp.ebeg[t]; m.rb = tp.rptr[t]; m.base = tp.base[t];
      m.E = (int)(tp.eptr[t + 1] - tp.eptr[t]); m.R = (int)(tp.rptr[t + 1] - m.rb); m.clid = tp.cl_id[t];
    };
    TileMeta cur, nxt;
    load_meta(tile_of(0), cur);
    int b = 0;                                            // element store of sequence number s
    for (int s = 0; tile_of(s) < tp.tile_end; ++s) {
      const double* Ks = sm + b * KSZ;
      const int E = cur.E, R = cur.R;
#if defined(MRH_WS_BULK)
      // slot / row lists of the tile by two bulk copies (TMA unit, one thread issues, completion on an mbarrier): whole 16-byte
      // groups around [eb, eb + E) and [rb, rb + R) land at the start of the staging areas, the lists start (eb & 3) / (rb & 3) in
      const unsigned* s_off = s_off0 + (cur.eb & 3);
      const int* s_rows = s_rows0 + (cur.rb & 3);
      if (tid == 0) {
        const unsigned nb_e = (unsigned)(((cur.eb & 3) + E + 3) & ~3) * 4u, nb_r = (unsigned)(((cur.rb & 3) + R + 3) & ~3) * 4u;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the previous tile's generic-proxy reads of the staging areas are done (BAR_SCATTER)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(nb_e + nb_r) : "memory");
        if (nb_e) asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                               ::"r"((unsigned)__cvta_generic_to_shared(s_off0)), "l"(tp.off + (cur.eb & ~(int64_t)3)), "r"(nb_e), "r"(mbar) : "memory");
        if (nb_r) asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                               ::"r"((unsigned)__cvta_generic_to_shared(s_rows0)), "l"(tp.rows + (cur.rb & ~(int64_t)3)), "r"(nb_r), "r"(mbar) : "memory");
      }
#else
      for (int k = tid; k < E; k += 128) cp_async4(s_off + k, tp.off + cur.eb + k);
      for (int k = tid; k < R; k += 128) cp_async4(s_rows + k, tp.rows + cur.rb + k);
      asm volatile("cp.async.commit_group;" ::: "memory");
#endif
      const size_t cid = (size_t)cur.clid * 128 + tid;
      const uint4* cl = reinterpret_cast<const uint4*>(tp.clist + cid * 72);
      uint4 wv = __ldg(cl);
      const int loc0 = tp.cbase[cid];
      const bool cont = (wv.x & 0x8000u) == 0;            // my first run started in the previous thread's chunk
      load_meta(tile_of(s + 1), nxt);                     // in flight during this tile's scatter
      bar_sync(BAR_FULL + b, NPAIR);                        // element store b complete
      {
        double sum = 0.0;
        int loc = loc0;
        double* dst = cont ? carry + tid : acc + loc;
#pragma unroll
        for (int c = 0; c < 9; ++c) {
          const unsigned w[4] = {wv.x, wv.y, wv.z, wv.w};
          if (c < 8) wv = __ldg(cl + c + 1);
          double v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const unsigned cc = (u & 1) ? (w[u >> 1] >> 16) : (w[u >> 1] & 0xffffu);
            const unsigned cell = cc & 0x7fffu;
            v[u] = Ks[cell + KPAD * (cell >> 7)];
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const unsigned cc = (u & 1) ? (w[u >> 1] >> 16) : (w[u >> 1] & 0xffffu);
            if ((c > 0 || u > 0) && (cc & 0x8000u)) { *dst = sum; ++loc; dst = acc + loc; sum = 0.0; }
            sum += v[u];
          }
        }
        *dst = sum;
      }
      bar_sync(BAR_SCATTER, 128);                         // run totals stored, element store no longer read
      bar_arrive(BAR_EMPTY + b, NPAIR);
      if (cont) atomicAdd(acc + loc0, carry[tid]);
#if defined(MRH_WS_BULK)
      {
        unsigned done = 0;
        while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(mbar), "r"(phase) : "memory");
        phase ^= 1u;
      }
#else
      asm volatile("cp.async.wait_group 0;" ::: "memory");
#endif
      bar_sync(BAR_SCATTER, 128);
      {
        double* jb = kp.jac + cur.base;
        for (int k0 = tid; k0 < E; k0 += 512) {
          unsigned o[4]; double v[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) { const int k = k0 + 128 * u; if (k < E) { o[u] = s_off[k]; v[u] = acc[k]; } }
#pragma unroll
          for (int u = 0; u < 4; ++u) { const int k = k0 + 128 * u; if (k < E) atomicAdd(jb + o[u], v[u]); }
        }
        for (int k = tid; k < R; k += 128) atomicAdd(kp.res + s_rows[k], acc[E + k]);
      }
      bar_sync(BAR_SCATTER, 128);                         // lists and accumulation array free again
      cur = nxt;
      b = (b + 1 == NBUF) ? 0 : b + 1;
    }
  }
  __device__ static void tiled_ws(const KParams& kp, double* sm) {
    static_assert(JAC, "the tile-staged scatter is used for Jacobian assembly only");
    static_assert(!CF::staged, "the warp-specialised kernel is built with NVRTC-printed coefficients");
    const TilePlan& tp = kp.tile;
    const int wg = threadIdx.x >> 7, tid = threadIdx.x & 127;
    constexpr int KSZ = NE * 128 + 2;
    double* acc = sm + WS_NBUF * KSZ;
    double* carry = acc + ((tp.max_acc + 2) & ~1);
    int* rowbase = reinterpret_cast<int*>(carry + 128);
    unsigned* s_off = reinterpret_cast<unsigned*>(rowbase + 2 * 128 * 16);
    int* s_rows = reinterpret_cast<int*>(s_off + tp.max_e);
    const int64_t tstep = 2 * (int64_t)gridDim.x;
    // The CTA's tiles in hand-over order: sequence number s = 2 i + g is tile i of compute group g; it lives in
    // element store s % WS_NBUF.  Named barriers (0 is __syncthreads): FULL[b] compute -> scatter, EMPTY[b] back.
    constexpr int BAR_FULL = 1, BAR_EMPTY = 1 + WS_NBUF, BAR_SCATTER = 1 + 2 * WS_NBUF;
    auto tile_of = [&](int s) -> int64_t { return tp.tile_begin + 2 * (int64_t)blockIdx.x + (s & 1) + (int64_t)(s >> 1) * tstep; };
    if (wg == 0) {
      asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(MRH_WS_SREG));
      ws_scatter_group<2, 0>(kp, sm, tid, 2 * 128 * 64);
    } else {
      // ================================= compute warpgroups =================================
      asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(MRH_WS_CREG));
      const int g = wg - 1;
      int* rowbuf = rowbase + (g * 128 + tid) * 16;         // my next element's 8 LIDs | 8 node ids
      auto element_of = [&](int64_t t) -> int { const int64_t e0 = t * 128 + tid; return (t < tp.tile_end && e0 < kp.nwork) ? tp.perm[e0] : -1; };
      auto fetch_rows = [&](int e) {
        if (e < 0) return;
        cp_async16(rowbuf, kp.lids + (int64_t)e * 8); cp_async16(rowbuf + 4, kp.lids + (int64_t)e * 8 + 4);
        cp_async16(rowbuf + 8, kp.e2n + (int64_t)e * 8); cp_async16(rowbuf + 12, kp.e2n + (int64_t)e * 8 + 4);
      };
      const int64_t tfirst = tile_of(g);
      // pipeline per thread: registers hold the gathered data of tile i, rowbuf the LID / node rows of tile i + 1.
      // (An explicit L2 prefetch of tile i + 2's vertices was measured slower: CCTL.PF2 stalls the issuing warp.)
      int e_cur = element_of(tfirst), e_next = element_of(tfirst + tstep), e_next2 = element_of(tfirst + 2 * tstep);
      double X[8][3], us[8], ud[8];
      {
        int lid[8];
        if (e_cur >= 0) gather(kp, kp.lids + (int64_t)e_cur * 8, kp.e2n + (int64_t)e_cur * 8, lid, X, us, ud);
      }
      fetch_rows(e_next);
      asm volatile("cp.async.commit_group;" ::: "memory");
      int s = g, b = g % WS_NBUF;
      for (int64_t t = tfirst; t < tp.tile_end; t += tstep) {
        double* Ks = sm + b * KSZ + tid;
        if (s >= WS_NBUF) bar_sync(BAR_EMPTY + b, 256);                  // the scatter group has consumed this store's previous tile
        if (e_cur >= 0) {
          // the element matrix is accumulated in place in the shared element store: no register copy of K, which
          // removes every spill from the point loop and leaves the scheduler room to overlap dependent chains
          double r[8];
          KSmem kacc{Ks};
          work(kp, X, us, ud, nullptr, 0, r, kacc);
#pragma unroll
          for (int k = 0; k < 8; ++k) Ks[(NK + k) * 128] = -r[k];
        }
        bar_arrive(BAR_FULL + b, 256);
        asm volatile("cp.async.wait_group 0;" ::: "memory");             // my rows of the next tile
        if (e_next >= 0) {
          int lid[8];
          gather(kp, rowbuf, rowbuf + 8, lid, X, us, ud);
        }
        const int e_next3 = element_of(t + 3 * tstep);
        fetch_rows(e_next2);                                             // rowbuf was consumed by the gather above
        asm volatile("cp.async.commit_group;" ::: "memory");
        e_cur = e_next; e_next = e_next2; e_next2 = e_next3;
        s += 2; b = (b + 2) % WS_NBUF;
      }
    }
  }
#if defined(MRH_WSK_CREG)
  // K-in-registers form of tiled_ws: the compute groups keep the element matrix in registers over the point loop and park the
  // monomial coefficients of the trilinear map in shared memory instead (24 reads per point against 72 read-modify-writes of the
  // element store, whose traffic made the shared-memory pipe the co-limiter of tiled_ws); the element store is written once per
  // element at hand-over, so two stores suffice.  Shared memory: Ks[2] | acc | carry | rowbuf[2][128][16] int | Cs[2][32 * 128] | off | rows
  __device__ static void tiled_wsk(const KParams& kp, double* sm) {
    static_assert(JAC, "the tile-staged scatter is used for Jacobian assembly only");
    static_assert(!CF::staged, "the warp-specialised kernel is built with NVRTC-printed coefficients");
    const TilePlan& tp = kp.tile;
    const int wg = threadIdx.x >> 7, tid = threadIdx.x & 127;
    constexpr int KSZ = NE * 128 + 2;
    double* acc = sm + WSK_NBUF * KSZ;
    double* carry = acc + ((tp.max_acc + 2) & ~1);
    int* rowbase = reinterpret_cast<int*>(carry + 128);
    double* csbase = reinterpret_cast<double*>(rowbase + 2 * 128 * 16);
    const int64_t tstep = 2 * (int64_t)gridDim.x;
    // The CTA's tiles in hand-over order: sequence number s = 2 i + g is tile i of compute group g; it lives in
    // element store s % WS_NBUF.  Named barriers (0 is __syncthreads): FULL[b] compute -> scatter, EMPTY[b] back.
    constexpr int BAR_FULL = 1, BAR_EMPTY = 1 + WSK_NBUF;
    auto tile_of = [&](int s) -> int64_t { return tp.tile_begin + 2 * (int64_t)blockIdx.x + (s & 1) + (int64_t)(s >> 1) * tstep; };
    if (wg == 0) {
      asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(MRH_WSK_SREG));
      ws_scatter_group<2, 0, WSK_NBUF>(kp, sm, tid, 2 * 128 * 64 + 2 * 32 * 128 * (int)sizeof(double));
    } else {
      // ================================= compute warpgroups =================================
      asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(MRH_WSK_CREG));
      double* cs = csbase + (size_t)(wg - 1) * 32 * 128 + tid;
      const int g = wg - 1;
      int* rowbuf = rowbase + (g * 128 + tid) * 16;         // my next element's 8 LIDs | 8 node ids
      auto element_of = [&](int64_t t) -> int { const int64_t e0 = t * 128 + tid; return (t < tp.tile_end && e0 < kp.nwork) ? tp.perm[e0] : -1; };
      auto fetch_rows = [&](int e) {
        if (e < 0) return;
        cp_async16(rowbuf, kp.lids + (int64_t)e * 8); cp_async16(rowbuf + 4, kp.lids + (int64_t)e * 8 + 4);
        cp_async16(rowbuf + 8, kp.e2n + (int64_t)e * 8); cp_async16(rowbuf + 12, kp.e2n + (int64_t)e * 8 + 4);
      };
      const int64_t tfirst = tile_of(g);
      // pipeline per thread: registers hold the gathered data of tile i, rowbuf the LID / node rows of tile i + 1.
      // (An explicit L2 prefetch of tile i + 2's vertices was measured slower: CCTL.PF2 stalls the issuing warp.)
      int e_cur = element_of(tfirst), e_next = element_of(tfirst + tstep), e_next2 = element_of(tfirst + 2 * tstep);
      double X[8][3], us[8], ud[8];
      {
        int lid[8];
        if (e_cur >= 0) gather(kp, kp.lids + (int64_t)e_cur * 8, kp.e2n + (int64_t)e_cur * 8, lid, X, us, ud);
      }
      fetch_rows(e_next);
      asm volatile("cp.async.commit_group;" ::: "memory");
      int s = g, b = g % WSK_NBUF;
      for (int64_t t = tfirst; t < tp.tile_end; t += tstep) {
        double* Ks = sm + b * KSZ + tid;
        double r[8], K[NK];
        if (e_cur >= 0) {
          KReg kacc{K};
          work<KReg, true>(kp, X, us, ud, nullptr, 0, r, kacc, cs);
        }
        if (s >= WSK_NBUF) bar_sync(BAR_EMPTY + b, 256);                 // the scatter group has consumed this store's previous tile
        if (e_cur >= 0) {
#pragma unroll
          for (int k = 0; k < NK; ++k) Ks[k * 128] = K[k];
#pragma unroll
          for (int k = 0; k < 8; ++k) Ks[(NK + k) * 128] = -r[k];
        }
        bar_arrive(BAR_FULL + b, 256);
        asm volatile("cp.async.wait_group 0;" ::: "memory");             // my rows of the next tile
        if (e_next >= 0) {
          int lid[8];
          gather(kp, rowbuf, rowbuf + 8, lid, X, us, ud);
        }
        const int e_next3 = element_of(t + 3 * tstep);
        fetch_rows(e_next2);                                             // rowbuf was consumed by the gather above
        asm volatile("cp.async.commit_group;" ::: "memory");
        e_cur = e_next; e_next = e_next2; e_next2 = e_next3;
        s += 2; b = (b + 2) % WSK_NBUF;
      }
    }
  }
#endif  // MRH_WSK_CREG
#if defined(MRH_WS8_NCG)
  // ---- lane-per-point form of the warp-specialised kernel (device only; NVRTC kernel) --------------------------
  // Eight consecutive lanes own one element: lane q gathers vertex q and evaluates quadrature point q.  Nothing is
  // accumulated over points in shared memory: every lane forms the 36 products of ITS point and the eight lanes combine them
  // with a transposing butterfly (shuffles over lane^4, ^2, ^1; each stage halves the entries a lane keeps), after which lane q
  // holds entries 5q .. 5q+4 of the element matrix and entry q of the residual and parks them in the shared element store.
  // Against tiled_ws this removes the 576 shared-memory read-modify-writes per element (its shared-memory pipe was 75 % busy),
  // needs ~130 instead of 200 registers per compute thread (MRH_WS8_NCG = 3 compute warpgroups) and makes every gather lane
  // load one vertex.  The element store uses an entry stride of 129 doubles so that the 8 x 4 lanes of a store instruction
  // fall into different banks; the scatter warpgroup is the same code as in tiled_ws.
  MRH_HD static constexpr int tri_i(int k) { int i = 0; while (k >= 8 - i) { k -= 8 - i; ++i; } return i; }
  MRH_HD static constexpr int tri_j(int k) { int i = 0; while (k >= 8 - i) { k -= 8 - i; ++i; } return i + k; }
  __host__ __device__ static constexpr size_t ws8_scratch_bytes() { return ws8_scratch_bytes_of(MRH_WS8_NCG); }
  __device__ __forceinline__ static double shx(double v, int mask) { return __shfl_xor_sync(0xffffffffu, v, mask); }
  // c ? a : b as an opaque select: written with ?: on array elements the compiler turns the pair into ONE dynamically indexed
  // local-memory array (a[j + 10 c]), which puts the butterfly's partial sums on the stack
  __device__ __forceinline__ static double sel(bool c, double a, double b) {
    double r;
    asm("{ .reg .pred p; setp.ne.s32 p, %3, 0; selp.f64 %0, %1, %2, p; }" : "=d"(r) : "d"(a), "d"(b), "r"((int)c));
    return r;
  }
  __device__ static void tiled_ws8(const KParams& kp, double* sm) {
    static_assert(JAC && !ADV, "lane-per-point kernel: symmetric element matrix, Jacobian assembly");
    static_assert(!CF::staged, "the warp-specialised kernels are built with NVRTC-printed coefficients");
    constexpr int NCG = MRH_WS8_NCG, KST = 129;
    constexpr int KSZ = NE * KST + 2;
    const TilePlan& tp = kp.tile;
    const int wg = threadIdx.x >> 7, tid = threadIdx.x & 127;
    constexpr int BAR_FULL = 1, BAR_EMPTY = 1 + WS_NBUF;
    if (wg == 0) {
      asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(MRH_WS8_SREG));
      ws_scatter_group<NCG, 1>(kp, sm, tid, (int)ws8_scratch_bytes());
      return;
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(MRH_WS8_CREG));
    const int g = wg - 1;
    const int q = tid & 7, el = tid >> 3;                   // my vertex / point, my element within a pass of 16
    const bool b0 = (q & 1) != 0, b1 = (q & 2) != 0, b2 = (q & 4) != 0;
    const bool hx = ((q ^ (q >> 1)) & 1) != 0;              // tensor x bit of shards vertex q (y bit = b1, z bit = b2)
    const int mono = (hx ? 1 : 0) + (b1 ? 2 : 0) + (b2 ? 4 : 0);   // monomial this lane holds after the vertex butterfly
    double* acc = sm + WS_NBUF * KSZ;
    double* scr = acc + ((tp.max_acc + 2) & ~1) + 128 + (size_t)(g * 16 + el) * W8_STR;
    // constants of point q (descending 2-point rule: bit 0 -> +g): abscissae, half products of the 1-D shape functions
    const double gq = HexQ1::g;
    const double sx = b0 ? -gq : gq, sy = b1 ? -gq : gq, sz = b2 ? -gq : gq;
    const double sxy = sx * sy, syz = sy * sz, sxz = sx * sz, sxyz = sxy * sz;
    const double lx[2] = {0.5 - 0.5 * sx, 0.5 + 0.5 * sx}, ly[2] = {0.5 - 0.5 * sy, 0.5 + 0.5 * sy}, lz[2] = {0.5 - 0.5 * sz, 0.5 + 0.5 * sz};
    double hyz[2][2], hxz[2][2], hxy[2][2];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) { hyz[a][b] = 0.5 * ly[a] * lz[b]; hxz[a][b] = 0.5 * lx[a] * lz[b]; hxy[a][b] = 0.5 * lx[a] * ly[b]; }
    const double au = TRANSIENT ? kp.tp.alpha_u : 1.0, at = TRANSIENT ? kp.tp.alpha_t : 0.0;
    const int64_t tstep = NCG * (int64_t)gridDim.x;
    const int64_t tfirst = tp.tile_begin + NCG * (int64_t)blockIdx.x + g;
    // item it = (tile it / 8 of this group, pass it % 8): element position tile * 128 + 16 pass + el
    auto elem_at = [&](int64_t it) -> int {
      const int64_t t = tfirst + (it >> 3) * tstep;
      if (t >= tp.tile_end) return -1;
      const int64_t pos = t * 128 + (it & 7) * 16 + el;
      return pos < kp.nwork ? tp.perm[pos] : -1;
    };
    // software pipeline over items: vertex values of item it + 1 and indices of item it + 2 are in flight during item it
    int eC = elem_at(0), eN = elem_at(1), eN2 = elem_at(2);
    int lidN = 0, ndN = 0;
    double xC0 = 0.0, xC1 = 0.0, xC2 = 0.0, uC = 0.0, dC = 0.0, xN0 = 0.0, xN1 = 0.0, xN2 = 0.0, uN = 0.0, dN = 0.0;
#define MRH_W8_VALS(E, LID, ND, X0, X1, X2, U, D)                                                          \
    if ((E) >= 0) {                                                                                        \
      const double* p_ = kp.nodes + (int64_t)(ND) * 3; X0 = p_[0]; X1 = p_[1]; X2 = p_[2];                  \
      const double uval_ = kp.u[LID];                                                                      \
      if (!TRANSIENT) { U = uval_; D = 0.0; }                                                              \
      else {                                                                                               \
        const TimeParams& tq_ = kp.tp;                                                                     \
        const double prev0_ = kp.u_prev[(size_t)(LID) * tq_.nsteps];                                       \
        double bu_ = (1.0 - tq_.alpha_u) * prev0_;                                                         \
        for (int k_ = 0; k_ < tq_.stage; ++k_) bu_ += tq_.a_over_b[k_] * (kp.u_stage[(size_t)(LID) * tq_.nstages + k_] - prev0_); \
        double bt_ = 0.0;                                                                                  \
        for (int k_ = 1; k_ <= tq_.nsteps; ++k_) bt_ += tq_.bdf[k_] * kp.u_prev[(size_t)(LID) * tq_.nsteps + (k_ - 1)]; \
        U = tq_.alpha_u * uval_ + bu_; D = tq_.alpha_t * uval_ + bt_ * tq_.timewt;                         \
      }                                                                                                    \
    }
    if (eC >= 0) { const int lid_ = kp.lids[(int64_t)eC * 8 + q], nd_ = kp.e2n[(int64_t)eC * 8 + q]; MRH_W8_VALS(eC, lid_, nd_, xC0, xC1, xC2, uC, dC) }
    if (eN >= 0) { lidN = kp.lids[(int64_t)eN * 8 + q]; ndN = kp.e2n[(int64_t)eN * 8 + q]; }
    int sq = g;                                             // hand-over sequence number of the tile being filled
    for (int64_t it = 0; tfirst + (it >> 3) * tstep < tp.tile_end; ++it) {
      const int pass = (int)(it & 7);
      const int b = sq % WS_NBUF;
      double* Kst = sm + (size_t)b * KSZ + pass * 16 + el;  // my element's column of the element store
      if (pass == 0 && sq >= WS_NBUF) bar_sync(BAR_EMPTY + b, 256);       // the scatter group has consumed this store's previous tile
      MRH_W8_VALS(eN, lidN, ndN, xN0, xN1, xN2, uN, dN)
      int lidN2 = 0, ndN2 = 0;
      if (eN2 >= 0) { lidN2 = kp.lids[(int64_t)eN2 * 8 + q]; ndN2 = kp.e2n[(int64_t)eN2 * 8 + q]; }
      const int eN3 = elem_at(it + 3);
      {
        // ---- vertex butterfly over the three tensor bits: lane holds 8 x the monomial coefficient `mono` of x, y, z, u (, du/dt)
        double v0 = xC0, v1 = xC1, v2 = xC2, v3 = uC, v4 = dC;
#define MRH_W8_BF(MASK, HI) { const double o0 = shx(v0, MASK), o1 = shx(v1, MASK), o2 = shx(v2, MASK), o3 = shx(v3, MASK); \
          v0 = (HI) ? v0 - o0 : v0 + o0; v1 = (HI) ? v1 - o1 : v1 + o1; v2 = (HI) ? v2 - o2 : v2 + o2; v3 = (HI) ? v3 - o3 : v3 + o3; \
          if (TRANSIENT) { const double o4 = shx(v4, MASK); v4 = (HI) ? v4 - o4 : v4 + o4; } }
        MRH_W8_BF(1, hx) MRH_W8_BF(3, b1) MRH_W8_BF(4, b2)
#undef MRH_W8_BF
        __syncwarp();                                       // the previous item's reads of the scratch are complete
        *reinterpret_cast<double2*>(scr + mono * 4) = make_double2(0.125 * v0, 0.125 * v1);
        *reinterpret_cast<double2*>(scr + mono * 4 + 2) = make_double2(0.125 * v2, 0.125 * v3);
        if (TRANSIENT) scr[32 + mono] = 0.125 * v4;
        __syncwarp();
        // ---- Jacobian of the trilinear map, point coordinates, reference gradient of u: all from the monomial form
        double J[3][3], pt[3], g0, g1, g2, edot = 0.0;
        {
          const double2* c2 = reinterpret_cast<const double2*>(scr);
          double C[8][3], A[8];
#pragma unroll
          for (int m = 0; m < 8; ++m) { const double2 a = c2[2 * m], bb = c2[2 * m + 1]; C[m][0] = a.x; C[m][1] = a.y; C[m][2] = bb.x; A[m] = bb.y; }
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            J[i][0] = C[1][i] + C[3][i] * sy + C[5][i] * sz + C[7][i] * syz;
            J[i][1] = C[2][i] + C[3][i] * sx + C[6][i] * sz + C[7][i] * sxz;
            J[i][2] = C[4][i] + C[5][i] * sx + C[6][i] * sy + C[7][i] * sxy;
            pt[i] = C[0][i] + C[1][i] * sx + C[2][i] * sy + C[4][i] * sz + C[3][i] * sxy + C[5][i] * sxz + C[6][i] * syz + C[7][i] * sxyz;
          }
          g0 = A[1] + A[3] * sy + A[5] * sz + A[7] * syz;
          g1 = A[2] + A[3] * sx + A[6] * sz + A[7] * sxz;
          g2 = A[4] + A[5] * sx + A[6] * sy + A[7] * sxy;
          if (TRANSIENT) {
            const double* Dm = scr + 32;
            edot = Dm[0] + Dm[1] * sx + Dm[2] * sy + Dm[4] * sz + Dm[3] * sxy + Dm[5] * sxz + Dm[6] * syz + Dm[7] * sxyz;
          }
        }
        double Ji[3][3];
        const double det = det_inv<3>(J, Ji);
        const double w = fabs(det);                         // reference weight 1
        double gx[8], gy[8], gz[8];
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          const int a = HexQ1::bx(s), bq = HexQ1::by(s), c = HexQ1::bz(s);
          const double d0 = a ? hyz[bq][c] : -hyz[bq][c], d1 = bq ? hxz[a][c] : -hxz[a][c], d2 = c ? hxy[a][bq] : -hxy[a][bq];
          gx[s] = Ji[0][0] * d0 + Ji[1][0] * d1 + Ji[2][0] * d2;
          gy[s] = Ji[0][1] * d0 + Ji[1][1] * d1 + Ji[2][1] * d2;
          gz[s] = Ji[0][2] * d0 + Ji[1][2] * d1 + Ji[2][2] * d2;
        }
        const double ex = Ji[0][0] * g0 + Ji[1][0] * g1 + Ji[2][0] * g2;
        const double ey = Ji[0][1] * g0 + Ji[1][1] * g1 + Ji[2][1] * g2;
        const double ez = Ji[0][2] * g0 + Ji[1][2] * g1 + Ji[2][2] * g2;
        const double tm = kp.tp.time;
        const double srcq = CF::source(pt[0], pt[1], pt[2], tm), kapq = CF::diff(pt[0], pt[1], pt[2], tm);
        const double rcpq = TRANSIENT ? CF::rhocp(pt[0], pt[1], pt[2], tm) : 0.0;
        const double kw = kapq * w;
        double f0 = -srcq * w;
        if (TRANSIENT) f0 += rcpq * edot * w;
        const double fx = kw * ex, fy = kw * ey, fz = kw * ez;
        // ---- residual: 8 partial entries per lane, reduced over the 8 points by a halving butterfly; lane q keeps entry q
        double rq;
        {
          double r4[4], r2[2];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const double lo = f0 * (2.0 * hxy[HexQ1::bx(j)][HexQ1::by(j)] * lz[0]) + fx * gx[j] + fy * gy[j] + fz * gz[j];
            const double hi = f0 * (2.0 * hxy[HexQ1::bx(j + 4)][HexQ1::by(j + 4)] * lz[1]) + fx * gx[j + 4] + fy * gy[j + 4] + fz * gz[j + 4];
            r4[j] = sel(b2, hi, lo) + shx(sel(b2, lo, hi), 4);
          }
#pragma unroll
          for (int j = 0; j < 2; ++j) r2[j] = sel(b1, r4[j + 2], r4[j]) + shx(sel(b1, r4[j], r4[j + 2]), 2);
          rq = sel(b0, r2[1], r2[0]) + shx(sel(b0, r2[0], r2[1]), 1);
        }
        // ---- element matrix: 40 slots (36 upper entries + 4 unused); stage 1 pairs slot j with slot 20 + j
        const double kwa = kw * au;
        const double mw = TRANSIENT ? rcpq * at * w : 0.0;
        double a1[20];
        // (written out pair by pair: the upper-triangle index arithmetic must be compile-time for gx / gy / gz to stay in registers)
#define MRH_W8_NQ(S) (2.0 * hxy[HexQ1::bx(S)][HexQ1::by(S)] * lz[HexQ1::bz(S)])
#define MRH_W8_P(I, J) (kwa * (gx[I] * gx[J] + gy[I] * gy[J] + gz[I] * gz[J]) + (TRANSIENT ? mw * MRH_W8_NQ(I) * MRH_W8_NQ(J) : 0.0))
#define MRH_W8_PAIR(SLOT, PLO, PHI) { const double plo = PLO, phi = PHI; a1[SLOT] = sel(b2, phi, plo) + shx(sel(b2, plo, phi), 4); }
        MRH_W8_PAIR(0, MRH_W8_P(0, 0), MRH_W8_P(2, 7))
        MRH_W8_PAIR(1, MRH_W8_P(0, 1), MRH_W8_P(3, 3))
        MRH_W8_PAIR(2, MRH_W8_P(0, 2), MRH_W8_P(3, 4))
        MRH_W8_PAIR(3, MRH_W8_P(0, 3), MRH_W8_P(3, 5))
        MRH_W8_PAIR(4, MRH_W8_P(0, 4), MRH_W8_P(3, 6))
        MRH_W8_PAIR(5, MRH_W8_P(0, 5), MRH_W8_P(3, 7))
        MRH_W8_PAIR(6, MRH_W8_P(0, 6), MRH_W8_P(4, 4))
        MRH_W8_PAIR(7, MRH_W8_P(0, 7), MRH_W8_P(4, 5))
        MRH_W8_PAIR(8, MRH_W8_P(1, 1), MRH_W8_P(4, 6))
        MRH_W8_PAIR(9, MRH_W8_P(1, 2), MRH_W8_P(4, 7))
        MRH_W8_PAIR(10, MRH_W8_P(1, 3), MRH_W8_P(5, 5))
        MRH_W8_PAIR(11, MRH_W8_P(1, 4), MRH_W8_P(5, 6))
        MRH_W8_PAIR(12, MRH_W8_P(1, 5), MRH_W8_P(5, 7))
        MRH_W8_PAIR(13, MRH_W8_P(1, 6), MRH_W8_P(6, 6))
        MRH_W8_PAIR(14, MRH_W8_P(1, 7), MRH_W8_P(6, 7))
        MRH_W8_PAIR(15, MRH_W8_P(2, 2), MRH_W8_P(7, 7))
        MRH_W8_PAIR(16, MRH_W8_P(2, 3), 0.0)
        MRH_W8_PAIR(17, MRH_W8_P(2, 4), 0.0)
        MRH_W8_PAIR(18, MRH_W8_P(2, 5), 0.0)
        MRH_W8_PAIR(19, MRH_W8_P(2, 6), 0.0)
#undef MRH_W8_PAIR
#undef MRH_W8_P
#undef MRH_W8_NQ
        double a2[10];
#pragma unroll
        for (int j = 0; j < 10; ++j) a2[j] = sel(b1, a1[j + 10], a1[j]) + shx(sel(b1, a1[j], a1[j + 10]), 2);
        double a3[5];
#pragma unroll
        for (int j = 0; j < 5; ++j) a3[j] = sel(b0, a2[j + 5], a2[j]) + shx(sel(b0, a2[j], a2[j + 5]), 1);
        if (eC >= 0) {
#pragma unroll
          for (int j = 0; j < 5; ++j) { const int k = 5 * q + j; if (k < NK) Kst[k * KST] = a3[j]; }
          Kst[(NK + q) * KST] = -rq;
        }
      }
      if (pass == 7) { bar_arrive(BAR_FULL + b, 256); sq += NCG; }
      eC = eN; xC0 = xN0; xC1 = xN1; xC2 = xN2; uC = uN; dC = dN;
      eN = eN2; lidN = lidN2; ndN = ndN2; eN2 = eN3;
    }
#undef MRH_W8_VALS
  }
#endif  // MRH_WS8_NCG
#endif  // MRH_WS_CREG
#endif
};

}  // namespace mrh
