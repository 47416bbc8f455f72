// Specialised assembly for thermal on Q1 quadrilaterals with the 4-point Gauss rule (BASELINE.json config C1:
// regression/thermal/2D_verification at scale).  One thread owns one element; vertices, DOFs, the 4x4 element matrix
// (10 upper entries, 16 with advection) and the residual live in registers.  Same conventions as the hex kernel
// (assemble_fast.h): shards vertex order, descending Gauss abscissae (point 0 = (+,+)), residual subtracted, Jacobian
// added, fixed rows skipped (thermal.cpp:71-165, assemblyManager.cpp:4031-4145).  The bilinear map and the solution are
// held in monomial form (x = c0 + c1 xi + c2 eta + c3 xi eta), the general bilinear Jacobian is evaluated at every
// point (no affine shortcut).
#pragma once
#include "assemble_fast.h"

namespace mrh {

#if !defined(__CUDACC_RTC__)
inline bool thermal_quad_q1_eligible(int phys, int dim, int order, int nq, bool state_independent, int lump_mass) {
  return phys == PHYS_THERMAL && dim == 2 && order == 1 && nq == 4 && state_independent && !lump_mass;
}
#endif

template <bool TRANSIENT, bool ADV, bool JAC, class CF = InterpCoefs>
struct ThermalQuadQ1 {
  static constexpr int NK = ADV ? 16 : 10;
  MRH_HD static constexpr int tri(int i, int j) { return i * 4 - (i * (i - 1)) / 2 + (j - i); }      // i <= j, 10 entries
  MRH_HD static constexpr int bx(int s) { return (s == 1 || s == 2) ? 1 : 0; }                        // shards vertex -> tensor bits
  MRH_HD static constexpr int by(int s) { return (s == 2 || s == 3) ? 1 : 0; }
  MRH_HD static constexpr int vtx(int a, int b) { return b ? (a ? 2 : 3) : (a ? 1 : 0); }

  // coefficient `slot` at a point: the NVRTC policy inlines the expression, the default policy runs the bytecode
  template <int SLOT>
  MRH_HDN static double coef(const KParams& kp, double x, double y, double t) {
    if constexpr (CF::staged) {
      PointState<double> ps;
      ps.x[0] = x; ps.x[1] = y; ps.x[2] = 0.0; ps.time = t; ps.fields = nullptr;
      for (int c = 0; c < 3; ++c) { ps.nrm[c] = 0.0; ps.tng[c] = 0.0; }
      return eval_prog<double>(kp.progs[SLOT], kp.code, kp.cpool, ps);
    } else {
      if constexpr (SLOT == TH_SOURCE) return CF::source(x, y, 0.0, t);
      else if constexpr (SLOT == TH_DIFF) return CF::diff(x, y, 0.0, t);
      else if constexpr (SLOT == TH_CP) return CF::cp_(x, y, 0.0, t);
      else if constexpr (SLOT == TH_RHO) return CF::rho_(x, y, 0.0, t);
      else if constexpr (SLOT == TH_BX) return CF::bx(x, y, 0.0, t);
      else return CF::by(x, y, 0.0, t);
    }
  }

  MRH_HDN static void element(const KParams& kp, int64_t e) {
    int lid[4];
    double r[4], K[NK];
    compute(kp, e, lid, r, K);
    scatter_direct(kp, e, lid, r, K);
  }

  // gather + element residual r (not yet negated) and element matrix K of element e
  MRH_HDN static void compute(const KParams& kp, int64_t e, int (&lid)[4], double (&r)[4], double (&K)[NK]) {
    double X[4][2], us[4], ud[4];
    gather(kp, e, lid, X, us, ud);
    work(kp, X, us, ud, r, K);
  }

#if defined(__CUDACC__)
  // grid-stride loop of the direct-scatter kernel, software-pipelined one element deep: the index rows and the vertex / solution
  // values of the thread's NEXT element are in flight while the current one is computed and scattered (ncu of the plain loop: 3.4
  // long-scoreboard stall cycles per issued instruction at 12 warps per SM - the gather's two dependent load levels were exposed)
  __device__ static void run_direct(const KParams& kp) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= kp.nwork) return;
    int lid[4]; double X[4][2], us[4], ud[4];
    gather(kp, e, lid, X, us, ud);
    for (;;) {
      const int64_t en = e + stride;
      const bool more = en < kp.nwork;
      int lidn[4]; double Xn[4][2], usn[4], udn[4];
      if (more) gather(kp, en, lidn, Xn, usn, udn);
      double r[4], K[NK];
      work(kp, X, us, ud, r, K);
      scatter_direct(kp, e, lid, r, K);
      if (!more) break;
#pragma unroll
      for (int s = 0; s < 4; ++s) { lid[s] = lidn[s]; X[s][0] = Xn[s][0]; X[s][1] = Xn[s][1]; us[s] = usn[s]; ud[s] = udn[s]; }
      e = en;
    }
  }
#endif

  MRH_HDN static void gather(const KParams& kp, int64_t e, int (&lid)[4], double (&X)[4][2], double (&us)[4], double (&ud)[4]) {
    {
      const int4 a = *reinterpret_cast<const int4*>(kp.lids + e * 4);
      lid[0] = a.x; lid[1] = a.y; lid[2] = a.z; lid[3] = a.w;
    }
    {
      const int4 a = *reinterpret_cast<const int4*>(kp.e2n + e * 4);
      const int nd[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
      for (int s = 0; s < 4; ++s) { const double* p = kp.nodes + (int64_t)nd[s] * 2; X[s][0] = p[0]; X[s][1] = p[1]; }
    }
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const double uval = kp.u[lid[s]];
      if (!TRANSIENT) { us[s] = uval; ud[s] = 0.0; }
      else {                                                // Workset::computeSolnTransientSeeded (workset.cpp:594-620)
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

  // element residual r (not yet negated) and element matrix K from the gathered vertices X and solution us / ud
  MRH_HDN static void work(const KParams& kp, const double (&X)[4][2], const double (&us)[4], const double (&ud)[4], double (&r)[4], double (&K)[NK]) {
    const double au = TRANSIENT ? kp.tp.alpha_u : 1.0, at = TRANSIENT ? kp.tp.alpha_t : 0.0;

    // ---- monomial coefficients, index m = i + 2 j for xi^i eta^j ------------------------------------------------
    double C[4][2], A[4];
    {
      double v[3][4];
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int a = 0; a < 2; ++a) { const int sv = vtx(a, b), m = a + 2 * b; v[0][m] = X[sv][0]; v[1][m] = X[sv][1]; v[2][m] = us[sv]; }
#pragma unroll
      for (int d = 0; d < 3; ++d)
#pragma unroll
        for (int bit = 1; bit < 4; bit <<= 1)
#pragma unroll
          for (int m = 0; m < 4; ++m)
            if (!(m & bit)) { const double lo = v[d][m], hi = v[d][m | bit]; v[d][m] = hi + lo; v[d][m | bit] = hi - lo; }
#pragma unroll
      for (int m = 0; m < 4; ++m) { C[m][0] = 0.25 * v[0][m]; C[m][1] = 0.25 * v[1][m]; A[m] = 0.25 * v[2][m]; }
    }

#pragma unroll
    for (int s = 0; s < 4; ++s) r[s] = 0.0;
#pragma unroll
    for (int k = 0; k < NK; ++k) K[k] = 0.0;

    // ---- point loop -------------------------------------------------------------------------------------------------
#if defined(MRH_Q2D_UNROLL)                // tuning experiment: rolled point loop (fewer registers, more resident warps)
#pragma unroll MRH_Q2D_UNROLL
#else
#pragma unroll
#endif
    for (int q = 0; q < 4; ++q) {
      const double g = HexQ1::g;
      const double sx = (q & 1) ? -g : g, sy = (q & 2) ? -g : g;
      const double lx[2] = {0.5 - 0.5 * sx, 0.5 + 0.5 * sx}, ly[2] = {0.5 - 0.5 * sy, 0.5 + 0.5 * sy};
      const double J00 = C[1][0] + C[3][0] * sy, J01 = C[2][0] + C[3][0] * sx;       // d x / d (xi, eta)
      const double J10 = C[1][1] + C[3][1] * sy, J11 = C[2][1] + C[3][1] * sx;       // d y / d (xi, eta)
      const double det = J00 * J11 - J01 * J10;
      const double idet = 1.0 / det;
      const double Ji00 = J11 * idet, Ji01 = -J01 * idet, Ji10 = -J10 * idet, Ji11 = J00 * idet;   // inverse Jacobian
      const double w = fabs(det);                            // reference weights are 1
      double gx[4], gy[4], Nq[4];
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const int a = bx(s), b = by(s);
        const double d0 = (a ? 0.5 : -0.5) * ly[b], d1 = (b ? 0.5 : -0.5) * lx[a];
        Nq[s] = lx[a] * ly[b];
        gx[s] = Ji00 * d0 + Ji10 * d1;                       // grad phi = J^-T grad_ref phi
        gy[s] = Ji01 * d0 + Ji11 * d1;
      }
      const double g0 = A[1] + A[3] * sy, g1 = A[2] + A[3] * sx;
      const double ex = Ji00 * g0 + Ji10 * g1, ey = Ji01 * g0 + Ji11 * g1;
      double edot = 0.0;
      if (TRANSIENT) {
#pragma unroll
        for (int s = 0; s < 4; ++s) edot += ud[s] * Nq[s];
      }
      const double xe = sx * sy;
      const double x = C[0][0] + C[1][0] * sx + C[2][0] * sy + C[3][0] * xe;
      const double y = C[0][1] + C[1][1] * sx + C[2][1] * sy + C[3][1] * xe;
      const double t = kp.tp.time;
      const double srcq = coef<TH_SOURCE>(kp, x, y, t), kapq = coef<TH_DIFF>(kp, x, y, t);
      double rcpq = 0.0, bxq = 0.0, byq = 0.0;
      if (TRANSIENT) rcpq = coef<TH_RHO>(kp, x, y, t) * coef<TH_CP>(kp, x, y, t);
      if (ADV) { bxq = coef<TH_BX>(kp, x, y, t); byq = coef<TH_BY>(kp, x, y, t); }
      const double kw = kapq * w;
      double f0 = -srcq * w;
      if (TRANSIENT) f0 += rcpq * edot * w;
      if (ADV) f0 += (bxq * ex + byq * ey) * w;
      const double fx = kw * ex, fy = kw * ey;
#pragma unroll
      for (int s = 0; s < 4; ++s) r[s] += f0 * Nq[s] + fx * gx[s] + fy * gy[s];
      if (JAC) {
        const double kwa = kw * au;
        const double mw = TRANSIENT ? rcpq * at * w : 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const double ax = kwa * gx[i], ay = kwa * gy[i];
          const double mi = TRANSIENT ? mw * Nq[i] : 0.0;
          if (!ADV) {
#pragma unroll
            for (int j = i; j < 4; ++j) {
              double v = K[tri(i, j)] + ax * gx[j] + ay * gy[j];
              if (TRANSIENT) v += mi * Nq[j];
              K[tri(i, j)] = v;
            }
          } else {
            const double ci = au * w * Nq[i];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              double v = K[i * 4 + j] + ax * gx[j] + ay * gy[j];
              v += ci * (bxq * gx[j] + byq * gy[j]);
              if (TRANSIENT) v += mi * Nq[j];
              K[i * 4 + j] = v;
            }
          }
        }
      }
    }

  }

  MRH_HDN static void scatter_direct(const KParams& kp, int64_t e, const int (&lid)[4], double (&r)[4], double (&K)[NK]) {
    // ---- scatter: one RED per entry through the slot-pattern dictionary --------------------------------------------------
    // The kernel is RED-bound (20 per element).  Consecutive lanes hold consecutive elements; where the next lane's left
    // edge (local nodes 0, 3) is this element's right edge (local nodes 1, 2) its contributions to the shared rows and
    // columns are added here in registers first (6 of 20 REDs saved on structured rows; the LIDs are compared, so any
    // numbering is handled correctly).
    bool skip03 = false;                  // this lane's (0,3) x (0,3) block and r0, r3 were handed to the previous lane
#if defined(__CUDA_ARCH__)
    {
      const unsigned mask = __activemask();
      const int lane = threadIdx.x & 31;
      const int n0 = __shfl_down_sync(mask, lid[0], 1), n3 = __shfl_down_sync(mask, lid[3], 1);
      const bool take = lane < 31 && ((mask >> (lane + 1)) & 1u) && n0 == lid[1] && n3 == lid[2];
      const int gave = __shfl_up_sync(mask, (int)take, 1);
      skip03 = lane > 0 && ((mask >> (lane - 1)) & 1u) && gave != 0;
      const double q0 = __shfl_down_sync(mask, r[0], 1), q3 = __shfl_down_sync(mask, r[3], 1);
      if (take) { r[1] += q0; r[2] += q3; }
      if (JAC) {
        if (!ADV) {
          const double k00 = __shfl_down_sync(mask, K[tri(0, 0)], 1), k03 = __shfl_down_sync(mask, K[tri(0, 3)], 1),
                       k33 = __shfl_down_sync(mask, K[tri(3, 3)], 1);
          if (take) { K[tri(1, 1)] += k00; K[tri(1, 2)] += k03; K[tri(2, 2)] += k33; }
        } else {
          const double k00 = __shfl_down_sync(mask, K[0 * 4 + 0], 1), k03 = __shfl_down_sync(mask, K[0 * 4 + 3], 1),
                       k30 = __shfl_down_sync(mask, K[3 * 4 + 0], 1), k33 = __shfl_down_sync(mask, K[3 * 4 + 3], 1);
          if (take) { K[1 * 4 + 1] += k00; K[1 * 4 + 2] += k03; K[2 * 4 + 1] += k30; K[2 * 4 + 2] += k33; }
        }
      }
    }
#endif
    bool fx_[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      fx_[s] = kp.fixed[lid[s]] != 0;
      if (!fx_[s] && !(skip03 && (s == 0 || s == 3))) atomic_add(kp.res + lid[s], -r[s]);
    }
    if (!JAC) return;
    uint16_t pat[16];
    {
      const uint4* pp = reinterpret_cast<const uint4*>(kp.pat + (size_t)kp.pat_id[e] * 16);
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const uint4 v = pp[k];
        pat[k * 8 + 0] = (uint16_t)(v.x & 0xffff); pat[k * 8 + 1] = (uint16_t)(v.x >> 16);
        pat[k * 8 + 2] = (uint16_t)(v.y & 0xffff); pat[k * 8 + 3] = (uint16_t)(v.y >> 16);
        pat[k * 8 + 4] = (uint16_t)(v.z & 0xffff); pat[k * 8 + 5] = (uint16_t)(v.z >> 16);
        pat[k * 8 + 6] = (uint16_t)(v.w & 0xffff); pat[k * 8 + 7] = (uint16_t)(v.w >> 16);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (fx_[i]) continue;
      double* jrow = kp.jac + kp.rowptr[lid[i]];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const double v = ADV ? K[i * 4 + j] : (i <= j ? K[tri(i, j)] : K[tri(j, i)]);
        if (skip03 && (i == 0 || i == 3) && (j == 0 || j == 3)) continue;
        atomic_add(jrow + pat[i * 4 + j], v);
      }
    }
  }

#if defined(__CUDACC__)
  // ---- tile-staged scatter (device only; same scheme as ThermalHexQ1::tiled, engine_tiles.inl builds the plan with NN = 4) --------
  // A CTA takes tiles of 128 elements (16 x 8 blocks on the structured mesh).  Every thread parks its element matrix and residual in
  // the shared element store (entry-major), then sums 24 positions of the tile's sorted contribution list in registers and stores
  // the run totals into the accumulation array; the CTA flushes that array with REDs whose lanes walk consecutive CRS slots.
  // Against the direct scatter: ~10 instead of 14-20 REDs per element, and 8-9 lanes per row share 32-byte sectors, which is what
  // bounds RED.F64 throughput (profiles/r1_red_microbench.txt: 105 G/s at one lane per sector, 690 G/s coalesced).
  // dynamic shared memory: Ks[(NK + 4) * 128] | zero cell | pad | acc[(max_acc + 2) & ~1] | carry[128] | off[max_e] uint32 | rows[max_r] int
  static constexpr int NE = NK + 4;
  static constexpr int NCP = 24;                            // stored contributions per thread (20 + 4 pads: three uint4 words)
  __host__ __device__ static constexpr size_t tiled_smem_bytes(int max_acc, int max_e, int max_r) {
    return ((size_t)NE * 128 + 2 + (size_t)((max_acc + 2) & ~1) + 128) * sizeof(double) + ((size_t)max_e + (size_t)max_r) * 4 + 16;
  }
  __device__ static void tiled(const KParams& kp, double* sm) {
    static_assert(JAC, "the tile-staged scatter is used for Jacobian assembly only");
    typedef ThermalHexQ1<TRANSIENT, ADV, JAC, CF> Hex;       // (its cp.async helper)
    const TilePlan& tp = kp.tile;
    const int tid = threadIdx.x;
    double* Ks = sm;
    double* acc = sm + NE * 128 + 2;
    double* carry = acc + ((tp.max_acc + 2) & ~1);
    unsigned* s_off = reinterpret_cast<unsigned*>(carry + 128);
    int* s_rows = reinterpret_cast<int*>(s_off + tp.max_e);
    if (tid == 0) Ks[NE * 128] = 0.0;                       // masked contributions and pads read this cell
    for (int64_t t = tp.tile_begin + blockIdx.x; t < tp.tile_end; t += gridDim.x) {
      const int64_t eb = tp.ebeg[t], rb = tp.rptr[t];
      const int E = (int)(tp.eptr[t + 1] - tp.eptr[t]), R = (int)(tp.rptr[t + 1] - rb);
      for (int k = tid; k < E; k += 128) Hex::cp_async4(s_off + k, tp.off + eb + k);
      for (int k = tid; k < R; k += 128) Hex::cp_async4(s_rows + k, tp.rows + rb + k);
      asm volatile("cp.async.commit_group;" ::: "memory");
      const size_t cid = (size_t)tp.cl_id[t] * 128 + tid;
      unsigned w[NCP / 2];
      {
        const uint4* cl = reinterpret_cast<const uint4*>(tp.clist + cid * NCP);
#pragma unroll
        for (int i = 0; i < NCP / 8; ++i) { const uint4 v = __ldg(cl + i); w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w; }
      }
      const int loc0 = tp.cbase[cid];
      const int64_t e0 = t * 128 + tid;
      if (e0 < kp.nwork) {
        int lid[4];
        double r[4], K[NK];
        compute(kp, (int64_t)tp.perm[e0], lid, r, K);
#pragma unroll
        for (int k = 0; k < NK; ++k) Ks[k * 128 + tid] = K[k];
#pragma unroll
        for (int s = 0; s < 4; ++s) Ks[(NK + s) * 128 + tid] = -r[s];
      }
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      const bool cont = (w[0] & 0x8000u) == 0;              // my first run started in the previous thread's chunk
      __syncthreads();                                      // element store and list copies complete
      {
        double sum = 0.0;
        int loc = loc0;
        double* dst = cont ? carry + tid : acc + loc;
#pragma unroll
        for (int g = 0; g < NCP; g += 8) {
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
      __syncthreads();                                      // run totals stored
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
      __syncthreads();                                      // lists, element store and accumulation array free for the next tile
    }
  }
#endif
};

}  // namespace mrh
