// ORACLE (test infrastructure, NOT product code).
// Reference-element data and reference->physical maps, restating the Intrepid2 calls made at
//   /root/reference/src/interfaces/discretizationInterface.cpp:346-478 (getBasis, getQuadrature)
//   :483-665 (setReferenceData), :732-776 (getPhysicalIntegrationData),
//   :898-981 (getPhysicalVolumetricBasis), :1608-1789 (getPhysicalBoundaryIntegrationData),
//   :1810-1887 (getPhysicalBoundaryBasis).
// Intrepid2 itself is absent (Trilinos, unpinned).  Pinned by regression/discretization/HGRAD/mrhyde.gold:
// tensor dof order (x fastest), 2-point Gauss order (abscissae DESCENDING, x fastest), J^-T scaling.
// 3-point Gauss ordering and all side data are unpinned (summation-order effects only).
#pragma once
#include <cmath>
#include <stdexcept>
#include <vector>

namespace orc {

// Gauss-Legendre rule with n points on [-1,1], abscissae descending.
inline void gauss_legendre(int n, std::vector<double>& x, std::vector<double>& w) {
  x.assign(n, 0.0); w.assign(n, 0.0);
  for (int i = 0; i < n; ++i) {
    double z = std::cos(M_PI * (i + 0.75) / (n + 0.5));   // descending initial guesses
    double pp = 0.0;
    for (int it = 0; it < 100; ++it) {
      double p1 = 1.0, p2 = 0.0;
      for (int j = 0; j < n; ++j) { const double p3 = p2; p2 = p1; p1 = ((2.0 * j + 1.0) * z * p2 - j * p3) / (j + 1.0); }
      pp = n * (z * p1 - p2) / (z * z - 1.0);
      const double z1 = z; z = z1 - p1 / pp;
      if (std::fabs(z - z1) < 1e-16) break;
    }
    // one more evaluation for the weight at the converged root
    double p1 = 1.0, p2 = 0.0;
    for (int j = 0; j < n; ++j) { const double p3 = p2; p2 = p1; p1 = ((2.0 * j + 1.0) * z * p2 - j * p3) / (j + 1.0); }
    pp = n * (z * p1 - p2) / (z * z - 1.0);
    x[i] = z; w[i] = 2.0 / ((1.0 - z * z) * pp * pp);
  }
  if (n % 2 == 1) x[n / 2] = 0.0;
  if (n == 2) { x[0] = 1.0 / std::sqrt(3.0); x[1] = -x[0]; w[0] = w[1] = 1.0; }
}

// 1-D Lagrange basis of order p on equispaced nodes of [-1,1]
inline void lagrange1d(int p, double xi, double* val, double* der) {
  const int m = p + 1;
  std::vector<double> nd(m);
  for (int i = 0; i < m; ++i) nd[i] = -1.0 + 2.0 * i / p;
  for (int i = 0; i < m; ++i) {
    double v = 1.0, dsum = 0.0;
    for (int j = 0; j < m; ++j) if (j != i) v *= (xi - nd[j]) / (nd[i] - nd[j]);
    for (int k = 0; k < m; ++k) if (k != i) {
      double t = 1.0 / (nd[i] - nd[k]);
      for (int j = 0; j < m; ++j) if (j != i && j != k) t *= (xi - nd[j]) / (nd[i] - nd[j]);
      dsum += t;
    }
    val[i] = v; der[i] = dsum;
  }
}

// HGRAD_{QUAD,HEX}_Cn: value and reference gradient of every basis function at one point.
inline void hgrad_eval(int dim, int p, const double* xi, double* val /*nb*/, double* grad /*nb*dim*/) {
  const int m = p + 1;
  double v[3][8], d[3][8];
  for (int c = 0; c < dim; ++c) lagrange1d(p, xi[c], v[c], d[c]);
  const int nb = (dim == 3) ? m * m * m : m * m;
  for (int o = 0; o < nb; ++o) {
    const int a = o % m, b = (o / m) % m, c = (dim == 3) ? o / (m * m) : 0;
    if (dim == 2) {
      val[o] = v[0][a] * v[1][b];
      grad[o * 2 + 0] = d[0][a] * v[1][b];
      grad[o * 2 + 1] = v[0][a] * d[1][b];
    } else {
      val[o] = v[0][a] * v[1][b] * v[2][c];
      grad[o * 3 + 0] = d[0][a] * v[1][b] * v[2][c];
      grad[o * 3 + 1] = v[0][a] * d[1][b] * v[2][c];
      grad[o * 3 + 2] = v[0][a] * v[1][b] * d[2][c];
    }
  }
}

// vertex (geometry) shape functions in shards vertex order
inline void vertex_eval(int dim, const double* xi, double* val, double* grad) {
  double tv[8], tg[24];
  hgrad_eval(dim, 1, xi, tv, tg);
  static const int t2s_q[4] = {0, 1, 3, 2}, t2s_h[8] = {0, 1, 3, 2, 4, 5, 7, 6};
  const int nv = (dim == 3) ? 8 : 4;
  for (int o = 0; o < nv; ++o) {
    const int s = (dim == 3) ? t2s_h[o] : t2s_q[o];
    val[s] = tv[o];
    for (int c = 0; c < dim; ++c) grad[s * dim + c] = tg[o * dim + c];
  }
}

struct RefData {
  int dim = 3, p = 1, nb = 8, nv = 8, nq = 8, nqs = 4, nsides = 6;
  std::vector<double> ip, wts;                 // (nq,dim), (nq)
  std::vector<double> B, G;                    // (nb,nq), (nb,nq,dim)   field basis
  std::vector<double> VB, VG;                  // (nv,nq), (nv,nq,dim)   vertex basis
  // per side s: side ips mapped to the cell (nqs,dim), weights, field/vertex basis there, tangents
  std::vector<std::vector<double>> sip, sB, sG, sVB, sVG;
  std::vector<double> swts;
  double tanU[6][3], tanV[6][3];               // 3-D face tangents; 2-D edge tangent in tanU

  void build(int dim_, int p_, int quad_degree, int side_degree) {
    dim = dim_; p = p_;
    const int m = p + 1;
    nb = (dim == 3) ? m * m * m : m * m;
    nv = (dim == 3) ? 8 : 4;
    nsides = (dim == 3) ? 6 : 4;
    const int n1 = (quad_degree + 2) / 2;          // ceil((deg+1)/2)
    std::vector<double> gx, gw; gauss_legendre(n1, gx, gw);
    nq = (dim == 3) ? n1 * n1 * n1 : n1 * n1;
    ip.resize((size_t)nq * dim); wts.resize(nq);
    for (int q = 0; q < nq; ++q) {
      const int a = q % n1, b = (q / n1) % n1, c = (dim == 3) ? q / (n1 * n1) : 0;
      ip[q * dim + 0] = gx[a]; ip[q * dim + 1] = gx[b];
      wts[q] = gw[a] * gw[b];
      if (dim == 3) { ip[q * dim + 2] = gx[c]; wts[q] *= gw[c]; }
    }
    tabulate(ip, nq, B, G, VB, VG);
    // side rule
    const int s1 = (side_degree + 2) / 2;
    std::vector<double> sx, sw; gauss_legendre(s1, sx, sw);
    nqs = (dim == 3) ? s1 * s1 : s1;
    swts.resize(nqs);
    std::vector<double> uv((size_t)nqs * 2, 0.0);
    for (int q = 0; q < nqs; ++q) {
      const int a = q % s1, b = q / s1;
      uv[q * 2 + 0] = sx[a]; swts[q] = sw[a];
      if (dim == 3) { uv[q * 2 + 1] = sx[b]; swts[q] *= sw[b]; }
    }
    sip.assign(nsides, {}); sB.assign(nsides, {}); sG.assign(nsides, {}); sVB.assign(nsides, {}); sVG.assign(nsides, {});
    for (int s = 0; s < nsides; ++s) {
      sip[s].resize((size_t)nqs * dim);
      for (int q = 0; q < nqs; ++q) {
        const double u = uv[q * 2], v = uv[q * 2 + 1];
        double* x = &sip[s][(size_t)q * dim];
        if (dim == 2) {   // Quadrilateral_4 edge parametrisations
          switch (s) { case 0: x[0] = u; x[1] = -1; break; case 1: x[0] = 1; x[1] = u; break;
                       case 2: x[0] = -u; x[1] = 1; break; default: x[0] = -1; x[1] = -u; }
        } else {          // Hexahedron_8 face parametrisations
          switch (s) { case 0: x[0] = u; x[1] = -1; x[2] = v; break; case 1: x[0] = 1; x[1] = u; x[2] = v; break;
                       case 2: x[0] = -u; x[1] = 1; x[2] = v; break; case 3: x[0] = -1; x[1] = -u; x[2] = v; break;
                       case 4: x[0] = v; x[1] = u; x[2] = -1; break; default: x[0] = u; x[1] = v; x[2] = 1; }
        }
      }
      tabulate(sip[s], nqs, sB[s], sG[s], sVB[s], sVG[s]);
    }
    static const double e2[4][2] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}};
    static const double fU[6][3] = {{1,0,0},{0,1,0},{-1,0,0},{0,-1,0},{0,1,0},{1,0,0}};
    static const double fV[6][3] = {{0,0,1},{0,0,1},{0,0,1},{0,0,1},{1,0,0},{0,1,0}};
    for (int s = 0; s < nsides; ++s)
      for (int c = 0; c < 3; ++c) {
        if (dim == 2) { tanU[s][c] = (c < 2) ? e2[s][c] : 0.0; tanV[s][c] = 0.0; }
        else { tanU[s][c] = fU[s][c]; tanV[s][c] = fV[s][c]; }
      }
  }

  void tabulate(const std::vector<double>& pts, int npt, std::vector<double>& b, std::vector<double>& g,
                std::vector<double>& vb, std::vector<double>& vg) const {
    b.resize((size_t)nb * npt); g.resize((size_t)nb * npt * dim);
    vb.resize((size_t)nv * npt); vg.resize((size_t)nv * npt * dim);
    std::vector<double> tv(nb), tg((size_t)nb * dim), wv(nv), wg((size_t)nv * dim);
    for (int q = 0; q < npt; ++q) {
      hgrad_eval(dim, p, &pts[(size_t)q * dim], tv.data(), tg.data());
      vertex_eval(dim, &pts[(size_t)q * dim], wv.data(), wg.data());
      for (int o = 0; o < nb; ++o) {
        b[(size_t)o * npt + q] = tv[o];
        for (int c = 0; c < dim; ++c) g[((size_t)o * npt + q) * dim + c] = tg[o * dim + c];
      }
      for (int o = 0; o < nv; ++o) {
        vb[(size_t)o * npt + q] = wv[o];
        for (int c = 0; c < dim; ++c) vg[((size_t)o * npt + q) * dim + c] = wg[o * dim + c];
      }
    }
  }
};

// ---- per-element physical data ------------------------------------------------------------------
inline double det_inv(int dim, const double* J, double* Ji) {
  if (dim == 2) {
    const double det = J[0] * J[3] - J[1] * J[2];
    Ji[0] = J[3] / det; Ji[1] = -J[1] / det; Ji[2] = -J[2] / det; Ji[3] = J[0] / det;
    return det;
  }
  const double c0 = J[4] * J[8] - J[5] * J[7], c1 = J[5] * J[6] - J[3] * J[8], c2 = J[3] * J[7] - J[4] * J[6];
  const double det = J[0] * c0 + J[1] * c1 + J[2] * c2;
  Ji[0] = c0 / det; Ji[1] = (J[2] * J[7] - J[1] * J[8]) / det; Ji[2] = (J[1] * J[5] - J[2] * J[4]) / det;
  Ji[3] = c1 / det; Ji[4] = (J[0] * J[8] - J[2] * J[6]) / det; Ji[5] = (J[2] * J[3] - J[0] * J[5]) / det;
  Ji[6] = c2 / det; Ji[7] = (J[1] * J[6] - J[0] * J[7]) / det; Ji[8] = (J[0] * J[4] - J[1] * J[3]) / det;
  return det;
}

// J_{ij} = sum_a x_{a,i} dN_a/dxi_j at point q of a tabulated vertex-gradient table (nv,npt,dim)
inline void jacobian_at(int dim, int nv, int npt, int q, const double* xv /*nv,dim*/, const double* VG, double* J) {
  for (int i = 0; i < dim * dim; ++i) J[i] = 0.0;
  for (int a = 0; a < nv; ++a)
    for (int i = 0; i < dim; ++i)
      for (int j = 0; j < dim; ++j) J[i * dim + j] += xv[a * dim + i] * VG[((size_t)a * npt + q) * dim + j];
}

}  // namespace orc
