// Host-side reference-element tables uploaded once per engine: tensor Gauss-Legendre rules, HGRAD_{QUAD,HEX}_Cn
// Lagrange bases on equispaced points and the vertex (geometry) basis, at the volume rule and at each local
// side's rule.  Stands in for the Intrepid2 objects the reference creates in
// /root/reference/src/interfaces/discretizationInterface.cpp:346-478 (getBasis/getQuadrature) and the reference
// data it tabulates at :483-665.  Conventions pinned by regression/discretization/HGRAD/mrhyde.gold:
// basis ordinal = a + (p+1) b + (p+1)^2 c, Gauss abscissae descending, point index x fastest.
#pragma once
#include <cmath>
#include <stdexcept>
#include <vector>

namespace mrh {

struct PointTables {            // point-major layout consumed by the kernels (params.h RefTab)
  int nq = 0;
  std::vector<double> B, G, VB, VG, W, pts;
  std::vector<double> FR;         // [nq][ceil(nb / 8)][32]: (B | G) in DMMA A-fragment order, lane = 4 (j % 8) + a (a = 0: value, 1 + m: d/dxi_m)
  double tanU[3] = {0, 0, 0}, tanV[3] = {0, 0, 0};
};

class RefElement {
 public:
  int dim = 3, p = 1, nb = 8, nv = 8, nsides = 6;
  PointTables vol;
  std::vector<PointTables> side;

  static void gauss(int n, std::vector<double>& x, std::vector<double>& w) {
    x.assign(n, 0.0); w.assign(n, 0.0);
    for (int i = 0; i < n; ++i) {
      double z = std::cos(M_PI * (i + 0.75) / (n + 0.5)), pp = 1.0, p1 = 1.0;
      for (int it = 0; it < 100; ++it) {
        legendre(n, z, p1, pp);
        const double z1 = z; z = z1 - p1 / pp;
        if (std::fabs(z - z1) < 1e-16) break;
      }
      legendre(n, z, p1, pp);
      x[i] = z; w[i] = 2.0 / ((1.0 - z * z) * pp * pp);
    }
    if (n % 2 == 1) x[n / 2] = 0.0;
    if (n == 2) { x[0] = 1.0 / std::sqrt(3.0); x[1] = -x[0]; w[0] = w[1] = 1.0; }
  }

  void build(int dim_, int order, int quad_degree, int side_degree) {
    dim = dim_; p = order;
    if (p < 1 || p > 2) throw std::runtime_error("only HGRAD order 1 and 2 are built");
    const int m = p + 1;
    nb = (dim == 3) ? m * m * m : m * m; nv = (dim == 3) ? 8 : 4; nsides = (dim == 3) ? 6 : 4;
    const int n1 = (quad_degree + 2) / 2;
    std::vector<double> gx, gw; gauss(n1, gx, gw);
    vol.nq = (dim == 3) ? n1 * n1 * n1 : n1 * n1;
    vol.pts.assign((size_t)vol.nq * dim, 0.0); vol.W.assign(vol.nq, 0.0);
    for (int q = 0; q < vol.nq; ++q) {
      const int a = q % n1, b = (q / n1) % n1, c = (dim == 3) ? q / (n1 * n1) : 0;
      vol.pts[q * dim] = gx[a]; vol.pts[q * dim + 1] = gx[b]; vol.W[q] = gw[a] * gw[b];
      if (dim == 3) { vol.pts[q * dim + 2] = gx[c]; vol.W[q] *= gw[c]; }
    }
    tabulate(vol);
    const int s1 = (side_degree + 2) / 2;
    std::vector<double> sx, sw; gauss(s1, sx, sw);
    const int nqs = (dim == 3) ? s1 * s1 : s1;
    static const double e2[4][2] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}};
    static const double fU[6][3] = {{1,0,0},{0,1,0},{-1,0,0},{0,-1,0},{0,1,0},{1,0,0}};
    static const double fV[6][3] = {{0,0,1},{0,0,1},{0,0,1},{0,0,1},{1,0,0},{0,1,0}};
    side.assign(nsides, PointTables());
    for (int s = 0; s < nsides; ++s) {
      PointTables& S = side[s];
      S.nq = nqs; S.pts.assign((size_t)nqs * dim, 0.0); S.W.assign(nqs, 0.0);
      for (int q = 0; q < nqs; ++q) {
        const int a = q % s1, b = q / s1;
        const double u = sx[a], v = (dim == 3) ? sx[b] : 0.0;
        S.W[q] = sw[a] * ((dim == 3) ? sw[b] : 1.0);
        double* x = &S.pts[(size_t)q * dim];
        if (dim == 2) {       // shards Quadrilateral_4 edge parametrisations
          switch (s) { case 0: x[0] = u; x[1] = -1; break; case 1: x[0] = 1; x[1] = u; break;
                       case 2: x[0] = -u; x[1] = 1; break; default: x[0] = -1; x[1] = -u; }
        } else {              // shards Hexahedron_8 face parametrisations
          switch (s) { case 0: x[0] = u; x[1] = -1; x[2] = v; break; case 1: x[0] = 1; x[1] = u; x[2] = v; break;
                       case 2: x[0] = -u; x[1] = 1; x[2] = v; break; case 3: x[0] = -1; x[1] = -u; x[2] = v; break;
                       case 4: x[0] = v; x[1] = u; x[2] = -1; break; default: x[0] = u; x[1] = v; x[2] = 1; }
        }
      }
      tabulate(S);
      for (int c = 0; c < 3; ++c) {
        if (dim == 2) { S.tanU[c] = (c < 2) ? e2[s][c] : 0.0; S.tanV[c] = 0.0; }
        else { S.tanU[c] = fU[s][c]; S.tanV[c] = fV[s][c]; }
      }
    }
  }

 private:
  static void legendre(int n, double z, double& p1, double& pp) {
    p1 = 1.0; double p2 = 0.0;
    for (int j = 0; j < n; ++j) { const double p3 = p2; p2 = p1; p1 = ((2.0 * j + 1.0) * z * p2 - j * p3) / (j + 1.0); }
    pp = n * (z * p1 - p2) / (z * z - 1.0);
  }
  static void lagrange(int p, double xi, double* v, double* d) {
    const int m = p + 1; double nd[4];
    for (int i = 0; i < m; ++i) nd[i] = -1.0 + 2.0 * i / p;
    for (int i = 0; i < m; ++i) {
      double val = 1.0, ds = 0.0;
      for (int j = 0; j < m; ++j) if (j != i) val *= (xi - nd[j]) / (nd[i] - nd[j]);
      for (int k = 0; k < m; ++k) if (k != i) {
        double t = 1.0 / (nd[i] - nd[k]);
        for (int j = 0; j < m; ++j) if (j != i && j != k) t *= (xi - nd[j]) / (nd[i] - nd[j]);
        ds += t;
      }
      v[i] = val; d[i] = ds;
    }
  }
  void tensor(int order, const double* xi, double* val, double* grad) const {
    const int m = order + 1; double v[3][4], d[3][4];
    for (int c = 0; c < dim; ++c) lagrange(order, xi[c], v[c], d[c]);
    const int n = (dim == 3) ? m * m * m : m * m;
    for (int o = 0; o < n; ++o) {
      const int a = o % m, b = (o / m) % m, c = (dim == 3) ? o / (m * m) : 0;
      if (dim == 2) { val[o] = v[0][a] * v[1][b]; grad[o * 2] = d[0][a] * v[1][b]; grad[o * 2 + 1] = v[0][a] * d[1][b]; }
      else {
        val[o] = v[0][a] * v[1][b] * v[2][c];
        grad[o * 3] = d[0][a] * v[1][b] * v[2][c]; grad[o * 3 + 1] = v[0][a] * d[1][b] * v[2][c]; grad[o * 3 + 2] = v[0][a] * v[1][b] * d[2][c];
      }
    }
  }
  void tabulate(PointTables& T) const {
    static const int t2s[8] = {0, 1, 3, 2, 4, 5, 7, 6};     // tensor vertex ordinal -> shards vertex
    T.B.assign((size_t)T.nq * nb, 0.0); T.G.assign((size_t)T.nq * nb * dim, 0.0);
    T.VB.assign((size_t)T.nq * nv, 0.0); T.VG.assign((size_t)T.nq * nv * dim, 0.0);
    std::vector<double> tv(27), tg(81);
    const int nrt = (nb + 7) / 8;
    T.FR.assign((size_t)T.nq * nrt * 32, 0.0);
    for (int q = 0; q < T.nq; ++q) {
      tensor(p, &T.pts[(size_t)q * dim], tv.data(), tg.data());
      for (int o = 0; o < nb; ++o) { T.B[(size_t)q * nb + o] = tv[o]; for (int c = 0; c < dim; ++c) T.G[((size_t)q * nb + o) * dim + c] = tg[o * dim + c]; }
      for (int o = 0; o < nb; ++o) {
        double* fr = &T.FR[((size_t)q * nrt + o / 8) * 32 + (o % 8) * 4];
        fr[0] = tv[o]; for (int c = 0; c < dim; ++c) fr[1 + c] = tg[o * dim + c];
      }
      tensor(1, &T.pts[(size_t)q * dim], tv.data(), tg.data());
      for (int o = 0; o < nv; ++o) { const int s = t2s[o]; T.VB[(size_t)q * nv + s] = tv[o]; for (int c = 0; c < dim; ++c) T.VG[((size_t)q * nv + s) * dim + c] = tg[o * dim + c]; }
    }
  }
};

}  // namespace mrh
