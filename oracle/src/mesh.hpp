// ORACLE (test infrastructure, NOT product code).
// Structured brick mesh + DOF numbering + CRS graph, restating what the reference obtains from
// Panzer-STK / Panzer DOFManager / Tpetra (third party, absent here, unpinned version):
//   mesh numbering  : panzer_stk::SquareQuadMeshFactory / CubeHexMeshFactory, called at
//                     /root/reference/src/interfaces/meshInterface.cpp:128-171; the in-repo analogue is
//                     /root/reference/src/tools/simplemeshmanager.hpp:892-977 (node i+j(NX+1), cell i+j*NX)
//   DOF numbering   : /root/reference/src/interfaces/discretizationInterface.cpp:2324-2418 (buildDOFManagers)
//                     element slots = entities (vertices, edges, faces, cell; shards order) x fields interleaved;
//                     LIDs = first-encounter order over elements, then slots.  The Q1 quad offsets {0,1,3,2}
//                     are pinned by discretizationInterface.cpp:302.
//   fixed (DBC) dofs: discretizationInterface.cpp:2744-2822 (closure of Dirichlet sides, sorted unique)
//   CRS graph       : /root/reference/src/interfaces/linearAlgebraInterface.cpp:186-234
//                     (insertGlobalIndices(gid, elemGIDs) for every element row, fillComplete => sorted unique cols)
// Parity of the third-party conventions is UNPINNED by the reference's own tests.
#pragma once
#include <algorithm>
#include <array>
#include <cstdint>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace orc {

struct SideSet {
  std::string name;
  int local_side;                 // shards side ordinal shared by all entries of a brick side set
  std::vector<int> elems;         // element ids, ascending
};

struct BrickMesh {
  int dim = 3;
  int N[3] = {1, 1, 1};
  double lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
  int nvert = 8;                           // vertices per element
  int64_t nelem = 0, nnode = 0;
  std::vector<double> nodes;               // (nnode, dim)
  std::vector<int> elem2node;              // (nelem, nvert), shards vertex order
  std::vector<SideSet> sides;              // alphabetical, like an stk part map

  void build(int dim_, const int* N_, const double* lo_, const double* hi_) {
    dim = dim_;
    for (int d = 0; d < 3; ++d) { N[d] = (d < dim) ? N_[d] : 1; lo[d] = lo_[d]; hi[d] = hi_[d]; }
    const int nx = N[0] + 1, ny = N[1] + 1, nz = (dim == 3) ? N[2] + 1 : 1;
    nvert = (dim == 3) ? 8 : 4;
    nnode = (int64_t)nx * ny * nz;
    nelem = (int64_t)N[0] * N[1] * ((dim == 3) ? N[2] : 1);
    nodes.resize(nnode * dim);
    for (int k = 0; k < nz; ++k)
      for (int j = 0; j < ny; ++j)
        for (int i = 0; i < nx; ++i) {
          const int64_t id = i + (int64_t)nx * (j + (int64_t)ny * k);
          const int ijk[3] = {i, j, k};
          for (int d = 0; d < dim; ++d)
            nodes[id * dim + d] = lo[d] + (hi[d] - lo[d]) * ((double)ijk[d] / (double)N[d]);
        }
    elem2node.resize(nelem * nvert);
    const int kz = (dim == 3) ? N[2] : 1;
    for (int k = 0; k < kz; ++k)
      for (int j = 0; j < N[1]; ++j)
        for (int i = 0; i < N[0]; ++i) {
          const int64_t e = i + (int64_t)N[0] * (j + (int64_t)N[1] * k);
          auto nid = [&](int a, int b, int c) { return (int)((i + a) + (int64_t)nx * ((j + b) + (int64_t)ny * (k + c))); };
          int* v = &elem2node[e * nvert];
          v[0] = nid(0, 0, 0); v[1] = nid(1, 0, 0); v[2] = nid(1, 1, 0); v[3] = nid(0, 1, 0);
          if (dim == 3) { v[4] = nid(0, 0, 1); v[5] = nid(1, 0, 1); v[6] = nid(1, 1, 1); v[7] = nid(0, 1, 1); }
        }
    // side sets.  2-D (edges): bottom=0, right=1, top=2, left=3.
    // 3-D (faces), panzer_stk::CubeHexMeshFactory names: bottom(y-)=0, right(x+)=1, top(y+)=2, left(x-)=3, back(z-)=4, front(z+)=5
    // (bottom/top are the y planes: regression/maxwell/Multiblock/input_mesh.yaml pairs 'top;bottom' with periodic 'xz-all').
    sides.clear();
    auto add = [&](const char* nm, int ls, int axis, int at_hi) {
      SideSet s; s.name = nm; s.local_side = ls;
      for (int k = 0; k < kz; ++k)
        for (int j = 0; j < N[1]; ++j)
          for (int i = 0; i < N[0]; ++i) {
            const int ijk[3] = {i, j, k};
            if (ijk[axis] == (at_hi ? N[axis] - 1 : 0)) s.elems.push_back((int)(i + (int64_t)N[0] * (j + (int64_t)N[1] * k)));
          }
      sides.push_back(s);
    };
    if (dim == 2) { add("bottom", 0, 1, 0); add("left", 3, 0, 0); add("right", 1, 0, 1); add("top", 2, 1, 1); }
    else { add("back", 4, 2, 0); add("bottom", 0, 1, 0); add("front", 5, 2, 1); add("left", 3, 0, 0); add("right", 1, 0, 1); add("top", 2, 1, 1); }
  }
  int side_index(const std::string& nm) const {
    for (size_t i = 0; i < sides.size(); ++i) if (sides[i].name == nm) return (int)i;
    return -1;
  }
};

// ---------------------------------------------------------------------------------------------
// HGRAD_{QUAD,HEX}_Cn local layout: basis ordinal = a + (p+1) b + (p+1)^2 c (x fastest).
// slot_of(): position of that basis function in Panzer's entity-ordered element GID list.
// ---------------------------------------------------------------------------------------------
struct LocalLayout {
  int dim = 3, p = 1, nb = 8;            // nb = (p+1)^dim
  std::vector<int> slot;                  // basis ordinal -> entity slot
  std::vector<std::array<int, 3>> tix;    // basis ordinal -> (a,b,c)

  static int vertex_ord(int dim, int ha, int hb, int hc) {   // ha.. in {0,1}
    const int q = (hb == 0) ? ha : 3 - ha;                    // 0,1 / 3,2
    return (dim == 3) ? q + 4 * hc : q;
  }
  void build(int dim_, int p_) {
    dim = dim_; p = p_;
    if (p != 1 && p != 2) throw std::runtime_error("oracle: only HGRAD order 1 and 2 are restated");
    const int m = p + 1;
    nb = (dim == 3) ? m * m * m : m * m;
    slot.assign(nb, -1); tix.resize(nb);
    static const int hexedge[12][2] = {{0,1},{1,2},{2,3},{3,0},{4,5},{5,6},{6,7},{7,4},{0,4},{1,5},{2,6},{3,7}};
    static const int hexface[6][4] = {{0,1,5,4},{1,2,6,5},{2,3,7,6},{0,4,7,3},{0,3,2,1},{4,5,6,7}};
    static const int quadedge[4][2] = {{0,1},{1,2},{2,3},{3,0}};
    const int nv = (dim == 3) ? 8 : 4, ne = (dim == 3) ? 12 : 4, nf = (dim == 3) ? 6 : 0;
    for (int o = 0; o < nb; ++o) {
      const int a = o % m, b = (o / m) % m, c = (dim == 3) ? o / (m * m) : 0;
      tix[o] = {a, b, c};
      if (p == 1) { slot[o] = vertex_ord(dim, a, b, c); continue; }
      const int idx[3] = {a, b, c};
      int nmid = 0; for (int d = 0; d < dim; ++d) if (idx[d] == 1) ++nmid;
      // vertices spanned by this entity
      std::vector<int> vs;
      for (int hc = 0; hc < ((dim == 3) ? 2 : 1); ++hc)
        for (int hb = 0; hb < 2; ++hb)
          for (int ha = 0; ha < 2; ++ha) {
            const int h[3] = {ha, hb, hc}; bool ok = true;
            for (int d = 0; d < dim; ++d) if (idx[d] != 1 && idx[d] != 2 * h[d]) ok = false;
            if (ok) vs.push_back(vertex_ord(dim, ha, hb, hc));
          }
      std::sort(vs.begin(), vs.end());
      if (nmid == 0) slot[o] = vs[0];
      else if (nmid == dim) slot[o] = nv + ne + nf;
      else if (nmid == 1) {
        for (int e = 0; e < ne; ++e) {
          std::vector<int> ev = (dim == 3) ? std::vector<int>{hexedge[e][0], hexedge[e][1]} : std::vector<int>{quadedge[e][0], quadedge[e][1]};
          std::sort(ev.begin(), ev.end());
          if (ev == vs) slot[o] = nv + e;
        }
      } else {  // face of a hex
        for (int f = 0; f < nf; ++f) {
          std::vector<int> fv(hexface[f], hexface[f] + 4); std::sort(fv.begin(), fv.end());
          if (fv == vs) slot[o] = nv + ne + f;
        }
      }
      if (slot[o] < 0) throw std::runtime_error("oracle: slot lookup failed");
    }
  }
  // basis ordinals lying on local side `ls` (closure), ascending ordinal
  std::vector<int> side_dofs(int ls) const {
    int axis, at_hi;
    if (dim == 2) { const int ax[4] = {1, 0, 1, 0}, hi[4] = {0, 1, 1, 0}; axis = ax[ls]; at_hi = hi[ls]; }
    else { const int ax[6] = {1, 0, 1, 0, 2, 2}, hi[6] = {0, 1, 1, 0, 0, 1}; axis = ax[ls]; at_hi = hi[ls]; }
    std::vector<int> r;
    for (int o = 0; o < nb; ++o) if (tix[o][axis] == (at_hi ? p : 0)) r.push_back(o);
    return r;
  }
};

struct DofMap {
  int nvar = 1, nb = 8, ndof_elem = 8;
  int64_t nlocal = 0;
  std::vector<int> lids;      // (nelem, ndof_elem)  dof_lids
  std::vector<int> offsets;   // (nvar, nb)          offsets[var][basis dof] -> element slot
  std::vector<double> dofxyz; // (nlocal, dim)  coordinates of the DOF's node (for synthetic states)
  std::vector<int> dofvar;    // (nlocal)       variable index of each local dof

  void build(const BrickMesh& M, const LocalLayout& L, int nvar_) {
    nvar = nvar_; nb = L.nb; ndof_elem = nvar * nb;
    offsets.resize((size_t)nvar * nb);
    for (int f = 0; f < nvar; ++f) for (int o = 0; o < nb; ++o) offsets[(size_t)f * nb + o] = L.slot[o] * nvar + f;
    std::vector<int> ord_of_slot(nb);
    for (int o = 0; o < nb; ++o) ord_of_slot[L.slot[o]] = o;
    const int p = L.p;
    const int64_t gx = (int64_t)p * M.N[0] + 1, gy = (int64_t)p * M.N[1] + 1, gz = (M.dim == 3) ? (int64_t)p * M.N[2] + 1 : 1;
    std::vector<int> first((size_t)(gx * gy * gz), -1);   // refined-grid node -> LID of its field 0
    lids.resize((size_t)M.nelem * ndof_elem);
    int next = 0;
    const int kz = (M.dim == 3) ? M.N[2] : 1;
    for (int k = 0; k < kz; ++k)
      for (int j = 0; j < M.N[1]; ++j)
        for (int i = 0; i < M.N[0]; ++i) {
          const int64_t e = i + (int64_t)M.N[0] * (j + (int64_t)M.N[1] * k);
          for (int s = 0; s < nb; ++s) {
            const auto& t = L.tix[ord_of_slot[s]];
            const int64_t key = (p * i + t[0]) + gx * ((p * j + t[1]) + gy * (int64_t)(p * k + t[2]));
            if (first[key] < 0) {
              first[key] = next; next += nvar;
              for (int f = 0; f < nvar; ++f) {
                dofvar.push_back(f);
                const int g[3] = {p * i + t[0], p * j + t[1], p * k + t[2]};
                for (int d = 0; d < M.dim; ++d) dofxyz.push_back(M.lo[d] + (M.hi[d] - M.lo[d]) * ((double)g[d] / (double)(p * M.N[d])));
              }
            }
            for (int f = 0; f < nvar; ++f) lids[(size_t)e * ndof_elem + s * nvar + f] = first[key] + f;
          }
        }
    nlocal = next;
  }
};

struct CrsGraph {
  int64_t nrows = 0;
  std::vector<int64_t> rowptr;
  std::vector<int> colind;
  void build(const DofMap& D, int64_t nelem) {
    nrows = D.nlocal;
    std::vector<std::vector<int>> rows((size_t)nrows);
    const int n = D.ndof_elem;
    for (int64_t e = 0; e < nelem; ++e) {
      const int* l = &D.lids[(size_t)e * n];
      for (int i = 0; i < n; ++i) { auto& r = rows[l[i]]; r.insert(r.end(), l, l + n); }
    }
    rowptr.assign((size_t)nrows + 1, 0);
    for (int64_t r = 0; r < nrows; ++r) {
      auto& v = rows[r]; std::sort(v.begin(), v.end()); v.erase(std::unique(v.begin(), v.end()), v.end());
      rowptr[r + 1] = rowptr[r] + (int64_t)v.size();
    }
    colind.resize((size_t)rowptr[nrows]);
    for (int64_t r = 0; r < nrows; ++r) std::copy(rows[r].begin(), rows[r].end(), colind.begin() + rowptr[r]);
  }
  int64_t find(int row, int col) const {
    auto b = colind.begin() + rowptr[row], e = colind.begin() + rowptr[row + 1];
    auto it = std::lower_bound(b, e, col);
    if (it == e || *it != col) return -1;
    return it - colind.begin();
  }
};

}  // namespace orc
