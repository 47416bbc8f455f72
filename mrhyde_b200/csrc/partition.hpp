// Host side of the multi-GPU path: element-block partition of the structured brick and the plan of the additive
// interface exchange.  The reference partitions with panzer_stk's X/Y/Z Procs (/root/reference/src/interfaces/meshInterface.cpp:140-155),
// assembles every rank's own elements into the OVERLAPPED vector / matrix and folds the rows of shared DOFs onto their owner with
// Tpetra doExport(ADD) (/root/reference/src/interfaces/linearAlgebraInterface.hpp:296-337).  Here a DOF shared by several blocks is
// owned by the lowest rank that holds it; every other holder sends its row (residual entry + CRS values) to the owner each step.
// The plan is built from per-DOF global ids and owners, so a caller with Panzer's owned / ghosted GID lists can supply them
// (mrh_set_partition) instead of using the brick partition.  No CUDA here: exchange.inl adds the device side.
#pragma once
#include <algorithm>
#include <cstdint>
#include <map>
#include <stdexcept>
#include <vector>
#include "brick.hpp"

namespace mrh {

struct Partition {
  int rank = 0, world = 1;
  std::vector<int64_t> gids;      // (nlocal) global id of every local DOF
  std::vector<int> owner;         // (nlocal) owning rank
  bool active() const { return world > 1 && !gids.empty(); }
};

// element range of block coordinate c out of P along an axis with N elements (same split as an even panzer_stk decomposition)
inline void block_range(int N, int P, int c, int& e0, int& e1) { e0 = (int)((int64_t)N * c / P); e1 = (int)((int64_t)N * (c + 1) / P); }

// Local brick of `rank` in a Px x Py x Pz decomposition (rank = px + Px (py + Py pz)) of the global brick, with
// side sets restricted to the physical boundary, plus global ids / owners of the local DOFs.
inline void build_brick_part(HostMesh& M, HostDofs& D, HostCrs& G, Partition& Pt, const LocalLayout& L, int dim, int nvar, const int* Nglob,
                             const double* lo, const double* hi, const int* procs, int rank) {
  int P[3] = {procs[0], procs[1], dim == 3 ? procs[2] : 1}, c[3], e0[3] = {0, 0, 0}, e1[3] = {1, 1, 1}, Ng[3] = {Nglob[0], Nglob[1], dim == 3 ? Nglob[2] : 1};
  const int world = P[0] * P[1] * P[2];
  if (rank < 0 || rank >= world) throw std::runtime_error("rank outside the processor grid");
  c[0] = rank % P[0]; c[1] = (rank / P[0]) % P[1]; c[2] = rank / (P[0] * P[1]);
  int nloc[3] = {1, 1, 1}; double llo[3] = {0, 0, 0}, lhi[3] = {1, 1, 1};
  for (int d = 0; d < 3; ++d) {
    if (d < dim) {
      if (P[d] < 1 || P[d] > Ng[d]) throw std::runtime_error("processor grid does not fit the mesh");
      block_range(Ng[d], P[d], c[d], e0[d], e1[d]);
      nloc[d] = e1[d] - e0[d];
      // coordinates exactly as the global builder computes them: lo + (hi - lo) * (i / N)
      llo[d] = lo[d] + (hi[d] - lo[d]) * ((double)e0[d] / (double)Ng[d]);
      lhi[d] = lo[d] + (hi[d] - lo[d]) * ((double)e1[d] / (double)Ng[d]);
    }
  }
  build_brick(M, dim, nloc, llo, lhi);
  // node coordinates from the GLOBAL formula (bitwise identical on every rank that holds the node)
  {
    const int64_t nx = nloc[0] + 1, ny = nloc[1] + 1, nz = (dim == 3) ? nloc[2] + 1 : 1;
    for (int64_t k = 0; k < nz; ++k) for (int64_t j = 0; j < ny; ++j) for (int64_t i = 0; i < nx; ++i) {
      const int64_t id = i + nx * (j + ny * k), g[3] = {i + e0[0], j + e0[1], k + e0[2]};
      for (int d = 0; d < dim; ++d) M.nodes[id * dim + d] = lo[d] + (hi[d] - lo[d]) * ((double)g[d] / (double)Ng[d]);
    }
  }
  // side sets: keep the elements only where the local face lies on the global boundary
  static const int axis2[4] = {1, 0, 1, 0}, high2[4] = {0, 1, 1, 0}, axis3[6] = {1, 0, 1, 0, 2, 2}, high3[6] = {0, 1, 1, 0, 0, 1};
  for (auto& S : M.sides) {
    const int ax = (dim == 2) ? axis2[S.local_side] : axis3[S.local_side], hi_ = (dim == 2) ? high2[S.local_side] : high3[S.local_side];
    const bool physical = hi_ ? (c[ax] == P[ax] - 1) : (c[ax] == 0);
    if (!physical) S.elems.clear();
  }
  build_brick_dofs(D, M, L, nvar);
  // DOF coordinates from the global formula as well
  const int p = L.p;
  build_crs(G, D.lids.data(), M.nelem, D.N, D.nlocal, nvar);
  // global ids and owners: refined-grid node (I, J, K) of the global brick, id = node * nvar + var; owner = lowest rank holding it
  Pt.rank = rank; Pt.world = world; Pt.gids.assign((size_t)D.nlocal, -1); Pt.owner.assign((size_t)D.nlocal, rank);
  const int64_t GX = (int64_t)p * Ng[0] + 1, GY = (int64_t)p * Ng[1] + 1;
  std::vector<int> ord_of_slot(L.nb);
  for (int o = 0; o < L.nb; ++o) ord_of_slot[L.slot[o]] = o;
  const int ez = (dim == 3) ? nloc[2] : 1;
  for (int k = 0; k < ez; ++k) for (int j = 0; j < nloc[1]; ++j) for (int i = 0; i < nloc[0]; ++i) {
    const int64_t e = i + (int64_t)nloc[0] * (j + (int64_t)nloc[1] * k);
    for (int s = 0; s < L.nb; ++s) {
      const auto& t = L.tix[ord_of_slot[s]];
      const int64_t g[3] = {(int64_t)p * (i + e0[0]) + t[0], (int64_t)p * (j + e0[1]) + t[1], (int64_t)p * (k + e0[2]) + t[2]};
      int oc[3];
      for (int d = 0; d < 3; ++d) {
        // a node on the low face of this block (and not on the global boundary) also belongs to the block below
        const bool on_low = d < dim && g[d] == (int64_t)p * e0[d] && c[d] > 0;
        oc[d] = on_low ? c[d] - 1 : c[d];
      }
      const int own = oc[0] + P[0] * (oc[1] + P[1] * oc[2]);
      const int64_t node = g[0] + GX * (g[1] + GY * g[2]);
      for (int f = 0; f < nvar; ++f) {
        const int lid = D.lids[(size_t)e * D.N + s * nvar + f];
        Pt.gids[lid] = node * nvar + f; Pt.owner[lid] = own;
      }
    }
  }
  for (int64_t l = 0; l < D.nlocal; ++l) {
    const int64_t node = Pt.gids[l] / nvar;
    const int64_t g[3] = {node % GX, (node / GX) % GY, node / (GX * GY)};
    for (int d = 0; d < dim; ++d) D.dofxyz[(size_t)l * dim + d] = lo[d] + (hi[d] - lo[d]) * ((double)g[d] / (double)((int64_t)p * Ng[d]));
  }
}

// ---- exchange plan ---------------------------------------------------------------------------------------------------
// Message to owner p, per step:  [ residual of the rows sent to p | CRS values of those rows, row after row in CRS order ].
// Meta message (once): per sent row  [ gid(row), nnz, gid(col_0) ... gid(col_nnz-1) ].
struct ExchangePlan {
  struct Peer {
    int rank = -1;
    std::vector<int> send_rows;            // local rows sent to this peer, ascending gid
    std::vector<int64_t> send_slots;       // their CRS slots, row after row
    std::vector<int> recv_rows;            // local (owned) rows receiving from this peer, in the peer's send order
    std::vector<int64_t> recv_slots;       // local CRS slot of every received value, or -(k + 1) for ghost-column entry k
  };
  std::vector<Peer> peers;                 // every rank I send to or receive from, ascending rank
  // ghost-column block: entries of owned rows whose column is not a local DOF (they arrive from the neighbour and stay separate,
  // like the off-process columns of Tpetra's non-overlapped matrix)
  std::vector<int> ghost_row; std::vector<int64_t> ghost_colgid;
  bool complete = false;

  Peer& peer(int r) {
    for (auto& p : peers) if (p.rank == r) return p;
    Peer np; np.rank = r;
    auto it = peers.begin(); while (it != peers.end() && it->rank < r) ++it;
    return *peers.insert(it, np);
  }
  const Peer* find(int r) const { for (auto& p : peers) if (p.rank == r) return &p; return nullptr; }

  // rows I hold but do not own, grouped by owner
  void build_send(const Partition& Pt, const HostCrs& G) {
    peers.clear(); ghost_row.clear(); ghost_colgid.clear(); ghost_index_.clear(); gid2lid_.clear(); complete = false;
    std::vector<std::pair<int64_t, int>> order;
    for (int64_t l = 0; l < (int64_t)Pt.gids.size(); ++l) if (Pt.owner[l] != Pt.rank) order.push_back({Pt.gids[l], (int)l});
    std::sort(order.begin(), order.end());
    for (auto& gl : order) {
      Peer& p = peer(Pt.owner[gl.second]);
      p.send_rows.push_back(gl.second);
      for (int64_t s = G.rowptr[gl.second]; s < G.rowptr[gl.second + 1]; ++s) p.send_slots.push_back(s);
    }
  }
  std::vector<int64_t> meta_for(int r, const Partition& Pt, const HostCrs& G) const {
    std::vector<int64_t> m;
    const Peer* p = find(r);
    if (!p) return m;
    for (int row : p->send_rows) {
      m.push_back(Pt.gids[row]); m.push_back(G.rowptr[row + 1] - G.rowptr[row]);
      for (int64_t s = G.rowptr[row]; s < G.rowptr[row + 1]; ++s) m.push_back(Pt.gids[G.colind[s]]);
    }
    return m;
  }
  // meta received from rank r: where its values land
  void ingest(int r, const int64_t* m, int64_t n, const Partition& Pt, const HostCrs& G) {
    if (n == 0) return;
    if (gid2lid_.empty()) { gid2lid_.reserve(Pt.gids.size()); for (int64_t l = 0; l < (int64_t)Pt.gids.size(); ++l) gid2lid_.push_back({Pt.gids[l], (int)l}); std::sort(gid2lid_.begin(), gid2lid_.end()); }
    auto lid_of = [&](int64_t g) -> int {
      auto it = std::lower_bound(gid2lid_.begin(), gid2lid_.end(), std::make_pair(g, -1));
      return (it != gid2lid_.end() && it->first == g) ? it->second : -1;
    };
    Peer& p = peer(r);
    p.recv_rows.clear(); p.recv_slots.clear();
    for (int64_t k = 0; k < n;) {
      const int row = lid_of(m[k]); const int64_t nnz = m[k + 1];
      if (row < 0 || Pt.owner[row] != Pt.rank) throw std::runtime_error("exchange plan: received a row this rank does not own");
      p.recv_rows.push_back(row);
      for (int64_t c = 0; c < nnz; ++c) {
        const int64_t cg = m[k + 2 + c]; const int col = lid_of(cg);
        int64_t slot = -1;
        if (col >= 0) {
          const int* b = &G.colind[G.rowptr[row]]; const int* e = &G.colind[0] + G.rowptr[row + 1];
          const int* it = std::lower_bound(b, e, col);
          if (it != e && *it == col) slot = it - &G.colind[0];
        }
        if (slot < 0) {                                     // the same (row, column) may arrive from several neighbours: one ghost entry
          auto key = std::make_pair(row, cg);
          auto gi = ghost_index_.find(key);
          if (gi == ghost_index_.end()) { ghost_row.push_back(row); ghost_colgid.push_back(cg); gi = ghost_index_.emplace(key, (int64_t)ghost_row.size() - 1).first; }
          slot = -(gi->second + 1);
        }
        p.recv_slots.push_back(slot);
      }
      k += 2 + nnz;
    }
  }
 private:
  std::vector<std::pair<int64_t, int>> gid2lid_;
  std::map<std::pair<int, int64_t>, int64_t> ghost_index_;
};

}  // namespace mrh
