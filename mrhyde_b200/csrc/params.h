// Launch parameter block shared by every assembly kernel (plain data, passed by value).
#pragma once
#if !defined(__CUDACC_RTC__)
#include <cstdint>
#endif
#include "bytecode.h"

namespace mrh {

enum Phys { PHYS_THERMAL = 0, PHYS_CDR = 1, PHYS_LE = 2, PHYS_NS = 3, PHYS_STOKES = 4, PHYS_HELMHOLTZ = 5, PHYS_BURGERS = 6 };      // stokes, helmholtz: the next HGRAD modules of SURVEY.md 8f3
enum BC { BC_NONE = 0, BC_DIRICHLET = 1, BC_WEAK_DIRICHLET = 2, BC_NEUMANN = 3, BC_FLUX = 4 };

// program slots inside KParams::progs
enum { TH_SOURCE = 0, TH_DIFF, TH_CP, TH_RHO, TH_BX, TH_BY, TH_BZ };                      // thermal, "ip"
enum { CDR_SOURCE = 0, CDR_DIFF, CDR_CP, CDR_RHO, CDR_REACTION, CDR_XVEL, CDR_YVEL, CDR_ZVEL };
enum { LE_LAMBDA = 0, LE_MU, LE_SRC0, LE_SRC1, LE_SRC2 };
enum { NS_SRC_UX = 0, NS_SRC_PR, NS_SRC_UY, NS_SRC_UZ, NS_DENSITY, NS_VISCOSITY };
// helmholtz: c2r_x, c2i_x, c2r_y, c2i_y, c2r_z, c2i_z at 0..5 (volume and side launches alike); volume: sources in the pre-evaluated
// slots, omega2 beyond them (evaluated on the fly); side: Robin coefficients, then the side sources in the slots of the unused variables 2, 3
enum { BG_SRC = 0, BG_DIFF, BG_VEL, BG_C1 = 5, BG_C2, BG_SUPG_C, BG_SUPG_C1, BG_SUPG_C2, BG_SIDE_VEL = 0 };      // Burgers ("supg C1" / "supg C2" beyond the pre-evaluated slots)
enum { HZ_C2 = 0, HZ_SRC_R = 6, HZ_SRC_I = 7, HZ_OMEGA2R = 8, HZ_OMEGA2I = 9, HZ_SIDE_ALPHA_R = 6, HZ_SIDE_ALPHA_I = 7, HZ_SIDE_SRC_R = 10, HZ_SIDE_SRC_I = 11 };
enum { SIDE_DIFF = 0, SIDE_LAMBDA = 0, SIDE_MU = 1, SIDE_BCSRC = 8, SIDE_FLUX = 12 };     // "side ip"
enum { VOL_INIT = 8 };                                                                   // "initial <var>" at "ip" (projection launches)
enum { PROJ_NONE = 0, PROJ_DIRICHLET = 1, PROJ_INITIAL = 2 };

struct RefTab {          // reference-element tables of one point set (volume rule, or the rule of one local side)
  const double* B;       // [nq][NB]       field basis values
  const double* G;       // [nq][NB][DIM]  field basis reference gradients
  const double* VB;      // [nq][NV]       vertex (geometry) basis values, shards vertex order
  const double* VG;      // [nq][NV][DIM]  vertex basis reference gradients
  const double* W;       // [nq]           reference weights
  const double* FR;      // [nq][ceil(NB / 8)][32]  (B | G) in DMMA A-fragment order (batched kernel, bases that do not store T)
  int nq;
  int pad_;
  double tanU[3], tanV[3];   // reference side tangents (side launches only)
};

struct TimeParams {      // Workset::computeSolnTransientSeeded coefficients (workset.cpp:594-620)
  int transient, nsteps, nstages, stage;
  double alpha_u, alpha_t, timewt, time, dt;
  double a_over_b[8];    // butcher_A(stage,s)/butcher_b(s), s < stage
  double bdf[8];         // BDF weights
};

struct PhysParams {
  int have_advection;    // thermal "include advection"
  int useSUPG, usePSPG;  // navierstokes
  int bc[4];             // BC code per variable on the current side
  double formparam;      // thermal/LE "form_param"
  double le_penalty;     // LE "penalty"
  int proj;              // L2-projection launch (assemble.h projection): PROJ_DIRICHLET on a side set, PROJ_INITIAL over the volume
  int useLSIC;           // stokes "useLSIC"
  int useEvisc;          // Burgers "use entropy viscosity" (its "use SUPG" travels in useSUPG)
};

struct TilePlan {          // tile-staged scatter of the specialised kernels (built at finalize, engine_tiles.inl)
  const int* perm;           // (nelem)   element order: tile t = perm[128 t .. 128 t + 127]
  const int64_t* eptr;       // (ntiles+1) running count of distinct CRS slots: tile t has E_t = eptr[t + 1] - eptr[t] of them
  const int64_t* ebeg;       // (ntiles)  start of the tile's slot list in `off` (lists are de-duplicated: tiles of one shape share one)
  const int64_t* rptr;       // (ntiles+1) first residual row of each tile in `rows`
  const int64_t* base;       // (ntiles)  smallest CRS slot of the tile
  const uint32_t* off;       // (dictionary) distinct CRS slots of a tile relative to base, ascending; every list 16-byte aligned
  const int* rows;           // (sum R_t) distinct residual rows (LIDs) of each tile, ascending
  const uint16_t* clist;     // (nlist, 128, 72) the tile's contributions sorted by destination; thread t reads row t:
                             //   bits 0..14 = entry * 128 + thread in the shared element store, bit 15 = first of its destination
  const uint16_t* cbase;     // (nlist, 128) index in [slots | rows | masked] of the destination of thread t's first contribution
  const int* cl_id;          // (ntiles)  contribution list of each tile (lists are de-duplicated)
  int64_t ntiles;
  int64_t tile_begin, tile_end; // tiles this launch covers (the host-buffer entry point assembles in chunks to overlap the copy-out)
  int max_acc;               // largest E_t + R_t: accumulation doubles per CTA (+1 cell for masked contributions)
  int nk;                    // 36 (symmetric element matrix) or 64: entry numbering the lists were built for
  int max_e, max_r;          // largest E_t, largest R_t (staged slot / row lists)
};

struct KParams {
  const double* nodes;       // (nnode, DIM)
  const int* e2n;            // (nelem, NV)
  const int* lids;           // (nelem, N)     dof_lids
  const uint8_t* fixed;      // (nlocal)       isFixedDOF
  const int64_t* rowptr;     // (nlocal+1)
  const uint16_t* pat;       // (npattern, N*N) row-relative CRS slot of (row slot, col slot)
  const int* pat_id;         // (nelem)        pattern of each element
  const int* elem_list;      // elements of this launch (side sets); null = 0..nwork-1
  int64_t nwork;
  const double* u;           // (nlocal)
  const double* u_prev;      // (nlocal, nsteps)   transient only
  const double* u_stage;     // (nlocal, nstages)  transient only
  double* res;               // (nlocal)   += -r
  double* jac;               // (nnz)      += dr/du
  const Instr* code;
  const double* cpool;
  RefTab tab;
  TimeParams tp;
  PhysParams ph;
  Prog progs[MRH_MAXPROG];
  int offsets[4 * 27];       // offsets(var, basis dof) -> element slot
  int is_side, lump_mass, compute_jac, elems_per_block;
  TilePlan tile;
};

}  // namespace mrh
