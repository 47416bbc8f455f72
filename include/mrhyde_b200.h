/* mrhyde_b200 — C ABI of the B200-native finite-element assembly engine.
 *
 * The reference (dannys4/MrHyDE) has no FFI: its assembly path is C++ classes compiled into one
 * executable.  This header is the boundary a MrHyDE maintainer would bind instead of the Kokkos loops;
 * every entry point names the reference interface it replaces (paths relative to /root/reference).
 * INTEGRATION.md shows the reference-side glue (AssemblyManager / PhysicsBase / FunctionManager shims).
 *
 * Conventions
 *   - every function returns 0 on success or a negative MRH_E* code; mrh_last_error() gives the text;
 *     nothing throws across the ABI.
 *   - set-up arrays (mesh, LIDs, CRS graph, side sets) are HOST pointers, copied by the engine.
 *   - hot-path arrays (solution, residual, Jacobian values) are DEVICE pointers owned by the caller —
 *     the Tpetra local views the reference extracts at src/managers/assemblyManager.cpp:2172-2178 —
 *     except in the *_host convenience calls.
 *   - residual is SUBTRACTED (res += -r), Jacobian is ADDED (assemblyManager.cpp:4094-4141); the caller
 *     zeroes both beforehand (src/managers/solverManager.cpp:1528-1533).
 *   - one handle per GPU; a handle is not re-entrant (same contract as one AssemblyManager per rank).
 */
#ifndef MRHYDE_B200_H
#define MRHYDE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mrh_engine* mrh_handle;

enum {
  MRH_OK = 0,
  MRH_EINVAL = -1,      /* bad argument / call order */
  MRH_EPARSE = -2,      /* expression could not be decomposed (functionManager.cpp:431) */
  MRH_EUNSUPPORTED = -3,/* feature outside the hot-path scope */
  MRH_ECUDA = -4,       /* CUDA runtime error (no CPU fallback exists) */
  MRH_ENOMEM = -5
};

/* BC codes of var_bcs(var, side): src/interfaces/discretizationInterface.cpp:2556-2593 */
enum { MRH_BC_NONE = 0, MRH_BC_DIRICHLET = 1, MRH_BC_WEAK_DIRICHLET = 2, MRH_BC_NEUMANN = 3, MRH_BC_FLUX = 4 };

/* Replaces the slice of the YAML ParameterList the path reads (src/interfaces/userInterface.hpp):
 * Physics/modules, Discretization/order + quadrature, Solver/{use strong DBCs, fix zero rows, lump mass},
 * module settings (thermal.cpp:24-41, linearelasticity.cpp:21-61, navierstokes.cpp:20-62). */
typedef struct mrh_config {
  int32_t dim;                /* 2 or 3 */
  const char* physics;        /* "thermal" | "cdr" | "linearelasticity" | "navierstokes" | "stokes" | "helmholtz" | "burgers" */
  int32_t order;              /* HGRAD order, 1 or 2 (order 2 in 3-D: thermal and cdr only - the vector-valued modules return
                               * MRH_EUNSUPPORTED from mrh_assemble_*) */
  int32_t quadrature;         /* volume cubature degree; <0 = 2*order (discretizationInterface.cpp:166) */
  int32_t side_quadrature;    /* side cubature degree;   <0 = 2*order (:203) */
  int32_t use_strong_dbc;     /* Solver/use strong DBCs   (assemblyManager.cpp:34) */
  int32_t fix_zero_rows;      /* Solver/fix zero rows     (:37) */
  int32_t lump_mass;          /* Solver/lump mass         (:40) */
  int32_t include_advection;  /* thermal "include advection" */
  int32_t use_supg;           /* navierstokes "useSUPG" */
  int32_t use_pspg;           /* navierstokes "usePSPG" */
  double form_param;          /* thermal / LE "form_param" (default 1.0) */
  double le_penalty;          /* LE "penalty" (default 10.0) */
  int32_t device;             /* CUDA device ordinal */
  int32_t flags;              /* bit 0: force the generic kernel; bit 1: no NVRTC specialisation;
                               * bit 2: direct instead of tile-staged scatter; bit 3: no warp-specialised kernel; bit 4: no batched kernel;
                               * bit 5: NVRTC-specialise the batched kernel on small meshes too (tests); bit 6: no lane-per-point thermal kernel;
                               * bit 7: no exchange inside mrh_assemble_* when a communicator is attached;
                               * bit 8: tile-staged scatter for the Q1 quadrilateral kernel (off by default: measured slower than its direct scatter) */
  int32_t use_lsic;           /* stokes "useLSIC" (stokes.cpp:44) */
  int32_t use_evisc;          /* Burgers "use entropy viscosity" (burgers.cpp:24); its "use SUPG" travels in use_supg */
} mrh_config;

/* Workset time data: Workset::{time,deltat,isTransient,current_stage,butcher_A,butcher_b,BDF_wts}
 * (src/tools/workset.hpp; set by AssemblyManager::setWorksetButcher/BDF, updateStage, updateTimeStep). */
typedef struct mrh_state {
  const double* u;            /* device (nlocal)            current / stage solution   sol[set]       */
  const double* u_prev;       /* device (nlocal, nsteps)    previous steps, or NULL    sol_prev[set]  */
  const double* u_stage;      /* device (nlocal, nstages)   previous stages, or NULL   sol_stage[set] */
  int32_t is_transient;
  int32_t nstages, nsteps, stage;
  const double* butcher_A;    /* host (nstages,nstages) */
  const double* butcher_b;    /* host (nstages) */
  const double* bdf_wts;      /* host (nsteps+1) */
  double time, deltat;
  void* stream;               /* cudaStream_t, NULL = default stream */
} mrh_state;

void mrh_default_config(mrh_config* cfg);
int mrh_create(const mrh_config* cfg, mrh_handle* out);
int mrh_destroy(mrh_handle h);
const char* mrh_last_error(mrh_handle h);            /* h may be NULL: error of the last failed mrh_create */

/* ---- set-up: either build the synthetic brick ... ------------------------------------------------ */
/* Stands in for panzer_stk::{SquareQuad,CubeHex}MeshFactory (src/interfaces/meshInterface.cpp:128-171),
 * Panzer DOFManager (discretizationInterface.cpp:2324-2418) and LinearAlgebraInterface::setupLinearAlgebra
 * (src/interfaces/linearAlgebraInterface.cpp:145-239) on a uniform brick. */
int mrh_build_brick(mrh_handle h, const int32_t* nxyz, const double* lo, const double* hi);

/* ---- ... or hand over what Panzer/Tpetra already built ------------------------------------------- */
/* nodes (nnode,dim), elem2node (nelem, 4|8) shards vertex order: Group::nodes, assemblyManager.cpp:684-688 */
int mrh_set_mesh(mrh_handle h, const double* nodes, int64_t nnode, const int32_t* elem2node, int64_t nelem);
/* replace the vertex coordinates of the current mesh (same topology), before or after mrh_finalize: moving / distorted
 * meshes.  The reference recomputes its stored bases from Group::nodes (src/tools/group.cpp:137-245 computeBasis);
 * here nothing geometric is stored, so only the coordinate array is refreshed. */
int mrh_set_nodes(mrh_handle h, const double* nodes, int64_t nnode);
/* lids (nelem, nvar*nb) = disc->dof_lids, offsets (nvar, nb) = wkset->offsets (assemblyManager.cpp:656-669) */
int mrh_set_dofmap(mrh_handle h, const int32_t* lids, const int32_t* offsets, int64_t nlocal);
/* local graph of the overlapped matrix (linearAlgebraInterface.cpp:219-234); columns ascending per row */
int mrh_set_crs(mrh_handle h, const int64_t* rowptr, const int32_t* colind, int64_t nrows);
/* one boundary side set: BoundaryGroup key (sideset, local side id), assemblyManager.cpp:397-503 */
int mrh_add_side(mrh_handle h, const char* name, int32_t local_side, const int32_t* elems, int64_t n);

/* ---- boundary conditions and coefficient functions ----------------------------------------------- */
/* Dirichlet/Neumann/Flux condition of `var` on `side` ("all boundaries" allowed): registers the function
 * "<Dirichlet|Neumann|Flux> <var> <side>" at "side ip" (src/interfaces/physicsInterface.cpp:156-478) and
 * sets var_bcs; a Dirichlet entry with weak=0 marks the side's closure DOFs fixed (createFixedDOFs,
 * assemblyManager.cpp:185-265). */
int mrh_set_bc(mrh_handle h, const char* var, const char* side, int32_t bc_code, const char* expr);
/* FunctionManager::addFunction(name, expr, location) — src/managers/functionManager.cpp:49-69.
 * Entries of the user's Functions list are added at both "ip" and "side ip" (physicsInterface.cpp:466-483). */
int mrh_add_function(mrh_handle h, const char* name, const char* expr, const char* location);
/* Scalar / vector parameter of the input's Parameters block, usable as a leaf `name` (component 0) or `name(k)` / `name(kk)` of the
 * coefficient expressions (src/managers/functionManager.cpp:270-355).  During assembleJacRes the parameters carry no derivative
 * with respect to the solution, so they are compiled in as constants of the pass: call again, then mrh_finalize, to change them
 * (parameter sensitivities - computeParamSteadySeeded - are out of scope).  1 <= n <= 100. */
int mrh_set_parameter(mrh_handle h, const char* name, const double* values, int32_t n);

/* AssemblyManager::{createGroups, allocateGroupStorage, createWorkset, finalizeFunctions, createFixedDOFs}:
 * compiles the expressions, builds slot-pattern dictionary and fixed-DOF tables, uploads everything. */
int mrh_finalize(mrh_handle h);
/* Host-only part of mrh_finalize (expression compilation, fixed DOFs); needs no GPU.  mrh_finalize calls it. */
int mrh_finalize_host(mrh_handle h);

/* ---- the hot path --------------------------------------------------------------------------------- */
/* AssemblyManager::assembleRes  (assemblyManager.cpp:2916-3150): res += -r(u) */
int mrh_assemble_res(mrh_handle h, const mrh_state* st, double* res);
/* AssemblyManager::assembleJacRes (assemblyManager.cpp:2045-2665) incl. dofConstraints/updateJacDBC (:3158, :1167)
 * and fix-zero-rows (:2644-2663): res += -r(u), jac += dr/du on the CRS values aligned with colind. */
int mrh_assemble_jacres(mrh_handle h, const mrh_state* st, double* res, double* jac);
/* Same call with HOST buffers (pinned or pageable): copies u in, zeroes, assembles, copies res and jac out. */
int mrh_assemble_jacres_host(mrh_handle h, const mrh_state* st_host, double* res_host, double* jac_host);

/* ---- multi-GPU: element-block partition + additive interface exchange ---------------------------------- */
/* One process per GPU.  Every rank assembles its own elements into its OVERLAPPED vector / matrix; rows of DOFs it holds but does
 * not own are then added onto the owner: LinearAlgebraInterface::exportVectorFromOverlapped / exportMatrixFromOverlapped
 * (src/interfaces/linearAlgebraInterface.hpp:296-337, Tpetra doExport ADD).  With a communicator attached mrh_assemble_res /
 * mrh_assemble_jacres include that exchange (flag bit 7 of mrh_config.flags: leave it to an explicit mrh_exchange); the
 * tile-staged thermal kernels assemble the interface tiles first and overlap the exchange with the interior.  A DOF shared by
 * several ranks is owned by the lowest one; fixed rows shared by k ranks end with diagonal k (assemblyManager.cpp:1167-1180 + ADD). */
/* block `rank` of a procs[0] x procs[1] x procs[2] decomposition of the global brick: panzer_stk X/Y/Z Procs
 * (src/interfaces/meshInterface.cpp:140-155); replaces mrh_build_brick; side sets keep only faces on the physical boundary */
int mrh_build_brick_part(mrh_handle h, const int32_t* nxyz_global, const double* lo, const double* hi, const int32_t* procs, int32_t rank);
/* caller-supplied ownership (Panzer getOwnedAndGhostedIndices, discretizationInterface.cpp:2382-2397): global id and owner rank per local DOF */
int mrh_set_partition(mrh_handle h, const int64_t* gids, const int32_t* owner, int32_t rank, int32_t world);
int mrh_get_partition(mrh_handle h, int64_t* gids, int32_t* owner);
/* NCCL: rank 0 creates the 128-byte id and distributes it by any means; every rank then calls mrh_comm_init (collective) */
int mrh_comm_unique_id(void* id128);
int mrh_comm_init(mrh_handle h, const void* id128, int32_t rank, int32_t world);
/* the exchange alone, on `stream` (jac may be NULL: residual only); device pointers */
int mrh_exchange(mrh_handle h, double* res, double* jac, void* stream);
/* entries of owned rows whose column is not a local DOF (the off-process columns of Tpetra's non-overlapped matrix): local row, global
 * column id and - after an exchange - value of each; returns their number (arguments may be NULL) */
int64_t mrh_get_ghost_block(mrh_handle h, int32_t* rows, int64_t* col_gids, double* values_host);
/* exchange-plan metadata for a transport of the caller's choice (mrh_comm_init moves it over NCCL itself; the CPU tests use gloo):
 * words to send to `peer`, the words, and the words received from `peer` */
int64_t mrh_exchange_meta_size(mrh_handle h, int32_t peer);
int mrh_exchange_meta_get(mrh_handle h, int32_t peer, int64_t* out);
int mrh_exchange_meta_put(mrh_handle h, int32_t peer, const int64_t* in, int64_t n);
/* plan tables of `peer` - which: 0 send rows, 1 send CRS slots, 2 receive rows, 3 receive slots (-(k+1): ghost entry k); 4 / 5: ghost
 * rows / ghost column ids.  Returns the entry count (out may be NULL). */
int64_t mrh_get_exchange_table(mrh_handle h, int32_t peer, int32_t which, int64_t* out);

/* ---- the direct caller of the hot path, on the device ----------------------------------------------- */
/* Solver sublist of the input (src/managers/solverManager.cpp:55-70, src/interfaces/linearAlgebraInterface.cpp:47-50) */
typedef struct mrh_solver_opts {
  int32_t max_nonlinear_iters;   /* "max nonlinear iters" (10) */
  int32_t use_relative_tol;      /* "use relative TOL" (true): stop when |res|_inf / |res_0|_inf < nonlinear_tol */
  int32_t use_absolute_tol;      /* "use absolute TOL" (false) */
  int32_t allow_backtracking;    /* "allow backtracking" (true): halve the step while the scaled residual exceeds 1.1 */
  double nonlinear_tol;          /* "nonlinear TOL" (1e-6) */
  double nonlinear_abs_tol;      /* "absolute nonlinear TOL" */
  int32_t linear_method;         /* 0: restarted GMRES(krylov_vectors), 1: conjugate gradients; both Jacobi-preconditioned */
  int32_t max_linear_iters;      /* "max linear iters" (100) */
  int32_t krylov_vectors;        /* "krylov vectors" (100) */
  int32_t verbosity;             /* > 1 prints the reference's per-iteration residual lines */
  double linear_tol;             /* "linear TOL" (1e-7), relative to |rhs|_2 */
} mrh_solver_opts;
typedef struct mrh_solver_info {
  int32_t nonlinear_iters, linear_iters, converged, time_steps;
  double resnorm_first, resnorm_last, last_linear_relres, final_time;
} mrh_solver_info;
void mrh_default_solver_opts(mrh_solver_opts* opts);
/* LinearAlgebraInterface::linearSolver(set, J, res, du): jac x = rhs on the engine's CRS graph, x zeroed first; device pointers.
 * (The reference uses Belos GMRES + MueLu/Amesos2; here GMRES(m) / CG with a Jacobi preconditioner, written as CUDA kernels.) */
int mrh_linear_solve(mrh_handle h, const double* jac, const double* rhs, double* x, const mrh_solver_opts* opts, mrh_solver_info* info, void* stream);
/* SolverManager::nonlinearSolver (solverManager.cpp:1465-1709): Newton on `u` (device, nlocal, in/out; Dirichlet values of the
 * fixed DOFs must already be in u), time data from st (st->u is ignored); res (nlocal) and jac (nnz) are the caller's work arrays.
 * Returns 0 when converged, 1 when max_nonlinear_iters was reached (the reference's status), < 0 on errors. */
int mrh_nonlinear_solve(mrh_handle h, const mrh_state* st, double* u, double* res, double* jac, const mrh_solver_opts* opts, mrh_solver_info* info);
/* SolverManager::transientSolver, forward branch (solverManager.cpp:1229-1340) with setButcherTableau / setBackwardDifference
 * (:490-830): steps u from start_time to end_time with fixed deltat; tableau "BWE" | "DIRK-1,1" | "CN" | "DIRK-1,2" | "DIRK-2,2" |
 * "DIRK-2,3" | "DIRK-3,3", BDF order 1..4 (orders > 1 start up with lower orders). */
int mrh_transient_solve(mrh_handle h, double* u, double start_time, double end_time, double deltat, const char* butcher_tableau,
                        int32_t bdf_order, double* res, double* jac, const mrh_solver_opts* opts, mrh_solver_info* info, void* stream);

/* ---- mass-matrix / L2-projection assembly for Dirichlet and initial data ------------------------------ */
/* AssemblyManager::setDirichlet (src/managers/assemblyManager.cpp:1855-1948 with getDirichletBoundary / getMassBoundary :6288-6420;
 * kind 0) and AssemblyManager::setInitial (:1185-1310; kind 1): the L2-projection system of the strong Dirichlet data onto the
 * fixed DOFs (side mass matrix + side moments of "Dirichlet <var> <side>", identity on every other row) or of "initial <var>"
 * (registered with mrh_add_function at "ip"; absent = 0) over the volume.  rhs (nlocal) and mass (nnz, on the CRS graph) are
 * device arrays the caller zeroes; both are accumulated into. */
int mrh_assemble_projection(mrh_handle h, int32_t kind, double time, double* rhs, double* mass, void* stream);
/* SolverManager::projectDirichlet + setDirichlet (src/managers/solverManager.cpp:1060-1098, 1876-1935): assembles the system into the
 * caller's work arrays, solves it (conjugate gradients) and writes the values of the fixed DOFs into u (device, in/out). */
int mrh_set_dirichlet(mrh_handle h, double* u, double time, double* rhs_work, double* mass_work, const mrh_solver_opts* opts, mrh_solver_info* info, void* stream);
/* SolverManager::setInitial (L2 projection of the initial data): u = projection of "initial <var>" */
int mrh_set_initial(mrh_handle h, double* u, double* rhs_work, double* mass_work, const mrh_solver_opts* opts, mrh_solver_info* info, void* stream);

/* Allocation-free variant of the host-buffer residual assembly.  It has NO counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it): copies u in, zeroes, assembles res += -r(u), copies res out, with the device copies of u
 * and res held in a file-scope `static __device__` array of 2 x MRH_STATIC_SCRATCH_MAX_DOFS doubles (16 MiB, resident on every
 * GPU that loads the library until exit) - no cudaMalloc / cudaFree / allocating library call is made.  Steady state only.
 * Meshes with more than MRH_STATIC_SCRATCH_MAX_DOFS local DOFs are rejected with MRH_EINVAL and a message.  Every call (and
 * every handle on the same GPU) shares the one array, which keeps the previous call's contents (the residual part is cleared
 * at the start of each call): calls that overlap on the GPU - different streams or host threads - race; serialise them. */
#define MRH_STATIC_SCRATCH_MAX_DOFS 1048576
int mrh_assemble_res_host_static(mrh_handle h, const mrh_state* st_host, double* res_host);

/* ---- introspection (parity tests, bindings) ------------------------------------------------------ */
/* out[16]: dim, nelem, nnode, nvert, nvar, nb, ndof_elem, nlocal, nnz, nq, nqs, nsides, npatterns, nfixed, 0, 0 */
int mrh_get_sizes(mrh_handle h, int64_t* out);
int mrh_get_mesh(mrh_handle h, double* nodes, int32_t* elem2node);
int mrh_get_dofmap(mrh_handle h, int32_t* lids, int32_t* offsets, uint8_t* is_fixed, double* dofxyz, int32_t* dofvar);
int mrh_get_crs(mrh_handle h, int64_t* rowptr, int32_t* colind);
int mrh_get_side(mrh_handle h, int32_t s, char* name, int32_t namelen, int32_t* local_side, int32_t* elems);
int mrh_get_bc(mrh_handle h, int32_t var, int32_t side);
const char* mrh_get_varname(mrh_handle h, int32_t var);
/* evaluate a registered "ip" function on host doubles at given points (parser known-answer tests) */
int mrh_eval_function_host(mrh_handle h, const char* name, const char* location, int64_t npts, const double* xyz,
                           double time, double* out);
/* which volume kernel the last assemble call used:
 *   thermal Q1 hexahedra   "nvrtc-specialised+tiled-ws" (default with a Jacobian; "-ws8" / "-wsk": experimental variants behind
 *                          MRH_WS8=1 / MRH_WSK=1), "nvrtc-specialised+tiled", "nvrtc-specialised" (residual-only passes, or no tile
 *                          plan), "specialised+tiled", "specialised" (ahead-of-time build with the in-kernel bytecode interpreter)
 *   thermal Q1 quadrilaterals   "nvrtc-specialised-2d+tiled", "nvrtc-specialised-2d", "specialised-2d+tiled", "specialised-2d"
 *   every other module / order  "nvrtc-batched", "batched", and "generic" as the last resort */
const char* mrh_last_kernel_path(mrh_handle h);
/* number of kernels launched by the last assemble call (bench.py "gpu_launches") */
int64_t mrh_last_launch_count(mrh_handle h);
/* device time of the dominant (volume) kernel of the last assemble call, ms (CUDA events on the caller's stream) */
double mrh_last_volume_kernel_ms(mrh_handle h);
/* micro-benchmarks used for the roofline denominators: FP64 FMA peak (TFLOP/s) and RED.F64 throughput (G/s) */
int mrh_microbench(int32_t device, double* dfma_tflops, double* red_gops, double* copy_gbs);

#ifdef __cplusplus
}
#endif
#endif /* MRHYDE_B200_H */
