// TEST PROGRAM (not product code): drives the engine through the C++ facade include/mrhyde_b200.hpp the way MrHyDE's
// SolverManager drives AssemblyManager (solverManager.cpp:1528-1560), compiled with plain g++ against include/ and linked
// to libmrhyde_b200.so; the result is compared with the oracle's C API (oracle/src/capi.cpp) on the same state.
//   facade_test --layout            : prints the C-ABI struct layout (compared with the ctypes mirror; needs no GPU)
//   facade_test <C2|C4|C5|C2T> <n>  : parity run on the GPU, one JSON line, exit code 0 on success
//   facade_test --yaml <deck> <pre> : MrHyDE-format input deck in, Matrix Market dumps of J / res / u out
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <cuda_runtime_api.h>
#include "mrhyde_b200.hpp"

extern "C" {      // oracle C API (checker)
void* orc_create(const char* text);
void orc_destroy(void* h);
const char* orc_last_error();
void orc_sizes(void* h, int64_t* out);
void orc_get_crs(void* h, int64_t* rowptr, int* colind);
void orc_get_dofmap(void* h, int* lids, int* offsets, uint8_t* is_fixed, double* dofxyz, int* dofvar);
void orc_set_time(void* h, int nstages, const double* A, const double* b, int nbdf, const double* bdf, double dt, double time, int stage);
int orc_assemble(void* h, const double* u, const double* u_prev, const double* u_stage, int compute_jacobian, double* res, double* jac, int nthreads);
}

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA error %s at %s\n", cudaGetErrorString(e_), #call); return 2; } } while (0)

static int layout() {
  printf("{\"sizeof_config\": %zu, \"sizeof_state\": %zu, \"config\": {", sizeof(mrh_config), sizeof(mrh_state));
#define F(S, f) printf("\"%s\": %zu, ", #f, offsetof(S, f))
  F(mrh_config, dim); F(mrh_config, physics); F(mrh_config, order); F(mrh_config, quadrature); F(mrh_config, side_quadrature);
  F(mrh_config, use_strong_dbc); F(mrh_config, fix_zero_rows); F(mrh_config, lump_mass); F(mrh_config, include_advection);
  F(mrh_config, use_supg); F(mrh_config, use_pspg); F(mrh_config, form_param); F(mrh_config, le_penalty); F(mrh_config, device);
  F(mrh_config, flags); F(mrh_config, use_lsic);
  printf("\"use_evisc\": %zu}, \"state\": {", offsetof(mrh_config, use_evisc));
  F(mrh_state, u); F(mrh_state, u_prev); F(mrh_state, u_stage); F(mrh_state, is_transient); F(mrh_state, nstages); F(mrh_state, nsteps);
  F(mrh_state, stage); F(mrh_state, butcher_A); F(mrh_state, butcher_b); F(mrh_state, bdf_wts); F(mrh_state, time); F(mrh_state, deltat);
  printf("\"stream\": %zu}}\n", offsetof(mrh_state, stream));
#undef F
  return 0;
}

int main(int argc, char** argv) {
  if (argc > 1 && !strcmp(argv[1], "--layout")) return layout();
  if (argc > 3 && !strcmp(argv[1], "--yaml")) {
    // facade_test --yaml <input.yaml> <prefix>: read a MrHyDE-format deck, assemble residual + Jacobian at a synthetic state and dump
    // J, res and the state as Matrix Market files (what linalg->writeToFile leaves behind, solverManager.cpp:1653)
    using namespace mrh;
    try {
      auto settings = readInput(std::string(argv[2]));
      std::vector<BlockData> blocks(1);
      AssemblyManager<> assembler(settings, blocks);
      assembler.createGroups(); assembler.allocateGroupStorage(); assembler.createWorkset(); assembler.finalizeFunctions();
      const auto S = assembler.sizes(0);
      std::vector<int64_t> rp; std::vector<int32_t> ci, lids, offs, dofvar; std::vector<uint8_t> fixed; std::vector<double> xyz;
      assembler.getGraph(0, rp, ci); assembler.getDOFs(0, lids, offs, fixed, xyz, dofvar);
      std::vector<double> u(S.nlocal);
      for (int64_t i = 0; i < S.nlocal; ++i) { double v = 1.0; for (int d = 0; d < S.dim; ++d) v *= sin(2 * M_PI * xyz[i * S.dim + d]); u[i] = v + 1e-3 * sin(12.9898 * i + 0.1 * dofvar[i]); }
      auto sol0 = std::make_shared<Vector>(); sol0->nlocal = S.nlocal;
      auto res = std::make_shared<Vector>(); res->nlocal = S.nlocal;
      auto J = std::make_shared<CrsMatrix>(); J->rowptr = rp.data(); J->colind = ci.data(); J->nrows = S.nlocal;
      CK(cudaMalloc((void**)&sol0->data, S.nlocal * sizeof(double))); CK(cudaMemcpy(sol0->data, u.data(), S.nlocal * sizeof(double), cudaMemcpyHostToDevice));
      CK(cudaMalloc((void**)&res->data, S.nlocal * sizeof(double))); CK(cudaMemset(res->data, 0, S.nlocal * sizeof(double)));
      CK(cudaMalloc((void**)&J->values, S.nnz * sizeof(double))); CK(cudaMemset(J->values, 0, S.nnz * sizeof(double)));
      std::vector<vector_RCP> sol{sol0}, none; vector_RCP Psol;
      assembler.assembleJacRes(0, 0, sol, none, none, none, none, none, true, false, false, res, J, false, 0.0, false, false, 0, Psol, false, 1.0);
      CK(cudaDeviceSynchronize());
      std::vector<double> r(S.nlocal), jv(S.nnz);
      CK(cudaMemcpy(r.data(), res->data, r.size() * sizeof(double), cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(jv.data(), J->values, jv.size() * sizeof(double), cudaMemcpyDeviceToHost));
      const std::string pre(argv[3]);
      writeMatrixMarket(pre + "_J.mtx", rp.data(), ci.data(), jv.data(), S.nlocal, S.nlocal);
      writeMatrixMarket(pre + "_res.mtx", r.data(), S.nlocal);
      writeMatrixMarket(pre + "_u.mtx", u.data(), S.nlocal);
      printf("{\"nlocal\": %lld, \"nnz\": %lld, \"path\": \"%s\"}\n", (long long)S.nlocal, (long long)S.nnz, assembler.kernelPath(0));
      return 0;
    } catch (const std::exception& e) { fprintf(stderr, "exception: %s\n", e.what()); return 2; }
  }
  const std::string which = argc > 1 ? argv[1] : "C2";
  const int n = argc > 2 ? atoi(argv[2]) : 12;
  using namespace mrh;
  try {
    // ---- the input deck, laid out like regression/<case>/input.yaml -----------------------------------------------
    auto settings = std::make_shared<ParameterList>();
    ParameterList& mesh = settings->sublist("Mesh");
    mesh.set("dimension", 3); mesh.set("NX", n); mesh.set("NY", n + 1); mesh.set("NZ", n - 1);
    ParameterList& phys = settings->sublist("Physics").sublist("eblock-0_0_0");
    ParameterList& disc = settings->sublist("Discretization").sublist("eblock-0_0_0");
    ParameterList& fns = settings->sublist("Functions");
    std::string otext = "dim: 3\nNX: " + std::to_string(n) + "\nNY: " + std::to_string(n + 1) + "\nNZ: " + std::to_string(n - 1) + "\n";
    bool transient = false;
    if (which == "C2" || which == "C2T") {
      transient = which == "C2T";
      phys.set("modules", "thermal");
      phys.sublist("Dirichlet conditions").sublist("e").set("all boundaries", "0.0");
      phys.sublist("Neumann conditions").sublist("e").set("top", "1.0+x");           // overrides the Dirichlet entry on that side
      disc.sublist("order").set("e", 1);
      fns.set("thermal source", "12*(pi*pi)*sin(2*pi*x)*sin(2*pi*y)*sin(2*pi*z)");
      fns.set("thermal diffusion", "1.0+0.5*x*y");
      otext += "physics: thermal\norder: 1\nfunction.thermal source: 12*(pi*pi)*sin(2*pi*x)*sin(2*pi*y)*sin(2*pi*z)\n"
               "function.thermal diffusion: 1.0+0.5*x*y\ndirichlet.e.all boundaries: 0.0\nneumann.e.top: 1.0+x\n";
      if (transient) { fns.set("density", "2.0"); otext += "function.density: 2.0\ntransient: true\n"; }
    } else if (which == "C4") {
      phys.set("modules", "linearelasticity");
      for (const char* v : {"dx", "dy", "dz"}) {
        phys.sublist("Dirichlet conditions").sublist(v).set("all boundaries", "0.0");
        disc.sublist("order").set(v, 1);
        otext += std::string("dirichlet.") + v + ".all boundaries: 0.0\n";
      }
      fns.set("lambda", "1.0"); fns.set("mu", "1.0+x"); fns.set("source dx", "sin(pi*x)*y"); fns.set("source dz", "z");
      otext += "physics: linearelasticity\norder: 1\nfunction.lambda: 1.0\nfunction.mu: 1.0+x\nfunction.source dx: sin(pi*x)*y\nfunction.source dz: z\n";
    } else if (which == "C5") {
      phys.set("modules", "navierstokes"); phys.set("useSUPG", true); phys.set("usePSPG", true);
      for (const char* v : {"ux", "uy", "uz"}) {
        for (const char* s : {"bottom", "top", "front", "back"}) { phys.sublist("Dirichlet conditions").sublist(v).set(s, "0.0"); otext += std::string("dirichlet.") + v + "." + s + ": 0.0\n"; }
      }
      for (const char* v : {"ux", "pr", "uy", "uz"}) disc.sublist("order").set(v, 1);
      fns.set("source ux", "1.0");
      otext += "physics: navierstokes\norder: 1\nfunction.source ux: 1.0\nsetting.useSUPG: true\nsetting.usePSPG: true\n";
    } else { fprintf(stderr, "unknown case %s\n", which.c_str()); return 2; }

    // ---- MrHyDE's driver sequence (main.cpp: assembler->createGroups(), allocateGroupStorage(), createWorkset(), finalizeFunctions())
    std::vector<BlockData> blocks(1);
    AssemblyManager<> assembler(settings, blocks);
    assembler.createGroups();
    assembler.allocateGroupStorage();
    assembler.createWorkset();
    assembler.finalizeFunctions();
    assembler.createFixedDOFs();
    const auto S = assembler.sizes(0);

    // ---- oracle on the same problem; indexing must be identical ----------------------------------------------------
    void* orc = orc_create(otext.c_str());
    if (!orc) { fprintf(stderr, "oracle: %s\n", orc_last_error()); return 2; }
    int64_t os[15]; orc_sizes(orc, os);
    if (os[7] != S.nlocal || os[8] != S.nnz || os[1] != S.nelem) { fprintf(stderr, "size mismatch\n"); return 1; }
    std::vector<int64_t> rp, orp(S.nlocal + 1); std::vector<int32_t> ci, oci(S.nnz);
    assembler.getGraph(0, rp, ci); orc_get_crs(orc, orp.data(), oci.data());
    std::vector<int32_t> lids, offs, dofvar, olids(S.nelem * S.ndof_elem); std::vector<uint8_t> fixed, ofixed(S.nlocal); std::vector<double> xyz;
    assembler.getDOFs(0, lids, offs, fixed, xyz, dofvar);
    orc_get_dofmap(orc, olids.data(), nullptr, ofixed.data(), nullptr, nullptr);
    const bool index_ok = rp == orp && ci == oci && lids == olids && fixed == ofixed;

    // ---- state, device vectors (the stand-ins of the Tpetra overlapped vector / matrix) ----------------------------
    const int nsteps = 2, nstages = 2, stage = 1;
    std::vector<double> u(S.nlocal), uprev((size_t)S.nlocal * nsteps), ustage((size_t)S.nlocal * nstages);
    for (int64_t i = 0; i < S.nlocal; ++i) {
      const double* x = &xyz[i * 3];
      u[i] = sin(2 * M_PI * x[0]) * sin(2 * M_PI * x[1]) * sin(2 * M_PI * x[2]) + 1e-3 * sin(12.9898 * i + 0.1 * dofvar[i]);
      for (int k = 0; k < nsteps; ++k) uprev[i * nsteps + k] = 0.9 * u[i] + 0.01 * cos(i + k);
      for (int k = 0; k < nstages; ++k) ustage[i * nstages + k] = 1.1 * u[i] - 0.02 * sin(i + 2 * k);
    }
    auto dvec = [&](const std::vector<double>& h, int ncols) { auto v = std::make_shared<Vector>(); v->nlocal = S.nlocal; v->ncols = ncols;
      cudaMalloc((void**)&v->data, h.size() * sizeof(double)); cudaMemcpy(v->data, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice); return v; };
    std::vector<vector_RCP> sol{dvec(u, 1)}, sol_prev{dvec(uprev, nsteps)}, sol_stage{dvec(ustage, nstages)}, phi, phi_stage, phi_prev;
    vector_RCP res = dvec(std::vector<double>(S.nlocal, 0.0), 1), Psol;
    auto J = std::make_shared<CrsMatrix>(); J->rowptr = rp.data(); J->colind = ci.data(); J->nrows = S.nlocal;
    CK(cudaMalloc((void**)&J->values, S.nnz * sizeof(double))); CK(cudaMemset(J->values, 0, S.nnz * sizeof(double)));
    const std::vector<double> A{0.25, 0.0, 0.5, 0.25}, b{0.5, 0.5}, c{0.25, 0.75}, bdf{1.5, -2.0, 0.5};
    const double dt = 0.02, time = 0.1;
    if (transient) {
      assembler.setWorksetButcher(0, 0, A, b, c); assembler.setWorksetBDF(0, 0, bdf); assembler.updateStage(stage, time, dt);
      orc_set_time(orc, nstages, A.data(), b.data(), (int)bdf.size(), bdf.data(), dt, time + c[stage] * dt, stage);   // the oracle takes the workset (stage) time
    }
    // ---- the hot call, with the reference's argument list (solverManager.cpp:1536-1547) -----------------------------
    assembler.assembleJacRes(0, transient ? stage : 0, sol, sol_stage, sol_prev, phi, phi_stage, phi_prev, true, false, false, res, J,
                             transient, transient ? time : 0.0, false, false, 0, Psol, false, transient ? dt : 1.0);
    assembler.dofConstraints(0, J, res, 0.0, true, false);
    CK(cudaDeviceSynchronize());
    std::vector<double> r(S.nlocal), jv(S.nnz), orr(S.nlocal, 0.0), ojv(S.nnz, 0.0);
    CK(cudaMemcpy(r.data(), res->data, r.size() * sizeof(double), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(jv.data(), J->values, jv.size() * sizeof(double), cudaMemcpyDeviceToHost));
    if (orc_assemble(orc, u.data(), transient ? uprev.data() : nullptr, transient ? ustage.data() : nullptr, 1, orr.data(), ojv.data(), 4) != 0) { fprintf(stderr, "oracle: %s\n", orc_last_error()); return 2; }
    auto relerr = [](const std::vector<double>& a, const std::vector<double>& b) { double d = 0, s = 0; for (size_t i = 0; i < a.size(); ++i) { d = std::max(d, std::fabs(a[i] - b[i])); s = std::max(s, std::fabs(b[i])); } return d / (s > 0 ? s : 1.0); };
    const double er = relerr(r, orr), ej = relerr(jv, ojv);
    // residual-only pass through assembleRes must reproduce the residual
    CK(cudaMemset(res->data, 0, S.nlocal * sizeof(double)));
    assembler.assembleRes(0, transient ? stage : 0, sol, sol_stage, sol_prev, phi, phi_stage, phi_prev, Psol, res, J, transient, transient ? time : 0.0, transient ? dt : 1.0);
    CK(cudaDeviceSynchronize());
    std::vector<double> r2(S.nlocal);
    CK(cudaMemcpy(r2.data(), res->data, r2.size() * sizeof(double), cudaMemcpyDeviceToHost));
    const double er2 = relerr(r2, orr);
    const bool ok = index_ok && er <= 1e-12 && ej <= 1e-12 && er2 <= 1e-12;
    printf("{\"case\": \"%s\", \"n\": %d, \"nlocal\": %lld, \"nnz\": %lld, \"indexing_bit_exact\": %s, \"res_rel\": %.3e, \"jac_rel\": %.3e, \"res_only_rel\": %.3e, "
           "\"path\": \"%s\", \"bc_top\": \"%s\", \"ok\": %s}\n", which.c_str(), n, (long long)S.nlocal, (long long)S.nnz, index_ok ? "true" : "false", er, ej, er2,
           assembler.kernelPath(0), assembler.wkset[0]->var_bcs[0].back().c_str(), ok ? "true" : "false");
    orc_destroy(orc);
    return ok ? 0 : 1;
  } catch (const std::exception& e) { fprintf(stderr, "exception: %s\n", e.what()); return 2; }
}
