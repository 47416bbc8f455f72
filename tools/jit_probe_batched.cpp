// compile-only check (no GPU needed): the NVRTC translation unit of the batched volume kernel builds for sm_100a
// usage: jit_probe_batched <phys 0..3> <dim> <order>
#include <nvrtc.h>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <string>
#include <vector>
#include "../mrhyde_b200/csrc/expr.hpp"
#include "../mrhyde_b200/csrc/jit.hpp"
using namespace mrh;
int main(int argc, char** argv) {
  const int phys = argc > 1 ? atoi(argv[1]) : 2, dim = argc > 2 ? atoi(argv[2]) : 3, order = argc > 3 ? atoi(argv[3]) : 1;
  ExprCompiler ec; ec.varlist = {"dx", "dy", "dz"}; ec.dim = 3;
  ec.addFunction("lambda", "1.0", "ip");
  ec.addFunction("mu", "1.0", "ip");
  ec.addFunction("A", "1.0", "ip");
  ec.addFunction("B", "2.0", "ip");
  ec.addFunction("C", "3.0", "ip");
  ec.addFunction("dxxx", "(A*pi)*(A*pi)*sin(A*pi*x)*sin(A*pi*y)*sin(A*pi*z)", "ip");
  ec.addFunction("dxyy", "(A*pi)*(A*pi)*sin(A*pi*x)*sin(A*pi*y)*sin(A*pi*z)", "ip");
  ec.addFunction("dxzz", "(A*pi)*(A*pi)*sin(A*pi*x)*sin(A*pi*y)*sin(A*pi*z)", "ip");
  ec.addFunction("dxxy", "-1.0*(A*pi)*(A*pi)*cos(A*pi*x)*cos(A*pi*y)*sin(A*pi*z)", "ip");
  ec.addFunction("dxxz", "-1.0*(A*pi)*(A*pi)*cos(A*pi*x)*sin(A*pi*y)*cos(A*pi*z)", "ip");
  ec.addFunction("dxyz", "-1.0*(A*pi)*(A*pi)*sin(A*pi*x)*cos(A*pi*y)*cos(A*pi*z)", "ip");
  ec.addFunction("dyxx", "(B*pi)*(B*pi)*sin(B*pi*x)*sin(B*pi*y)*sin(B*pi*z)", "ip");
  ec.addFunction("dyyy", "(B*pi)*(B*pi)*sin(B*pi*x)*sin(B*pi*y)*sin(B*pi*z)", "ip");
  ec.addFunction("dyzz", "(B*pi)*(B*pi)*sin(B*pi*x)*sin(B*pi*y)*sin(B*pi*z)", "ip");
  ec.addFunction("dyxy", "-1.0*(B*pi)*(B*pi)*cos(B*pi*x)*cos(B*pi*y)*sin(B*pi*z)", "ip");
  ec.addFunction("dyxz", "-1.0*(B*pi)*(B*pi)*cos(B*pi*x)*sin(B*pi*y)*cos(B*pi*z)", "ip");
  ec.addFunction("dyyz", "-1.0*(B*pi)*(B*pi)*sin(B*pi*x)*cos(B*pi*y)*cos(B*pi*z)", "ip");
  ec.addFunction("dzxx", "(C*pi)*(C*pi)*sin(C*pi*x)*sin(C*pi*y)*sin(C*pi*z)", "ip");
  ec.addFunction("dzyy", "(C*pi)*(C*pi)*sin(C*pi*x)*sin(C*pi*y)*sin(C*pi*z)", "ip");
  ec.addFunction("dzzz", "(C*pi)*(C*pi)*sin(C*pi*x)*sin(C*pi*y)*sin(C*pi*z)", "ip");
  ec.addFunction("dzxy", "-1.0*(C*pi)*(C*pi)*cos(C*pi*x)*cos(C*pi*y)*sin(C*pi*z)", "ip");
  ec.addFunction("dzxz", "-1.0*(C*pi)*(C*pi)*cos(C*pi*x)*sin(C*pi*y)*cos(C*pi*z)", "ip");
  ec.addFunction("dzyz", "-1.0*(C*pi)*(C*pi)*sin(C*pi*x)*cos(C*pi*y)*cos(C*pi*z)", "ip");
  ec.addFunction("source dx", "(lambda+2.0*mu)*dxxx + mu*(dxyy+dyxy+dxzz+dzxz) + lambda*(dyxy+dzxz)", "ip");
  ec.addFunction("source dy", "(lambda+2.0*mu)*dyyy + mu*(dyxx+dxxy+dyzz+dzyz) + lambda*(dxxy+dzyz)", "ip");
  ec.addFunction("source dz", "(lambda+2.0*mu)*dzzz + mu*(dzxx+dxxz+dzyy+dyyz) + lambda*(dxxz+dyyz)", "ip");
  std::string src, preamble;
  int natoms = 0;
  if (!getenv("MRH_PROBE_NOATOMS")) {
    Prog progs[MRH_MAXPROG] = {};
    for (int k = 2; k < 5; ++k) progs[k] = ec.compile(k == 2 ? "source dx" : k == 3 ? "source dy" : "source dz", "ip");
    src = jit_batched_atom_source(progs, 8, ec.code, ec.cpool, 32, natoms, &preamble);
    printf("atoms: %d\n", natoms);
  }
  if (src.empty())
  for (int k = 0; k < 8; ++k) {
    const std::string name = "p" + std::to_string(k);
    if (k >= 2 && k < 5) src += jit_function_source(name, ec.compile(k == 2 ? "source dx" : k == 3 ? "source dy" : "source dz", "ip"), ec.code, ec.cpool);
    else src += "  static __device__ __forceinline__ double " + name + "(double, double, double, double) { return 0.0; }\n";
  }
  const int thr = getenv("MRH_PROBE_THREADS") ? atoi(getenv("MRH_PROBE_THREADS")) : 192, mb = getenv("MRH_PROBE_MINB") ? atoi(getenv("MRH_PROBE_MINB")) : 0;
  JitCache jc; jc.setBatchedSource(src, phys, dim, order, thr, mb, natoms, natoms > 0 ? preamble : std::string());
  const std::string o1 = "-DMRH_BAT_BOUNDS=" + std::to_string(thr) + (mb > 0 ? "," + std::to_string(mb) : std::string()), o2 = "-DMRH_PROBE_X=1";
  // same steps as JitCache::batched(), with ptxas -v and without loading the module
  std::vector<const char*> hs, hn;
  for (int i = 0; i < k_num_embedded_headers; ++i) { hn.push_back(k_embedded_headers[i].name); hs.push_back(k_embedded_headers[i].text); }
  std::string full = jc.batchedSource();
  nvrtcProgram p; nvrtcCreateProgram(&p, full.c_str(), "probe.cu", (int)hs.size(), hs.data(), hn.data());
  const char* extra = getenv("MRH_PROBE_DEFINE") ? getenv("MRH_PROBE_DEFINE") : "-DMRH_PROBE_NONE=1";
  const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "-DMRH_JIT=1", "--ptxas-options=-v", extra, o1.c_str(), o2.c_str()};
  const auto t0 = std::chrono::steady_clock::now();
  nvrtcResult rc = nvrtcCompileProgram(p, 8, opts);
  const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  size_t ls; nvrtcGetProgramLogSize(p, &ls); std::string log(ls, ' '); nvrtcGetProgramLog(p, &log[0]);
  printf("rc=%d compile %.1f s\n%s\n", (int)rc, dt, log.c_str());
  if (rc == NVRTC_SUCCESS) { size_t cs; nvrtcGetCUBINSize(p, &cs); std::vector<char> c(cs); nvrtcGetCUBIN(p, c.data());
    if (FILE* f = fopen("mrhyde_b200/build/jit_probe_batched.cubin", "wb")) { fwrite(c.data(), 1, cs, f); fclose(f); } }
  return rc != NVRTC_SUCCESS;
}
