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
  ec.addFunction("s0", "3*(pi*pi)*sin(pi*x)*sin(pi*y)*sin(pi*z)-2*pi*pi*cos(pi*x)*cos(pi*y)*sin(pi*z)", "ip");
  ec.addFunction("s1", "exp(x)*cos(2*pi*y)+z^2", "ip");
  std::string src;
  for (int k = 0; k < 8; ++k) {
    const std::string name = "p" + std::to_string(k);
    if (k < 2) src += jit_function_source(name, ec.compile(k == 0 ? "s0" : "s1", "ip"), ec.code, ec.cpool);
    else src += "  static __device__ __forceinline__ double " + name + "(double, double, double, double) { return 0.0; }\n";
  }
  JitCache jc; jc.setBatchedSource(src, phys, dim, order);
  // same steps as JitCache::batched(), with ptxas -v and without loading the module
  std::vector<const char*> hs, hn;
  for (int i = 0; i < k_num_embedded_headers; ++i) { hn.push_back(k_embedded_headers[i].name); hs.push_back(k_embedded_headers[i].text); }
  std::string full = jc.batchedSource();
  nvrtcProgram p; nvrtcCreateProgram(&p, full.c_str(), "probe.cu", (int)hs.size(), hs.data(), hn.data());
  const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "-DMRH_JIT=1", "--ptxas-options=-v"};
  const auto t0 = std::chrono::steady_clock::now();
  nvrtcResult rc = nvrtcCompileProgram(p, 5, opts);
  const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  size_t ls; nvrtcGetProgramLogSize(p, &ls); std::string log(ls, ' '); nvrtcGetProgramLog(p, &log[0]);
  printf("rc=%d compile %.1f s\n%s\n", (int)rc, dt, log.c_str());
  if (rc == NVRTC_SUCCESS) { size_t cs; nvrtcGetCUBINSize(p, &cs); std::vector<char> c(cs); nvrtcGetCUBIN(p, c.data());
    if (FILE* f = fopen("mrhyde_b200/build/jit_probe_batched.cubin", "wb")) { fwrite(c.data(), 1, cs, f); fclose(f); } }
  return rc != NVRTC_SUCCESS;
}
