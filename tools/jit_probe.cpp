// compile-only check (no GPU needed): the NVRTC translation unit of the specialised thermal kernel builds for sm_100a
#include <nvrtc.h>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include "../mrhyde_b200/csrc/expr.hpp"
#include "../mrhyde_b200/csrc/jit.hpp"
using namespace mrh;
int main() {
  ExprCompiler ec; ec.varlist = {"e"}; ec.dim = 3;
  ec.addFunction("source", "12*(pi*pi)*sin(2*pi*x)*sin(2*pi*y)*sin(2*pi*z)", "ip");
  ec.addFunction("diff", "1.0+x*y", "ip"); ec.addFunction("one", "1.0", "ip"); ec.addFunction("w", "max(x,0.5*t)^2-abs(z)", "ip");
  std::string src;
  src += jit_function_source("source", ec.compile("source", "ip"), ec.code, ec.cpool);
  src += jit_function_source("diff", ec.compile("diff", "ip"), ec.code, ec.cpool);
  src += jit_function_source("rho_", ec.compile("one", "ip"), ec.code, ec.cpool);
  src += jit_function_source("cp_", ec.compile("w", "ip"), ec.code, ec.cpool);
  src += "  static __device__ __forceinline__ double rhocp(double x, double y, double z, double t) { return rho_(x, y, z, t) * cp_(x, y, z, t); }\n";
  for (const char* n : {"bx", "by", "bz"}) src += jit_function_source(n, ec.compile("one", "ip"), ec.code, ec.cpool);
  printf("%s\n", src.c_str());
  const char* minb_env = getenv("MRH_JIT_MINB");
  const std::string minb = std::string("-DMRH_JIT_MINB=") + (minb_env ? minb_env : "2");
  std::string full = std::string("#include \"assemble_fast.h\"\nnamespace mrh {\nstruct JitCoefs {\n  static constexpr bool staged = false;\n") + src +
      "};\n}\nextern \"C\" __global__ void __launch_bounds__(128, MRH_JIT_MINB) k(const __grid_constant__ mrh::KParams kp) {\n"
      "  typedef mrh::ThermalHexQ1<false, false, true, mrh::JitCoefs> Kern;\n"
      "  extern __shared__ double dsm[];\n"
      "#if !defined(MRH_DEBUG_KSMEM_COMPUTE)\n  if (kp.tile.ntiles > 0) { Kern::tiled(kp, nullptr, 0, dsm); return; }\n#endif\n"
      "  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < kp.nwork; e += (long long)gridDim.x * blockDim.x)\n"
      "    Kern::element(kp, e, nullptr, 0);\n}\n"
      "extern \"C\" __global__ void __launch_bounds__(384, 1) kws(const __grid_constant__ mrh::KParams kp) {\n"
      "  typedef mrh::ThermalHexQ1<false, false, true, mrh::JitCoefs> Kern;\n"
      "  extern __shared__ double dsm[];\n  Kern::tiled_ws(kp, dsm);\n}\n"
      "extern \"C\" __global__ void __launch_bounds__(384, 1) kwsk(const __grid_constant__ mrh::KParams kp) {\n"
      "  typedef mrh::ThermalHexQ1<MRH_PROBE_TRANSIENT, false, true, mrh::JitCoefs> Kern;\n"
      "  extern __shared__ double dsm[];\n  Kern::tiled_wsk(kp, dsm);\n}\n"
      "extern \"C\" __global__ void __launch_bounds__(128 * (1 + MRH_WS8_NCG), 1) kws8(const __grid_constant__ mrh::KParams kp) {\n"
      "  typedef mrh::ThermalHexQ1<MRH_PROBE_TRANSIENT, false, true, mrh::JitCoefs> Kern;\n"
      "  extern __shared__ double dsm[];\n  Kern::tiled_ws8(kp, dsm);\n}\n";
  const char* extra_env = getenv("MRH_JIT_DEFINE"); const std::string extra = std::string("-D") + (extra_env ? extra_env : "MRH_JIT_EXTRA=0");
  const char* creg_env = getenv("MRH_WS_CREG"); const char* sreg_env = getenv("MRH_WS_SREG");
  const std::string creg = std::string("-DMRH_WS_CREG=") + (creg_env ? creg_env : "200"), sreg = std::string("-DMRH_WS_SREG=") + (sreg_env ? sreg_env : "96");
  if (getenv("MRH_PROBE_2D"))
    full = std::string("#include \"assemble_fast2d.h\"\nnamespace mrh {\nstruct JitCoefs {\n  static constexpr bool staged = false;\n") + src +
      "};\n}\nextern \"C\" __global__ void __launch_bounds__(128, MRH_JIT_MINB) k(const __grid_constant__ mrh::KParams kp) {\n"
      "  typedef mrh::ThermalQuadQ1<false, false, true, mrh::JitCoefs> Kern;\n"
      "#if defined(MRH_TILED2D_ON)\n  extern __shared__ double dsm[];\n  if (kp.tile.ntiles > 0) { Kern::tiled(kp, dsm); return; }\n#endif\n"
      "#if defined(MRH_Q2D_PIPE)\n  Kern::run_direct(kp);\n#else\n"
      "  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < kp.nwork; e += (long long)gridDim.x * blockDim.x)\n"
      "    Kern::element(kp, e);\n#endif\n}\n";
  std::vector<const char*> hs, hn;
  for (int i = 0; i < k_num_embedded_headers; ++i) { hn.push_back(k_embedded_headers[i].name); hs.push_back(k_embedded_headers[i].text); }
  nvrtcProgram p; nvrtcCreateProgram(&p, full.c_str(), "probe.cu", (int)hs.size(), hs.data(), hn.data());
  auto envor = [](const char* n, const char* d) { const char* e = getenv(n); return std::string(e ? e : d); };
  const std::string ncg8 = "-DMRH_WS8_NCG=" + envor("MRH_WS8_NCG", "3"), creg8 = "-DMRH_WS8_CREG=" + envor("MRH_WS8_CREG", "136"),
                    sreg8 = "-DMRH_WS8_SREG=" + envor("MRH_WS8_SREG", "96"), ptr = "-DMRH_PROBE_TRANSIENT=" + envor("MRH_PROBE_TRANSIENT", "false");
  const std::string cregk = "-DMRH_WSK_CREG=" + envor("MRH_WSK_CREG", "216"), sregk = "-DMRH_WSK_SREG=" + envor("MRH_WSK_SREG", "80");
  const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "-DMRH_JIT=1", "--ptxas-options=-v", minb.c_str(), creg.c_str(), sreg.c_str(), extra.c_str(),
                        ncg8.c_str(), creg8.c_str(), sreg8.c_str(), ptr.c_str(), cregk.c_str(), sregk.c_str()};
  nvrtcResult rc = nvrtcCompileProgram(p, 15, opts);
  size_t ls; nvrtcGetProgramLogSize(p, &ls); std::string log(ls, ' '); nvrtcGetProgramLog(p, &log[0]);
  printf("rc=%d\n%s\n", (int)rc, log.c_str());
  if (rc == NVRTC_SUCCESS) { size_t cs; nvrtcGetCUBINSize(p, &cs); std::vector<char> c(cs); nvrtcGetCUBIN(p, c.data()); FILE* f = fopen("/tmp/jit_probe.cubin", "wb"); fwrite(c.data(), 1, cs, f); fclose(f); printf("cubin %zu bytes\n", cs); }
  return rc != NVRTC_SUCCESS;
}
