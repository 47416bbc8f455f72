// Run-time specialisation of the coefficient expressions (host side).
// The reference interprets every coefficient string with one Kokkos launch per DAG node, per 100-element group
// (/root/reference/src/managers/functionManager.cpp:588-760).  Here the postfix programs produced by expr.hpp are
// printed back as straight-line CUDA C++ and compiled with NVRTC for sm_100a together with the hand-written
// specialised kernel (assemble_fast.h), so the expressions are inlined into the element's point loop.
// libnvrtc and libcuda are opened lazily with dlopen: the library still loads on a machine without a driver, and
// when NVRTC is unavailable the engine keeps using the in-kernel bytecode interpreter (still on the GPU).
#pragma once
#include <cuda.h>
#include <dlfcn.h>
#include <nvrtc.h>

#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

#include "bytecode.h"
#include "embedded_headers.inc"

namespace mrh {

struct JitApi {
  bool ok = false; std::string why;
  nvrtcResult (*CreateProgram)(nvrtcProgram*, const char*, const char*, int, const char* const*, const char* const*) = nullptr;
  nvrtcResult (*CompileProgram)(nvrtcProgram, int, const char* const*) = nullptr;
  nvrtcResult (*GetCUBINSize)(nvrtcProgram, size_t*) = nullptr;
  nvrtcResult (*GetCUBIN)(nvrtcProgram, char*) = nullptr;
  nvrtcResult (*GetProgramLogSize)(nvrtcProgram, size_t*) = nullptr;
  nvrtcResult (*GetProgramLog)(nvrtcProgram, char*) = nullptr;
  nvrtcResult (*DestroyProgram)(nvrtcProgram*) = nullptr;
  CUresult (*ModuleLoadData)(CUmodule*, const void*) = nullptr;
  CUresult (*ModuleUnload)(CUmodule) = nullptr;
  CUresult (*ModuleGetFunction)(CUfunction*, CUmodule, const char*) = nullptr;
  CUresult (*LaunchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, CUstream, void**, void**) = nullptr;
  CUresult (*OccupancyMaxActiveBlocks)(int*, CUfunction, int, size_t) = nullptr;
  CUresult (*FuncSetAttribute)(CUfunction, CUfunction_attribute, int) = nullptr;

  // which = 0: NVRTC found by soname (inside a PyTorch process the wheel-bundled 12.8) - the specialised thermal kernels
  //            were tuned against it (C2 256^3: 6.46 ms, 6.50 ms with the toolkit's compiler);
  // which = 1: the CUDA toolkit's NVRTC first - used for the batched kernel, where the 12.8 register allocation spills the
  //            software-pipelined gather state (212 bytes) and 12.9 keeps it in registers (C4 192^3: 66.5 -> 61.6 ms).
  // MRH_NVRTC_LIB overrides both.
  static JitApi& get(int which = 0) {
    static JitApi inst[2];
    static bool tried[2] = {false, false};
    JitApi& a = inst[which ? 1 : 0];
    if (tried[which ? 1 : 0]) return a;
    tried[which ? 1 : 0] = true;
    void* hn = nullptr;
    if (const char* ev = getenv("MRH_NVRTC_LIB")) hn = dlopen(ev, RTLD_NOW | RTLD_LOCAL);
    if (which) { if (!hn) hn = dlopen("/usr/local/cuda/lib64/libnvrtc.so.12", RTLD_NOW | RTLD_LOCAL); }
    for (const char* n : {"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12"}) if (!hn) hn = dlopen(n, RTLD_NOW | RTLD_LOCAL);
    void* hc = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
    if (!hn) { a.why = "libnvrtc not found"; return a; }
    if (!hc) { a.why = "libcuda not found"; return a; }
#define MRH_SYM(h, field, name) a.field = reinterpret_cast<decltype(a.field)>(dlsym(h, name)); if (!a.field) { a.why = std::string("missing symbol ") + name; return a; }
    MRH_SYM(hn, CreateProgram, "nvrtcCreateProgram") MRH_SYM(hn, CompileProgram, "nvrtcCompileProgram")
    MRH_SYM(hn, GetCUBINSize, "nvrtcGetCUBINSize") MRH_SYM(hn, GetCUBIN, "nvrtcGetCUBIN")
    MRH_SYM(hn, GetProgramLogSize, "nvrtcGetProgramLogSize") MRH_SYM(hn, GetProgramLog, "nvrtcGetProgramLog")
    MRH_SYM(hn, DestroyProgram, "nvrtcDestroyProgram")
    MRH_SYM(hc, ModuleLoadData, "cuModuleLoadData") MRH_SYM(hc, ModuleUnload, "cuModuleUnload")
    MRH_SYM(hc, ModuleGetFunction, "cuModuleGetFunction") MRH_SYM(hc, LaunchKernel, "cuLaunchKernel")
    MRH_SYM(hc, OccupancyMaxActiveBlocks, "cuOccupancyMaxActiveBlocksPerMultiprocessor")
    MRH_SYM(hc, FuncSetAttribute, "cuFuncSetAttribute")
#undef MRH_SYM
    a.ok = true;
    return a;
  }
};

// ---- postfix program -> C++ source of   static __device__ double name(double x, double y, double z, double t) ----
inline std::string jit_literal(double v) {
  char buf[64]; snprintf(buf, sizeof buf, "%.17g", v);
  std::string s(buf);
  if (s.find_first_of(".en") == std::string::npos) s += ".0";       // 'n' covers inf/nan (not valid literals anyway)
  return "(" + s + ")";
}

// fast = true: sin / cos go through mrh::trig_fast (pad.h), which clears `ok` for arguments outside its range
inline std::string jit_function_body(const Prog& pr, const std::vector<Instr>& code, const std::vector<double>& cpool, bool fast, bool* uses_trig) {
  std::string out;
  std::vector<std::string> st;
  int ntmp = 0;
  auto leaf = [&](uint8_t op, uint16_t cidx) -> std::string {
    switch (op) {
      case OP_CONST: return jit_literal(cpool[cidx]);
      case OP_X: return "x"; case OP_Y: return "y"; case OP_Z: return "z"; case OP_TIME: return "t";
      default: return "(0.0)";       // normals / tangents do not exist at volume points
    }
  };
  auto emit = [&](const std::string& rhs) { const std::string v = "v" + std::to_string(ntmp++); out += "    const double " + v + " = " + rhs + ";\n"; st.push_back(v); };
  for (int pc = pr.start; pc < pr.start + pr.len; ++pc) {
    const Instr in = code[pc];
    if (in.op == OP_FIELD) throw std::runtime_error("state-dependent coefficient cannot be specialised");
    if (in.op < OP_FIELD) { st.push_back(leaf(in.op, in.cidx)); continue; }
    if (in.op <= OP_MEAN) {
      std::string b;
      if (in.arg != 0) b = leaf(in.arg - 1, in.cidx); else { b = st.back(); st.pop_back(); }
      const std::string a = st.back(); st.pop_back();
      switch (in.op) {
        case OP_ADD: emit(a + " + " + b); break;
        case OP_SUB: emit(a + " + (-" + b + ")"); break;
        case OP_MUL: emit(a + " * " + b); break;
        case OP_DIV: emit(a + " / " + b); break;
        case OP_POW: emit("pow(" + a + ", " + b + ")"); break;
        case OP_LT: emit("(" + a + " < " + b + ") ? 1.0 : 0.0"); break;
        case OP_LTE: emit("(" + a + " <= " + b + ") ? 1.0 : 0.0"); break;
        case OP_GT: emit("(" + a + " > " + b + ") ? 1.0 : 0.0"); break;
        case OP_GTE: emit("(" + a + " >= " + b + ") ? 1.0 : 0.0"); break;
        case OP_MAX: emit("(" + a + " < " + b + ") ? " + b + " : " + a); break;
        case OP_MIN: emit("(" + a + " > " + b + ") ? " + b + " : " + a); break;
        default: emit("0.5 * " + a + " + 0.5 * " + b); break;
      }
      continue;
    }
    if (in.op == OP_NIP) { const std::string top = st.back(); st.pop_back(); st.pop_back(); st.push_back(top); continue; }
    const std::string a = st.back(); st.pop_back();
    switch (in.op) {
      case OP_SIN: *uses_trig = true; emit(fast ? "mrh::trig_fast(" + a + ", 0, ok)" : "sin(" + a + ")"); break;
      case OP_COS: *uses_trig = true; emit(fast ? "mrh::trig_fast(" + a + ", 1, ok)" : "cos(" + a + ")"); break;
      case OP_TAN: emit("tan(" + a + ")"); break;
      case OP_EXP: emit("exp(" + a + ")"); break;
      case OP_LOG: emit("log(" + a + ")"); break;
      case OP_SQRT: emit("(" + a + " <= 0.0) ? 0.0 : sqrt(" + a + ")"); break;     // functionManager.cpp:1062-1067
      case OP_SINH: emit("sinh(" + a + ")"); break;
      case OP_COSH: emit("cosh(" + a + ")"); break;
      default: emit("(" + a + " < 0.0) ? -" + a + " : " + a); break;
    }
  }
  return out + "    const double result = " + st.back() + ";\n";
}

// Expressions with sin / cos are emitted twice: a branch-free version on the fast range reduction (so that the
// compiler can interleave the polynomial chains of several calls) and an exact out-of-line version taken by the
// whole expression when any argument was outside the fast range.
inline std::string jit_function_source(const std::string& name, const Prog& pr, const std::vector<Instr>& code,
                                       const std::vector<double>& cpool) {
  const std::string sig = "(double x, double y, double z, double t) {\n";
  const std::string head = "  static __device__ __forceinline__ double " + name + sig;
  if (pr.kind == 0) return head + "    return " + jit_literal(pr.cval) + ";\n  }\n";
  if (pr.kind == 2) throw std::runtime_error("state-dependent coefficient cannot be specialised");
  bool trig = false;
  const std::string exact = jit_function_body(pr, code, cpool, false, &trig);
  if (!trig) return head + exact + "    return result;\n  }\n";
  const std::string fast = jit_function_body(pr, code, cpool, true, &trig);
  return "  static __device__ __noinline__ double " + name + "_exact" + sig + exact + "    return result;\n  }\n" +
         head + "    bool ok = true;\n" + fast + "    return ok ? result : " + name + "_exact(x, y, z, t);\n  }\n";
}

// ---- batched kernel: programs with their sin / cos calls hoisted into a table of ATOMS ----------------------------------------
// A manufactured source such as the linear-elasticity one calls sin / cos 63 times per point, 18 of them distinct; evaluated by one
// lane per point that was 42 % of the batched kernel's instruction stream (profiles/r1_batched_final_C4_ncu_summary.txt).  The printer
// therefore collects the distinct (function, argument expression) pairs of ALL programs of the launch; the kernel evaluates the atoms
// of all points with every lane busy and uniform code (affine arguments from a table row per atom - a switch over the atom index diverges -, then one trig_fast call), and the programs read
// them from shared memory.
struct JitAffine {            // c[0] + c[1] x + c[2] y + c[3] z + c[4] t, when the sub-expression is of that form
  bool ok = false; double c[5] = {0, 0, 0, 0, 0};
};
struct JitAtomTable {
  std::vector<std::string> arg; std::vector<int> shift;      // shift 0: sin, 1: cos
  std::vector<JitAffine> aff;                                 // affine form of the argument (all atoms affine: table-driven, uniform code)
  int find_or_add(const std::string& a, int sh, const JitAffine& f) {
    for (size_t i = 0; i < arg.size(); ++i) if (arg[i] == a && shift[i] == sh) return (int)i;
    arg.push_back(a); shift.push_back(sh); aff.push_back(f); return (int)arg.size() - 1;
  }
};
// program -> one fully parenthesised expression; sin / cos nodes become at[i]
inline std::string jit_expr_with_atoms(const Prog& pr, const std::vector<Instr>& code, const std::vector<double>& cpool, JitAtomTable& T) {
  std::vector<std::string> st;
  std::vector<JitAffine> sa;                       // affine form of every stack entry (ok = false: not affine)
  auto leaf = [&](uint8_t op, uint16_t cidx) -> std::string {
    switch (op) {
      case OP_CONST: return jit_literal(cpool[cidx]);
      case OP_X: return "x"; case OP_Y: return "y"; case OP_Z: return "z"; case OP_TIME: return "t";
      default: return "(0.0)";
    }
  };
  auto leaf_aff = [&](uint8_t op, uint16_t cidx) -> JitAffine {
    JitAffine f;
    switch (op) {
      case OP_CONST: f.ok = true; f.c[0] = cpool[cidx]; break;
      case OP_X: f.ok = true; f.c[1] = 1.0; break; case OP_Y: f.ok = true; f.c[2] = 1.0; break;
      case OP_Z: f.ok = true; f.c[3] = 1.0; break; case OP_TIME: f.ok = true; f.c[4] = 1.0; break;
      default: break;
    }
    return f;
  };
  auto is_const = [](const JitAffine& f) { return f.ok && f.c[1] == 0.0 && f.c[2] == 0.0 && f.c[3] == 0.0 && f.c[4] == 0.0; };
  for (int pc = pr.start; pc < pr.start + pr.len; ++pc) {
    const Instr in = code[pc];
    if (in.op == OP_FIELD) throw std::runtime_error("state-dependent coefficient cannot be specialised");
    if (in.op < OP_FIELD) { st.push_back(leaf(in.op, in.cidx)); sa.push_back(leaf_aff(in.op, in.cidx)); continue; }
    if (in.op <= OP_MEAN) {
      std::string b; JitAffine fb;
      if (in.arg != 0) { b = leaf(in.arg - 1, in.cidx); fb = leaf_aff(in.arg - 1, in.cidx); } else { b = st.back(); st.pop_back(); fb = sa.back(); sa.pop_back(); }
      const std::string a = st.back(); st.pop_back();
      const JitAffine fa = sa.back(); sa.pop_back();
      JitAffine fr;
      if (fa.ok && fb.ok) {
        if (in.op == OP_ADD || in.op == OP_SUB) { fr.ok = true; for (int i = 0; i < 5; ++i) fr.c[i] = in.op == OP_ADD ? fa.c[i] + fb.c[i] : fa.c[i] - fb.c[i]; }
        else if (in.op == OP_MUL && (is_const(fa) || is_const(fb))) { const JitAffine& k = is_const(fa) ? fa : fb; const JitAffine& v = is_const(fa) ? fb : fa; fr.ok = true; for (int i = 0; i < 5; ++i) fr.c[i] = k.c[0] * v.c[i]; }
        else if (in.op == OP_DIV && is_const(fb) && fb.c[0] != 0.0) { fr.ok = true; for (int i = 0; i < 5; ++i) fr.c[i] = fa.c[i] / fb.c[0]; }
      }
      sa.push_back(fr);
      std::string r;
      switch (in.op) {
        case OP_ADD: r = "(" + a + " + " + b + ")"; break;
        case OP_SUB: r = "(" + a + " + (-" + b + "))"; break;
        case OP_MUL: r = "(" + a + " * " + b + ")"; break;
        case OP_DIV: r = "(" + a + " / " + b + ")"; break;
        case OP_POW: r = "pow(" + a + ", " + b + ")"; break;
        case OP_LT: r = "((" + a + " < " + b + ") ? 1.0 : 0.0)"; break;
        case OP_LTE: r = "((" + a + " <= " + b + ") ? 1.0 : 0.0)"; break;
        case OP_GT: r = "((" + a + " > " + b + ") ? 1.0 : 0.0)"; break;
        case OP_GTE: r = "((" + a + " >= " + b + ") ? 1.0 : 0.0)"; break;
        case OP_MAX: r = "mrh_jmax(" + a + ", " + b + ")"; break;
        case OP_MIN: r = "mrh_jmin(" + a + ", " + b + ")"; break;
        default: r = "(0.5 * " + a + " + 0.5 * " + b + ")"; break;
      }
      st.push_back(r);
      continue;
    }
    if (in.op == OP_NIP) { const std::string top = st.back(); st.pop_back(); st.pop_back(); st.push_back(top);
                           const JitAffine ft = sa.back(); sa.pop_back(); sa.pop_back(); sa.push_back(ft); continue; }
    const std::string a = st.back(); st.pop_back();
    const JitAffine fa = sa.back(); sa.pop_back();
    sa.push_back(JitAffine());                       // the result of a unary function is not affine
    switch (in.op) {
      case OP_SIN: st.push_back("at[" + std::to_string(T.find_or_add(a, 0, fa)) + "]"); break;
      case OP_COS: st.push_back("at[" + std::to_string(T.find_or_add(a, 1, fa)) + "]"); break;
      case OP_TAN: st.push_back("tan(" + a + ")"); break;
      case OP_EXP: st.push_back("exp(" + a + ")"); break;
      case OP_LOG: st.push_back("log(" + a + ")"); break;
      case OP_SQRT: st.push_back("mrh_jsqrt0(" + a + ")"); break;
      case OP_SINH: st.push_back("sinh(" + a + ")"); break;
      case OP_COSH: st.push_back("cosh(" + a + ")"); break;
      default: st.push_back("fabs(" + a + ")"); break;
    }
  }
  return st.back();
}
// source of struct JitProgs' members for the batched kernel: NA, atom_arg, atom_shift, p0..p7(x, y, z, t, at).  An atom whose argument
// itself contains an atom (sin(cos(x))) cannot be hoisted: the caller then falls back to the plain printer (returns "").
inline std::string jit_batched_atom_source(const Prog* progs, int nprog, const std::vector<Instr>& code, const std::vector<double>& cpool, int max_atoms, int& natoms,
                                           std::string* preamble = nullptr) {
  JitAtomTable T;
  std::string fns;
  for (int k = 0; k < nprog; ++k) {
    const std::string name = "p" + std::to_string(k);
    const std::string sig = "  static __device__ __forceinline__ double " + name + "(double x, double y, double z, double t, const double* at) { return ";
    if (progs[k].kind == 1) fns += sig + jit_expr_with_atoms(progs[k], code, cpool, T) + "; }\n";
    else fns += sig + "0.0; }\n";
  }
  natoms = (int)T.arg.size();
  if (natoms > max_atoms || natoms > 64) return "";
  for (auto& a : T.arg) if (a.find("at[") != std::string::npos) return "";
  unsigned long long mask = 0;
  std::string out = "  static __device__ __forceinline__ double mrh_jmax(double a, double b) { return (a < b) ? b : a; }\n"
                    "  static __device__ __forceinline__ double mrh_jmin(double a, double b) { return (a > b) ? b : a; }\n"
                    "  static __device__ __forceinline__ double mrh_jsqrt0(double a) { return (a <= 0.0) ? 0.0 : sqrt(a); }\n"
                    "  static constexpr int NA = " + std::to_string(natoms) + ";\n";
  bool all_affine = preamble != nullptr && natoms > 0;
  for (auto& f : T.aff) all_affine = all_affine && f.ok;
  for (int i = 0; i < natoms; ++i) if (T.shift[i]) mask |= 1ull << i;
  if (all_affine) {
    // every argument is c0 + cx x + cy y + cz z + ct t: one table row per atom in global memory, the same five instructions for every
    // lane (a switch over the atom index diverges: measured 930 warp instructions per element of C4 in that switch alone)
    std::string tab = "namespace mrh { __device__ const double mrh_atom_tab[" + std::to_string(natoms * 5) + "] = {";
    for (int i = 0; i < natoms; ++i) for (int k = 0; k < 5; ++k) tab += jit_literal(T.aff[i].c[k]) + ((i == natoms - 1 && k == 4) ? "" : ", ");
    *preamble = tab + "}; }\n";
    out += "  static __device__ __forceinline__ double atom_arg(int a, double x, double y, double z, double t) {\n"
           "    const double* c = mrh_atom_tab + a * 5;\n    return fma(c[4], t, fma(c[3], z, fma(c[2], y, fma(c[1], x, c[0]))));\n  }\n";
  } else {
    if (preamble) preamble->clear();
    out += "  static __device__ __forceinline__ double atom_arg(int a, double x, double y, double z, double t) {\n    switch (a) {\n";
    for (int i = 0; i < natoms; ++i) out += "      case " + std::to_string(i) + ": return " + T.arg[i] + ";\n";
    out += "      default: return 0.0;\n    }\n  }\n";
  }
  out += "  static __device__ __forceinline__ int atom_shift(int a) { return (int)((" + std::to_string(mask) + "ull >> a) & 1ull); }\n";
  return out + fns;
}

// One NVRTC-compiled module holding the specialised thermal kernel for one (TRANSIENT, ADV, JAC) variant.
struct JitKernel {
  CUmodule mod = nullptr; CUfunction fn = nullptr, fn_ws = nullptr, fn_ws8 = nullptr, fn_wsk = nullptr; size_t ws_smem_set = 0, ws8_smem_set = 0, wsk_smem_set = 0; int blocks_per_sm = 1, tiled_blocks_per_sm = 1; size_t smem_set = 0; std::string source, log;
};

class JitCache {
 public:
  std::string last_error;
  bool tiled2d = false;        // the quadrilateral kernel is built with its tile-staged path (flag bit 8 / MRH_TILED2D=1)
  ~JitCache() { clear(); }
  void clear() {
    for (auto& kv : kernels_) if (kv.second.mod && JitApi::get().ok) JitApi::get().ModuleUnload(kv.second.mod);
    kernels_.clear(); coef_src_.clear();
    if (bat_.mod && JitApi::get().ok) JitApi::get().ModuleUnload(bat_.mod);
    bat_ = JitKernel(); bat_tried_ = false;
  }
  void setCoefficientSource(const std::string& s, int dim = 3) { clear(); coef_src_ = s; dim_ = dim; }
  bool hasSource() const { return !coef_src_.empty(); }

  // returns nullptr (and sets last_error) when the variant cannot be built
  JitKernel* thermal(bool transient, bool adv, bool jac) {
    const int key = (transient ? 4 : 0) | (adv ? 2 : 0) | (jac ? 1 : 0);
    auto it = kernels_.find(key);
    if (it != kernels_.end()) return it->second.fn ? &it->second : nullptr;
    JitKernel& K = kernels_[key];
    JitApi& api = JitApi::get();
    if (!api.ok) { last_error = api.why; return nullptr; }
    if (coef_src_.empty()) { last_error = "no coefficient source"; return nullptr; }
    auto B = [](bool b) { return b ? "true" : "false"; };
    if (dim_ == 2) {          // Q1 quadrilaterals (assemble_fast2d.h): tile-staged scatter when the launch carries a tile plan, direct scatter otherwise
      K.source = std::string("#include \"assemble_fast2d.h\"\nnamespace mrh {\nstruct JitCoefs {\n  static constexpr bool staged = false;\n") + coef_src_ +
          "};\n}\nextern \"C\" __global__ void __launch_bounds__(128, MRH_JIT_MINB2D) mrh_jit_thermal_hex_q1(const __grid_constant__ mrh::KParams kp) {\n"
          "  typedef mrh::ThermalQuadQ1<" + B(transient) + ", " + B(adv) + ", " + B(jac) + ", mrh::JitCoefs> Kern;\n" +
          (jac ? "#if defined(MRH_TILED2D_ON)\n  extern __shared__ double dsm[];\n  if (kp.tile.ntiles > 0) { Kern::tiled(kp, dsm); return; }\n#endif\n" : "") +
          "#if defined(MRH_Q2D_PIPE)\n  Kern::run_direct(kp);\n#else\n"      // software-pipelined gather: measured slower (2.01 / 1.53 ms against 1.60 / 1.37 ms at 160 / 128 registers)
          "  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < kp.nwork; e += (long long)gridDim.x * blockDim.x)\n"
          "    Kern::element(kp, e);\n#endif\n}\n";
    } else {
    K.source = std::string("#include \"assemble_fast.h\"\nnamespace mrh {\nstruct JitCoefs {\n  static constexpr bool staged = false;\n") + coef_src_ +
        "};\n}\nextern \"C\" __global__ void __launch_bounds__(128, MRH_JIT_MINB) mrh_jit_thermal_hex_q1(const __grid_constant__ mrh::KParams kp) {\n"
        "  typedef mrh::ThermalHexQ1<" + B(transient) + ", " + B(adv) + ", " + B(jac) + ", mrh::JitCoefs> Kern;\n"
        "  extern __shared__ double dsm[];\n" +
        (jac ? "#if !defined(MRH_DEBUG_KSMEM_COMPUTE)\n  if (kp.tile.ntiles > 0) { Kern::tiled(kp, nullptr, 0, dsm); return; }\n#endif\n" : "") +
        "  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < kp.nwork; e += (long long)gridDim.x * blockDim.x)\n"
        "    Kern::element(kp, e, nullptr, 0);\n}\n";
    if (jac)          // warp-specialised variant: 384 threads, one CTA per SM (assemble_fast.h tiled_ws)
      K.source += std::string("extern \"C\" __global__ void __launch_bounds__(384, 1) mrh_jit_thermal_hex_q1_ws(const __grid_constant__ mrh::KParams kp) {\n") +
          "  typedef mrh::ThermalHexQ1<" + B(transient) + ", " + B(adv) + ", true, mrh::JitCoefs> Kern;\n"
          "  extern __shared__ double dsm[];\n  Kern::tiled_ws(kp, dsm);\n}\n";
    if (jac)          // K-in-registers variant (assemble_fast.h tiled_wsk)
      K.source += std::string("extern \"C\" __global__ void __launch_bounds__(384, 1) mrh_jit_thermal_hex_q1_wsk(const __grid_constant__ mrh::KParams kp) {\n") +
          "  typedef mrh::ThermalHexQ1<" + B(transient) + ", " + B(adv) + ", true, mrh::JitCoefs> Kern;\n"
          "  extern __shared__ double dsm[];\n  Kern::tiled_wsk(kp, dsm);\n}\n";
    if (jac && !adv)  // lane-per-point variant: 128 (1 + MRH_WS8_NCG) threads, one CTA per SM (assemble_fast.h tiled_ws8)
      K.source += std::string("extern \"C\" __global__ void __launch_bounds__(128 * (1 + MRH_WS8_NCG), 1) mrh_jit_thermal_hex_q1_ws8(const __grid_constant__ mrh::KParams kp) {\n") +
          "  typedef mrh::ThermalHexQ1<" + B(transient) + ", false, true, mrh::JitCoefs> Kern;\n"
          "  extern __shared__ double dsm[];\n  Kern::tiled_ws8(kp, dsm);\n}\n";
    }
    std::vector<const char*> hdr_src, hdr_name;
    for (int i = 0; i < k_num_embedded_headers; ++i) { hdr_name.push_back(k_embedded_headers[i].name); hdr_src.push_back(k_embedded_headers[i].text); }
    nvrtcProgram prog;
    if (api.CreateProgram(&prog, K.source.c_str(), "mrh_jit_thermal.cu", (int)hdr_src.size(), hdr_src.data(), hdr_name.data()) != NVRTC_SUCCESS) { last_error = "nvrtcCreateProgram failed"; return nullptr; }
    std::string minb = "-DMRH_JIT_MINB=2";                                                          // CTAs of 128 threads per SM the register allocation targets
    if (const char* ev = getenv("MRH_JIT_MINB")) minb = std::string("-DMRH_JIT_MINB=") + ev;          // tuning experiments
    std::string minb2d = "-DMRH_JIT_MINB2D=4";                                                      // same for the quadrilateral kernel: 128 registers, 16 warps per SM (1.37 ms at 4096^2 against 1.60 ms at 160 registers / 12 warps)
    if (const char* ev = getenv("MRH_JIT_MINB2D")) minb2d = std::string("-DMRH_JIT_MINB2D=") + ev;    // tuning experiments
    std::string extra = "-DMRH_JIT_EXTRA=0";
    if (const char* ev = getenv("MRH_JIT_DEFINE")) extra = std::string("-D") + ev;                    // tuning experiments
    // slot / row lists of a tile staged by two bulk copies (TMA unit) instead of per-thread 4-byte cp.async: measured 6.28 vs 6.46 ms (C2)
    std::string qun = "-DMRH_WS_QUNROLL=2";                                                          // two points per trip of the hexahedron's point loop: 6.27 -> 6.16 ms (C2); 4: 6.79
    if (const char* ev = getenv("MRH_WS_QUNROLL")) qun = std::string("-DMRH_WS_QUNROLL=") + ev;
    std::string t2d = "-DMRH_TILED2D_OFF=1";                                                         // the quadrilateral kernel carries its tile-staged path only when asked to
    if (tiled2d) t2d = "-DMRH_TILED2D_ON=1";
    std::string bulk = "-DMRH_WS_BULK=1";
    if (const char* ev = getenv("MRH_WS_BULK")) { if (atoi(ev) == 0) bulk = "-DMRH_WS_NOBULK=1"; }
    std::string creg = "-DMRH_WS_CREG=208", sreg = "-DMRH_WS_SREG=88";                              // 256 * CREG + 128 * SREG <= 384 * 168 (208 / 88: 6.16 ms, 200 / 96: 6.18, 216 / 64: 6.46)
    if (const char* ev = getenv("MRH_WS_CREG")) creg = std::string("-DMRH_WS_CREG=") + ev;
    if (const char* ev = getenv("MRH_WS_SREG")) sreg = std::string("-DMRH_WS_SREG=") + ev;
    ws8_ncg = 3;                                                                                    // compute warpgroups of the lane-per-point kernel
    if (const char* ev = getenv("MRH_WS8_NCG")) ws8_ncg = std::max(1, std::min(3, atoi(ev)));
    std::string ncg8 = "-DMRH_WS8_NCG=" + std::to_string(ws8_ncg);
    // register budgets: 128 SREG + 128 NCG CREG <= 65536 / 1 CTA
    std::string creg8 = std::string("-DMRH_WS8_CREG=") + (ws8_ncg == 3 ? "136" : ws8_ncg == 2 ? "200" : "232"), sreg8 = "-DMRH_WS8_SREG=96";
    if (const char* ev = getenv("MRH_WS8_CREG")) creg8 = std::string("-DMRH_WS8_CREG=") + ev;
    if (const char* ev = getenv("MRH_WS8_SREG")) sreg8 = std::string("-DMRH_WS8_SREG=") + ev;
    std::string cregk = "-DMRH_WSK_CREG=216", sregk = "-DMRH_WSK_SREG=64";                          // 256 * 216 + 128 * 64 = 63488 <= 384 * 168 (the CTA's pool: a larger sum blocks setmaxnreg.inc forever)
    if (const char* ev = getenv("MRH_WSK_CREG")) cregk = std::string("-DMRH_WSK_CREG=") + ev;
    if (const char* ev = getenv("MRH_WSK_SREG")) sregk = std::string("-DMRH_WSK_SREG=") + ev;
    const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "-DMRH_JIT=1", minb.c_str(), extra.c_str(), creg.c_str(), sreg.c_str(),
                          ncg8.c_str(), creg8.c_str(), sreg8.c_str(), cregk.c_str(), sregk.c_str(), minb2d.c_str(), bulk.c_str(), t2d.c_str(), qun.c_str()};
    const nvrtcResult rc = api.CompileProgram(prog, 17, opts);
    size_t ls = 0; api.GetProgramLogSize(prog, &ls);
    if (ls > 1) { K.log.resize(ls); api.GetProgramLog(prog, &K.log[0]); }
    if (rc != NVRTC_SUCCESS) { last_error = "NVRTC compile failed: " + K.log.substr(0, 2000); api.DestroyProgram(&prog); return nullptr; }
    size_t cs = 0; api.GetCUBINSize(prog, &cs);
    std::vector<char> cubin(cs);
    api.GetCUBIN(prog, cubin.data());
    api.DestroyProgram(&prog);
    if (api.ModuleLoadData(&K.mod, cubin.data()) != CUDA_SUCCESS) { last_error = "cuModuleLoadData failed"; K.mod = nullptr; return nullptr; }
    if (api.ModuleGetFunction(&K.fn, K.mod, "mrh_jit_thermal_hex_q1") != CUDA_SUCCESS) { last_error = "kernel symbol not found"; K.fn = nullptr; return nullptr; }
    if (api.OccupancyMaxActiveBlocks(&K.blocks_per_sm, K.fn, 128, 0) != CUDA_SUCCESS || K.blocks_per_sm < 1) K.blocks_per_sm = 1;
    if (!jac || dim_ == 2 || api.ModuleGetFunction(&K.fn_ws, K.mod, "mrh_jit_thermal_hex_q1_ws") != CUDA_SUCCESS) K.fn_ws = nullptr;
    if (!jac || dim_ == 2 || api.ModuleGetFunction(&K.fn_wsk, K.mod, "mrh_jit_thermal_hex_q1_wsk") != CUDA_SUCCESS) K.fn_wsk = nullptr;
    if (!jac || adv || dim_ == 2 || api.ModuleGetFunction(&K.fn_ws8, K.mod, "mrh_jit_thermal_hex_q1_ws8") != CUDA_SUCCESS) K.fn_ws8 = nullptr;
    return &K;
  }

  // ---- batched volume kernel (assemble_batched.h) with the state-independent programs printed as functions p0..p7 ----
  // threads / minb: launch bounds of the kernel (minb > 0: register budget 65536 / (threads * minb); 0: ptxas' own choice)
  int bat_natoms = 0;          // sin / cos atoms hoisted out of the batched kernel's programs (0: plain printer)
  void setBatchedSource(const std::string& progs, int phys, int dim, int order, int threads = 192, int minb = 0, int natoms = 0, const std::string& preamble = "") {
    bat_natoms = natoms;
    bat_bounds_ = "-DMRH_BAT_BOUNDS=" + std::to_string(threads) + (minb > 0 ? "," + std::to_string(minb) : std::string()); bat_max_threads = threads;
    if (const char* ev = getenv("MRH_BAT_BOUNDS")) { bat_bounds_ = std::string("-DMRH_BAT_BOUNDS=") + ev; bat_max_threads = atoi(ev); }   // tuning experiments
    if (bat_.mod && JitApi::get().ok) JitApi::get().ModuleUnload(bat_.mod);
    bat_ = JitKernel(); bat_tried_ = false;
    bat_.source = std::string("#include \"assemble_batched.h\"\n") + preamble + "namespace mrh {\nstruct JitProgs {\n  static constexpr bool jit = true;\n" + progs +
        (natoms > 0 ?
        "  static __device__ __forceinline__ double eval(int k, double x, double y, double z, double t, const double* at) {\n    switch (k) {\n"
        "      case 0: return p0(x, y, z, t, at); case 1: return p1(x, y, z, t, at); case 2: return p2(x, y, z, t, at); case 3: return p3(x, y, z, t, at);\n"
        "      case 4: return p4(x, y, z, t, at); case 5: return p5(x, y, z, t, at); case 6: return p6(x, y, z, t, at); default: return p7(x, y, z, t, at);\n    }\n  }\n};\n}\n"
        :
        "  static constexpr int NA = 0;\n"
        "  static __device__ __forceinline__ double atom_arg(int, double, double, double, double) { return 0.0; }\n"
        "  static __device__ __forceinline__ int atom_shift(int) { return 0; }\n"
        "  static __device__ __forceinline__ double eval(int k, double x, double y, double z, double t, const double*) {\n    switch (k) {\n"
        "      case 0: return p0(x, y, z, t); case 1: return p1(x, y, z, t); case 2: return p2(x, y, z, t); case 3: return p3(x, y, z, t);\n"
        "      case 4: return p4(x, y, z, t); case 5: return p5(x, y, z, t); case 6: return p6(x, y, z, t); default: return p7(x, y, z, t);\n    }\n  }\n};\n}\n") +
        "extern \"C\" __global__ void __launch_bounds__(MRH_BAT_BOUNDS) mrh_jit_batched(const __grid_constant__ mrh::KParams kp) {\n"
        "  extern __shared__ __align__(16) double smem_all[];\n"
        "  mrh::batched_body<mrh::GenCfg<" + std::to_string(phys) + ", " + std::to_string(dim) + ", " + std::to_string(order) + ">, mrh::JitProgs, false>(kp, smem_all);\n}\n";       // volume launches only: side branches compiled away
  }
  bool hasBatchedSource() const { return !bat_.source.empty(); }
  const std::string& batchedSource() const { return bat_.source; }
  // returns nullptr (and sets last_error) when the kernel cannot be built; the build is attempted once
  JitKernel* batched() {
    if (bat_tried_) return bat_.fn ? &bat_ : nullptr;
    bat_tried_ = true;
    JitApi& api = JitApi::get(1);
    if (!api.ok) { last_error = api.why; return nullptr; }
    if (bat_.source.empty()) { last_error = "no program source"; return nullptr; }
    std::vector<const char*> hdr_src, hdr_name;
    for (int i = 0; i < k_num_embedded_headers; ++i) { hdr_name.push_back(k_embedded_headers[i].name); hdr_src.push_back(k_embedded_headers[i].text); }
    nvrtcProgram prog;
    if (api.CreateProgram(&prog, bat_.source.c_str(), "mrh_jit_batched.cu", (int)hdr_src.size(), hdr_src.data(), hdr_name.data()) != NVRTC_SUCCESS) { last_error = "nvrtcCreateProgram failed"; return nullptr; }
    std::string bat_extra = "-DMRH_BAT_EXTRA=0";
    if (const char* ev = getenv("MRH_BAT_DEFINE")) bat_extra = std::string("-D") + ev;               // tuning experiments
    const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "-DMRH_JIT=1", bat_bounds_.c_str(), bat_extra.c_str()};
    const nvrtcResult rc = api.CompileProgram(prog, 6, opts);
    size_t ls = 0; api.GetProgramLogSize(prog, &ls);
    if (ls > 1) { bat_.log.resize(ls); api.GetProgramLog(prog, &bat_.log[0]); }
    if (rc != NVRTC_SUCCESS) { last_error = "NVRTC compile failed: " + bat_.log.substr(0, 2000); api.DestroyProgram(&prog); return nullptr; }
    size_t cs = 0; api.GetCUBINSize(prog, &cs);
    std::vector<char> cubin(cs);
    api.GetCUBIN(prog, cubin.data());
    api.DestroyProgram(&prog);
    if (api.ModuleLoadData(&bat_.mod, cubin.data()) != CUDA_SUCCESS) { last_error = "cuModuleLoadData failed"; bat_.mod = nullptr; return nullptr; }
    if (api.ModuleGetFunction(&bat_.fn, bat_.mod, "mrh_jit_batched") != CUDA_SUCCESS) { last_error = "kernel symbol not found"; bat_.fn = nullptr; return nullptr; }
    return &bat_;
  }

 private:
  JitKernel bat_;
  std::string bat_bounds_ = "-DMRH_BAT_BOUNDS=192";
 public:
  int bat_max_threads = 192;
  int ws8_ncg = 3;
 private:
  bool bat_tried_ = false;
  std::map<int, JitKernel> kernels_;
  std::string coef_src_;
  int dim_ = 3;
};

}  // namespace mrh
