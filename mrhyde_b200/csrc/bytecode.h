// Coefficient programs: the reference evaluates run-time string expressions with one Kokkos launch per
// DAG node over (elem, ip) views (/root/reference/src/managers/functionManager.cpp:588-760, operators
// :762-1900).  Here each expression is compiled once on the host (expr.hpp) into a postfix program that a
// thread evaluates per quadrature point inside the fused assembly kernel, on doubles or on PAD<P> lanes.
#pragma once
#if !defined(__CUDACC_RTC__)
#include <cstdint>
#endif
#include "pad.h"

namespace mrh {

enum Op : uint8_t {
  OP_CONST = 0, OP_X, OP_Y, OP_Z, OP_TIME, OP_NX, OP_NY, OP_NZ, OP_TX, OP_TY, OP_TZ, OP_FIELD,
  OP_ADD, OP_SUB, OP_MUL, OP_DIV, OP_POW, OP_LT, OP_LTE, OP_GT, OP_GTE, OP_MAX, OP_MIN, OP_MEAN,
  OP_SIN, OP_COS, OP_TAN, OP_EXP, OP_LOG, OP_SQRT, OP_SINH, OP_COSH, OP_ABS, OP_NIP
};

// OP_FIELD: arg = field id (var*8+kind).  OP_CONST: cidx = constant pool index.
// Binary ops (OP_ADD..OP_MEAN): arg == 0 -> both operands on the stack; arg = 1 + leaf opcode (OP_CONST..OP_TZ)
// -> the right operand is that leaf, fused by the host compiler (cidx = constant index for OP_CONST).
struct Instr { uint8_t op; uint8_t arg; uint16_t cidx; };

// kind: 0 constant scalar (cval), 1 depends on x/y/z/t/normals only, 2 depends on the solution state
struct Prog { int start; int len; int kind; int pad_; double cval; };

constexpr int MRH_MAXSTACK = 12;
constexpr int MRH_MAXPROG = 16;
constexpr int FIELD_KINDS = 8;   // field id = var*8 + kind; kind 0 value, 1..3 d/dx,d/dy,d/dz, 4 d/dt

template <class S>
struct PointState {             // what a program / physics routine may read at one quadrature point
  double x[3], nrm[3], tng[3], time;
  const S* fields;              // [nvar*FIELD_KINDS]
};

template <class S>
MRH_HD double leaf_value(uint8_t op, uint16_t cidx, const double* cpool, const PointState<S>& ps) {
  switch (op) {
    case OP_CONST: return cpool[cidx];
    case OP_X: return ps.x[0];
    case OP_Y: return ps.x[1];
    case OP_Z: return ps.x[2];
    case OP_TIME: return ps.time;
    case OP_NX: return ps.nrm[0];
    case OP_NY: return ps.nrm[1];
    case OP_NZ: return ps.nrm[2];
    case OP_TX: return ps.tng[0];
    case OP_TY: return ps.tng[1];
    default: return ps.tng[2];
  }
}

template <class S>
MRH_HDN inline S eval_code(const Instr* code, const double* cpool, int start, int len, const PointState<S>& ps) {
  S st[MRH_MAXSTACK];
  int sp = 0;
  for (int pc = start; pc < start + len; ++pc) {
    const Instr in = code[pc];
    if (in.op >= OP_ADD && in.op <= OP_MEAN && in.arg != 0) st[sp++] = S(leaf_value(in.arg - 1, in.cidx, cpool, ps));
    switch (in.op) {
      case OP_CONST: st[sp++] = S(cpool[in.cidx]); break;
      case OP_X: st[sp++] = S(ps.x[0]); break;
      case OP_Y: st[sp++] = S(ps.x[1]); break;
      case OP_Z: st[sp++] = S(ps.x[2]); break;
      case OP_TIME: st[sp++] = S(ps.time); break;
      case OP_NX: st[sp++] = S(ps.nrm[0]); break;
      case OP_NY: st[sp++] = S(ps.nrm[1]); break;
      case OP_NZ: st[sp++] = S(ps.nrm[2]); break;
      case OP_TX: st[sp++] = S(ps.tng[0]); break;
      case OP_TY: st[sp++] = S(ps.tng[1]); break;
      case OP_TZ: st[sp++] = S(ps.tng[2]); break;
      case OP_FIELD: st[sp++] = ps.fields[in.arg]; break;
      case OP_ADD: --sp; st[sp - 1] = st[sp - 1] + st[sp]; break;
      case OP_SUB: --sp; st[sp - 1] = st[sp - 1] + (-st[sp]); break;     // reference: data += -t
      case OP_MUL: --sp; st[sp - 1] = st[sp - 1] * st[sp]; break;
      case OP_DIV: --sp; st[sp - 1] = st[sp - 1] / st[sp]; break;
      case OP_POW: --sp; st[sp - 1] = f_pow(st[sp - 1], st[sp]); break;
      case OP_LT: --sp; st[sp - 1] = S(val(st[sp - 1]) < val(st[sp]) ? 1.0 : 0.0); break;
      case OP_LTE: --sp; st[sp - 1] = S(val(st[sp - 1]) <= val(st[sp]) ? 1.0 : 0.0); break;
      case OP_GT: --sp; st[sp - 1] = S(val(st[sp - 1]) > val(st[sp]) ? 1.0 : 0.0); break;
      case OP_GTE: --sp; st[sp - 1] = S(val(st[sp - 1]) >= val(st[sp]) ? 1.0 : 0.0); break;
      case OP_MAX: --sp; if (val(st[sp - 1]) < val(st[sp])) st[sp - 1] = st[sp]; break;
      case OP_MIN: --sp; if (val(st[sp - 1]) > val(st[sp])) st[sp - 1] = st[sp]; break;
      case OP_MEAN: --sp; st[sp - 1] = 0.5 * st[sp - 1] + 0.5 * st[sp]; break;
      case OP_SIN: st[sp - 1] = f_sin(st[sp - 1]); break;
      case OP_COS: st[sp - 1] = f_cos(st[sp - 1]); break;
      case OP_TAN: st[sp - 1] = f_tan(st[sp - 1]); break;
      case OP_EXP: st[sp - 1] = f_exp(st[sp - 1]); break;
      case OP_LOG: st[sp - 1] = f_log(st[sp - 1]); break;
      case OP_SQRT: st[sp - 1] = f_sqrt0(st[sp - 1]); break;
      case OP_SINH: st[sp - 1] = f_sinh(st[sp - 1]); break;
      case OP_COSH: st[sp - 1] = f_cosh(st[sp - 1]); break;
      case OP_ABS: if (val(st[sp - 1]) < 0.0) st[sp - 1] = -st[sp - 1]; break;
      case OP_NIP: --sp; st[sp - 1] = st[sp]; break;                       // data = t
      default: break;
    }
  }
  return st[0];
}

// Evaluate program `pr`.  Constant programs cost nothing; state-independent ones run on plain doubles.
template <class S>
MRH_HDN inline S eval_prog(const Prog& pr, const Instr* code, const double* cpool, const PointState<S>& ps) {
  if (pr.kind == 0) return S(pr.cval);
  if (pr.kind == 1) {
    PointState<double> pd;
    for (int c = 0; c < 3; ++c) { pd.x[c] = ps.x[c]; pd.nrm[c] = ps.nrm[c]; pd.tng[c] = ps.tng[c]; }
    pd.time = ps.time; pd.fields = nullptr;
    return S(eval_code<double>(code, cpool, pr.start, pr.len, pd));
  }
  return eval_code<S>(code, cpool, pr.start, pr.len, ps);
}

}  // namespace mrh

namespace mrh {

// Vector form for state-independent programs: evaluates the program at NP points in one pass, top of stack
// in registers, so instruction fetch / dispatch is paid once per NP points (specialised kernels, assemble_fast.h).
template <int NP>
MRH_HDN inline void eval_code_vec(const Instr* code, const double* cpool, int start, int len,
                                  const double (&x)[NP], const double (&y)[NP], const double (&z)[NP], double time,
                                  double (&tos)[NP]) {
  double st[MRH_MAXSTACK][NP];
  int sp = 0;     // number of live entries, the newest one held in tos
  for (int pc = start; pc < start + len; ++pc) {
    const Instr in = code[pc];
    if (in.op <= OP_FIELD) {                                   // push a leaf
      if (sp > 0) {
#pragma unroll
        for (int k = 0; k < NP; ++k) st[sp - 1][k] = tos[k];
      }
      ++sp;
      switch (in.op) {
        case OP_X:
#pragma unroll
          for (int k = 0; k < NP; ++k) tos[k] = x[k];
          break;
        case OP_Y:
#pragma unroll
          for (int k = 0; k < NP; ++k) tos[k] = y[k];
          break;
        case OP_Z:
#pragma unroll
          for (int k = 0; k < NP; ++k) tos[k] = z[k];
          break;
        default: {
          const double c = (in.op == OP_CONST) ? cpool[in.cidx] : (in.op == OP_TIME ? time : 0.0);
#pragma unroll
          for (int k = 0; k < NP; ++k) tos[k] = c;
        }
      }
    } else if (in.op <= OP_MEAN) {                             // binary: a (below) op b (top or fused leaf)
      double a[NP], b[NP];
      if (in.arg != 0) {
        const uint8_t lop = in.arg - 1;
#pragma unroll
        for (int k = 0; k < NP; ++k) {
          a[k] = tos[k];
          b[k] = (lop == OP_X) ? x[k] : (lop == OP_Y) ? y[k] : (lop == OP_Z) ? z[k] : (lop == OP_CONST) ? cpool[in.cidx] : (lop == OP_TIME) ? time : 0.0;
        }
      } else {
        --sp;
#pragma unroll
        for (int k = 0; k < NP; ++k) { b[k] = tos[k]; a[k] = st[sp - 1][k]; }
      }
#pragma unroll
      for (int k = 0; k < NP; ++k) {
        double r;
        switch (in.op) {
          case OP_ADD: r = a[k] + b[k]; break;
          case OP_SUB: r = a[k] + (-b[k]); break;
          case OP_MUL: r = a[k] * b[k]; break;
          case OP_DIV: r = a[k] / b[k]; break;
          case OP_POW: r = pow(a[k], b[k]); break;
          case OP_LT: r = a[k] < b[k] ? 1.0 : 0.0; break;
          case OP_LTE: r = a[k] <= b[k] ? 1.0 : 0.0; break;
          case OP_GT: r = a[k] > b[k] ? 1.0 : 0.0; break;
          case OP_GTE: r = a[k] >= b[k] ? 1.0 : 0.0; break;
          case OP_MAX: r = a[k] < b[k] ? b[k] : a[k]; break;
          case OP_MIN: r = a[k] > b[k] ? b[k] : a[k]; break;
          default: r = 0.5 * a[k] + 0.5 * b[k]; break;
        }
        tos[k] = r;
      }
    } else if (in.op == OP_NIP) {
      --sp;
    } else {
#pragma unroll
      for (int k = 0; k < NP; ++k) {
        const double t = tos[k];
        double r;
        switch (in.op) {
          case OP_SIN: r = sin(t); break;
          case OP_COS: r = cos(t); break;
          case OP_TAN: r = tan(t); break;
          case OP_EXP: r = exp(t); break;
          case OP_LOG: r = log(t); break;
          case OP_SQRT: r = t <= 0.0 ? 0.0 : sqrt(t); break;
          case OP_SINH: r = sinh(t); break;
          case OP_COSH: r = cosh(t); break;
          default: r = t < 0.0 ? -t : t; break;
        }
        tos[k] = r;
      }
    }
  }
}

// program value at NP points of one element (kind 0: constant, kind 1: vector interpreter)
template <int NP>
MRH_HDN inline void eval_prog_vec(const Prog& pr, const Instr* code, const double* cpool, const double (&x)[NP],
                                  const double (&y)[NP], const double (&z)[NP], double time, double (&out)[NP]) {
  if (pr.kind == 0) {
#pragma unroll
    for (int k = 0; k < NP; ++k) out[k] = pr.cval;
    return;
  }
  eval_code_vec<NP>(code, cpool, pr.start, pr.len, x, y, z, time, out);
}

}  // namespace mrh
