// Part of engine.cu: the direct caller of the hot path, on the device.
//   SolverManager::nonlinearSolver   /root/reference/src/managers/solverManager.cpp:1465-1709  (Newton loop, inf-norm test,
//                                     relative / absolute tolerance, backtracking)
//   SolverManager::transientSolver   :1229-1460 (time loop, stage and previous-step bookkeeping)
//   setButcherTableau / setBackwardDifference  :490-830
//   LinearAlgebraInterface::linearSolver  /root/reference/src/interfaces/linearAlgebraInterface.cpp (Belos GMRES + MueLu there;
//                                     here restarted GMRES(m) with classical Gram-Schmidt x2, or CG, Jacobi-preconditioned,
//                                     written as CUDA kernels over the engine's CRS graph)
// Work vectors live in the handle and are allocated at the first solve; res / jac are the caller's (Tpetra-owned) arrays.

namespace {

constexpr int SOLVER_BLOCK = 256;

__global__ void spmv_kernel(const int64_t* __restrict__ rowptr, const int* __restrict__ colind, const double* __restrict__ vals,
                            const double* __restrict__ x, double* __restrict__ y, int64_t nrows) {
  const int lane = threadIdx.x & 31;
  const int64_t wpb = blockDim.x >> 5;
  for (int64_t r = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5); r < nrows; r += (int64_t)gridDim.x * wpb) {
    const int64_t b = rowptr[r], e = rowptr[r + 1];
    double s = 0.0;
    for (int64_t p = b + lane; p < e; p += 32) s += vals[p] * x[colind[p]];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) y[r] = s;
  }
}
__device__ __forceinline__ double block_sum(double s) {
  __shared__ double part[32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) part[w] = s;
  __syncthreads();
  s = (threadIdx.x < (blockDim.x >> 5)) ? part[threadIdx.x] : 0.0;
  if (w == 0) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  }
  return s;      // valid in thread 0
}
// out[k] += V_k . w for k = blockIdx.y (V: columns of length n, leading dimension ld); plain dot with gridDim.y == 1
__global__ void multi_dot_kernel(const double* __restrict__ V, int64_t ld, const double* __restrict__ w, int64_t n, double* out) {
  const double* v = V + (int64_t)blockIdx.y * ld;
  double s = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) s += v[i] * w[i];
  s = block_sum(s);
  if (threadIdx.x == 0) atomicAdd(out + blockIdx.y, s);
}
// w += sign * sum_k coef[k] V_k, k < nk
__global__ void multi_axpy_kernel(const double* __restrict__ V, int64_t ld, int nk, const double* __restrict__ coef, double sign, double* w, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double s = 0.0;
    for (int k = 0; k < nk; ++k) s += coef[k] * V[(int64_t)k * ld + i];
    w[i] += sign * s;
  }
}
__global__ void axpby_kernel(double a, const double* __restrict__ x, double b, double* y, int64_t n) {      // y = a x + b y
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] = a * x[i] + (b == 0.0 ? 0.0 : b * y[i]);
}
__global__ void pointwise_mul_kernel(const double* __restrict__ d, const double* __restrict__ x, double* y, int64_t n) {  // y = d .* x
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] = d[i] * x[i];
}
__global__ void inv_diag_kernel(const int64_t* rowptr, const int* colind, const double* vals, int64_t nrows, double* dinv) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nrows) return;
  const int s = rel_slot(rowptr, colind, (int)r, (int)r);
  const double d = s >= 0 ? vals[rowptr[r] + s] : 0.0;
  dinv[r] = (d != 0.0) ? 1.0 / d : 1.0;
}
__global__ void norminf_kernel(const double* __restrict__ x, int64_t n, unsigned long long* out) {
  double m = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) m = fmax(m, fabs(x[i]));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMax(out, (unsigned long long)__double_as_longlong(m));     // non-negative doubles order like their bit patterns
}
// interleaved column copy: dst[i * ldd + cd] = src[i * lds + cs]
__global__ void copy_col_kernel(const double* src, int lds, int cs, double* dst, int ldd, int cd, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) dst[i * ldd + cd] = src[i * lds + cs];
}

__global__ void sub_col_kernel(const double* src, int lds, int cs, double* y, int64_t n) {       // y[i] -= src[i * lds + cs]
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] -= src[i * lds + cs];
}

struct SolverWork {
  int64_t n = 0; int m = 0;
  DevBuf<double> V, w, z, dinv, coef, scal, du, r, p, q, ustage, uprev, ucur;
  DevBuf<unsigned long long> nrm;
  int nstages_alloc = 0, nsteps_alloc = 0;
  void ensure(int64_t n_, int m_) {
    if (n != n_) { n = n_; w.alloc(n); z.alloc(n); dinv.alloc(n); du.alloc(n); r.alloc(n); p.alloc(n); q.alloc(n); m = 0; }
    if (m < m_) { m = m_; V.alloc((size_t)(m + 1) * n); coef.alloc(2 * (m + 1)); }
    if (!scal.p) { scal.alloc(8); nrm.alloc(1); }
  }
};

struct Krylov {
  const DeviceImpl& D; SolverWork& W; cudaStream_t s; int64_t n; int grid;
  Krylov(const DeviceImpl& D_, SolverWork& W_, cudaStream_t s_) : D(D_), W(W_), s(s_), n(W_.n) {
    grid = (int)std::min<int64_t>((n + SOLVER_BLOCK - 1) / SOLVER_BLOCK, (int64_t)D.num_sms * 8);
    if (grid < 1) grid = 1;
  }
  void spmv(const double* A, const double* x, double* y) const {
    const int64_t blocks = std::min<int64_t>((n + 7) / 8, (int64_t)D.num_sms * 16);
    spmv_kernel<<<(unsigned)std::max<int64_t>(blocks, 1), 256, 0, s>>>(D.rowptr.p, D.colind.p, A, x, y, n);
  }
  // host gets nk dot products V_k . w (one launch, one copy)
  void dots(const double* V, int nk, const double* w, double* host) const {
    CUDA_CK(cudaMemsetAsync(W.coef.p, 0, nk * sizeof(double), s));
    multi_dot_kernel<<<dim3((unsigned)grid, (unsigned)nk), SOLVER_BLOCK, 0, s>>>(V, n, w, n, W.coef.p);
    CUDA_CK(cudaMemcpyAsync(host, W.coef.p, nk * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_CK(cudaStreamSynchronize(s));
  }
  double dot(const double* a, const double* b) const { double v; dots(a, 1, b, &v); return v; }
  double norm2(const double* a) const { return std::sqrt(dot(a, a)); }
  double norminf(const double* a) const {
    CUDA_CK(cudaMemsetAsync(W.nrm.p, 0, sizeof(unsigned long long), s));
    norminf_kernel<<<grid, SOLVER_BLOCK, 0, s>>>(a, n, W.nrm.p);
    unsigned long long bits = 0;
    CUDA_CK(cudaMemcpyAsync(&bits, W.nrm.p, sizeof(bits), cudaMemcpyDeviceToHost, s));
    CUDA_CK(cudaStreamSynchronize(s));
    double v; memcpy(&v, &bits, sizeof(v)); return v;
  }
  void axpby(double a, const double* x, double b, double* y) const { axpby_kernel<<<grid, SOLVER_BLOCK, 0, s>>>(a, x, b, y, n); }
  void precond(const double* x, double* y) const { pointwise_mul_kernel<<<grid, SOLVER_BLOCK, 0, s>>>(W.dinv.p, x, y, n); }
  // w += sign * V[:, 0..nk) h   (h on the host)
  void combine(const double* V, int nk, const double* h, double sign, double* w) const {
    CUDA_CK(cudaMemcpyAsync(W.coef.p + (W.m + 1), h, nk * sizeof(double), cudaMemcpyHostToDevice, s));
    multi_axpy_kernel<<<grid, SOLVER_BLOCK, 0, s>>>(V, n, nk, W.coef.p + (W.m + 1), sign, w, n);
    CUDA_CK(cudaStreamSynchronize(s));          // h may live on the caller's stack
  }
};

SolverWork& solver_work(DeviceImpl& D) {      // one per handle, created at the first solve
  if (!D.solver) D.solver = std::shared_ptr<void>(new SolverWork(), [](void* p) { delete static_cast<SolverWork*>(p); });
  return *static_cast<SolverWork*>(D.solver.get());
}

// J x = b on the device; x starts at zero (the reference zeroes du before every linear solve, solverManager.cpp:1648-1649).
// Returns the iteration count; *relres = ||b - J x||_2 / ||b||_2 as estimated by the method.
int linear_solve(mrh_engine* h, const double* A, const double* b, double* x, const mrh_solver_opts& o, cudaStream_t s, double* relres) {
  EngineHost& H = h->host; DeviceImpl& D = *(DeviceImpl*)h->impl;
  const int64_t n = H.crs.nrows;
  const int m = std::max(1, std::min(o.krylov_vectors, 500));
  SolverWork& W = solver_work(D);
  W.ensure(n, o.linear_method == 1 ? 1 : m);
  Krylov K(D, W, s);
  inv_diag_kernel<<<blocks_for(n, 256), 256, 0, s>>>(D.rowptr.p, D.colind.p, A, n, W.dinv.p);
  CUDA_CK(cudaGetLastError());
  CUDA_CK(cudaMemsetAsync(x, 0, n * sizeof(double), s));
  const double bnorm = K.norm2(b);
  *relres = 0.0;
  if (bnorm == 0.0) return 0;
  const double tol = o.linear_tol;
  int iters = 0;
  if (o.linear_method == 1) {
    // ---- preconditioned conjugate gradients (symmetric positive definite Jacobians: thermal, linear elasticity) ----
    double* r = W.r.p; double* z = W.z.p; double* p = W.p.p; double* q = W.q.p;
    K.axpby(1.0, b, 0.0, r);
    K.precond(r, z);
    K.axpby(1.0, z, 0.0, p);
    double rz = K.dot(r, z), rn = bnorm;
    while (iters < o.max_linear_iters && rn / bnorm > tol) {
      K.spmv(A, p, q);
      const double pq = K.dot(p, q);
      if (!(pq > 0.0)) break;                    // not positive definite along p: give up (the caller sees relres)
      const double alpha = rz / pq;
      K.axpby(alpha, p, 1.0, x);
      K.axpby(-alpha, q, 1.0, r);
      K.precond(r, z);
      double two[2];
      CUDA_CK(cudaMemsetAsync(W.coef.p, 0, 2 * sizeof(double), s));
      multi_dot_kernel<<<dim3((unsigned)K.grid, 1), SOLVER_BLOCK, 0, s>>>(r, n, z, n, W.coef.p);
      multi_dot_kernel<<<dim3((unsigned)K.grid, 1), SOLVER_BLOCK, 0, s>>>(r, n, r, n, W.coef.p + 1);
      CUDA_CK(cudaMemcpyAsync(two, W.coef.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
      CUDA_CK(cudaStreamSynchronize(s));
      const double beta = two[0] / rz;
      rz = two[0]; rn = std::sqrt(two[1]);
      K.axpby(1.0, z, beta, p);
      ++iters;
    }
    *relres = rn / bnorm;
    return iters;
  }
  // ---- restarted GMRES(m), right preconditioning: J M^-1 y = b, x = M^-1 y; classical Gram-Schmidt applied twice ----
  double* V = W.V.p; double* w = W.w.p; double* z = W.z.p;
  std::vector<double> Hm((size_t)(m + 1) * m, 0.0), cs(m), sn(m), g(m + 1), hcol(m + 1), h2(m + 1), y(m);
  double beta = bnorm;
  K.axpby(1.0 / beta, b, 0.0, V);                                  // r0 = b (x0 = 0)
  while (iters < o.max_linear_iters) {
    std::fill(g.begin(), g.end(), 0.0); g[0] = beta;
    int j = 0;
    double res_est = beta;
    for (; j < m && iters < o.max_linear_iters; ++j) {
      K.precond(V + (size_t)j * n, z);
      K.spmv(A, z, w);
      K.dots(V, j + 1, w, hcol.data());
      K.combine(V, j + 1, hcol.data(), -1.0, w);
      K.dots(V, j + 1, w, h2.data());                              // second pass restores orthogonality lost to cancellation
      K.combine(V, j + 1, h2.data(), -1.0, w);
      for (int k = 0; k <= j; ++k) hcol[k] += h2[k];
      const double hn = K.norm2(w);
      hcol[j + 1] = hn;
      if (hn > 0.0) K.axpby(1.0 / hn, w, 0.0, V + (size_t)(j + 1) * n);
      for (int k = 0; k < j; ++k) {                                // previous Givens rotations
        const double t = cs[k] * hcol[k] + sn[k] * hcol[k + 1];
        hcol[k + 1] = -sn[k] * hcol[k] + cs[k] * hcol[k + 1]; hcol[k] = t;
      }
      const double den = std::hypot(hcol[j], hcol[j + 1]);
      cs[j] = den > 0.0 ? hcol[j] / den : 1.0; sn[j] = den > 0.0 ? hcol[j + 1] / den : 0.0;
      hcol[j] = den; hcol[j + 1] = 0.0;
      g[j + 1] = -sn[j] * g[j]; g[j] = cs[j] * g[j];
      for (int k = 0; k <= j; ++k) Hm[(size_t)k * m + j] = hcol[k];
      ++iters;
      res_est = std::fabs(g[j + 1]);
      if (res_est / bnorm <= tol || hn == 0.0) { ++j; break; }
    }
    for (int k = j - 1; k >= 0; --k) {                             // back substitution
      double t = g[k];
      for (int l = k + 1; l < j; ++l) t -= Hm[(size_t)k * m + l] * y[l];
      y[k] = t / Hm[(size_t)k * m + k];
    }
    CUDA_CK(cudaMemsetAsync(w, 0, n * sizeof(double), s));
    K.combine(V, j, y.data(), 1.0, w);                             // w = V y
    K.precond(w, z);
    K.axpby(1.0, z, 1.0, x);                                       // x += M^-1 V y
    K.spmv(A, x, w);
    K.axpby(1.0, b, -1.0, w);                                      // w = b - J x  (true residual)
    beta = K.norm2(w);
    *relres = beta / bnorm;
    if (beta / bnorm <= tol || iters >= o.max_linear_iters) break;
    K.axpby(1.0 / beta, w, 0.0, V);
  }
  return iters;
}

// SolverManager::nonlinearSolver: u is updated in place; res / jac are work arrays of the caller (overlapped vector / matrix)
int nonlinear_solve(mrh_engine* h, const mrh_state& st_in, double* u, double* res, double* jac, const mrh_solver_opts& o, mrh_solver_info& info) {
  EngineHost& H = h->host; DeviceImpl& D = *(DeviceImpl*)h->impl;
  if (!D.ready) throw std::runtime_error("mrh_finalize has not been called");
  CUDA_CK(cudaSetDevice(D.device));
  const int64_t n = H.crs.nrows; const size_t nnz = H.crs.colind.size();
  cudaStream_t s = (cudaStream_t)st_in.stream;
  SolverWork& W = solver_work(D);
  W.ensure(n, o.linear_method == 1 ? 1 : std::max(1, std::min(o.krylov_vectors, 500)));
  Krylov K(D, W, s);
  mrh_state st = st_in; st.u = u;
  double* du = W.du.p;
  CUDA_CK(cudaMemsetAsync(du, 0, n * sizeof(double), s));
  int NLiter = 0, status = 0;
  double resnorm_first = 10.0 * o.nonlinear_tol, resnorm = resnorm_first, resnorm_scaled = resnorm_first, alpha = 1.0;
  bool proceed = true;
  info.linear_iters = 0; info.converged = 0; info.last_linear_relres = 0.0;
  while (proceed) {
    // residual-only pass first (the reference's "autotune" order, solverManager.cpp:1541-1551), Jacobian only if a solve follows
    CUDA_CK(cudaMemsetAsync(res, 0, n * sizeof(double), s));
    device_assemble(h, &st, res, nullptr);
    resnorm = K.norminf(res);
    bool solve = true;
    if (NLiter == 0) { resnorm_first = resnorm; resnorm_scaled = 1.0; }
    else resnorm_scaled = resnorm / resnorm_first;
    if (o.verbosity > 1) printf("***** Iteration: %d\n***** Norm of nonlinear residual: %.12e\n***** Scaled Norm of nonlinear residual: %.12e\n", NLiter, resnorm, resnorm_scaled);
    if (o.allow_backtracking && resnorm_scaled > 1.1) {
      solve = false;
      alpha *= 0.5;
      K.axpby(-alpha, du, 1.0, u);
    } else {
      if (o.use_relative_tol) {
        if (resnorm_scaled < o.nonlinear_tol || resnorm < 1.0e-100) { solve = false; proceed = false; info.converged = 1; }
      } else if (o.use_absolute_tol && resnorm < o.nonlinear_abs_tol) { solve = false; proceed = false; info.converged = 1; }
    }
    if (solve) {
      CUDA_CK(cudaMemsetAsync(res, 0, n * sizeof(double), s));
      CUDA_CK(cudaMemsetAsync(jac, 0, nnz * sizeof(double), s));
      device_assemble(h, &st, res, jac);
      double relres = 0.0;
      info.linear_iters += linear_solve(h, jac, res, du, o, s, &relres);
      info.last_linear_relres = relres;
      alpha = 1.0;
      K.axpby(alpha, du, 1.0, u);
    }
    ++NLiter;
    if (NLiter >= o.max_nonlinear_iters) proceed = false;
  }
  CUDA_CK(cudaStreamSynchronize(s));
  info.nonlinear_iters = NLiter; info.resnorm_first = resnorm_first; info.resnorm_last = resnorm;
  if (!info.converged) status = 1;
  return status;
}

// Butcher tableaux and BDF weights by the reference's names (solverManager.cpp:490-600, 744-830)
void butcher_tableau(const std::string& name, std::vector<double>& A, std::vector<double>& b, std::vector<double>& c) {
  auto set = [&](int ns) { A.assign((size_t)ns * ns, 0.0); b.assign(ns, 0.0); c.assign(ns, 0.0); };
  if (name == "BWE" || name == "DIRK-1,1") { set(1); A[0] = 1.0; b[0] = 1.0; c[0] = 1.0; }
  else if (name == "CN") { set(2); A[2] = 0.5; A[3] = 0.5; b[0] = 0.5; b[1] = 0.5; c[1] = 1.0; }
  else if (name == "DIRK-1,2") { set(1); A[0] = 0.5; b[0] = 1.0; c[0] = 0.5; }
  else if (name == "DIRK-2,2") { set(2); A[0] = 0.25; A[2] = 0.5; A[3] = 0.25; b[0] = 0.5; b[1] = 0.5; c[0] = 0.25; c[1] = 0.75; }
  else if (name == "DIRK-2,3") { set(2); const double g = 0.5 + std::sqrt(3.0) / 6.0; A[0] = g; A[2] = -std::sqrt(3.0) / 3.0; A[3] = g; b[0] = 0.5; b[1] = 0.5; c[0] = g; c[1] = 0.5 - std::sqrt(3.0) / 6.0; }
  else if (name == "DIRK-3,3") {
    set(3); const double p = 0.4358665215;
    A[0] = p; A[3] = (1.0 - p) / 2.0; A[4] = p; A[6] = -3.0 * p * p / 2.0 + 4.0 * p - 0.25; A[7] = 3.0 * p * p / 2.0 - 5.0 * p + 1.25; A[8] = p;
    b[0] = A[6]; b[1] = A[7]; b[2] = p; c[0] = p; c[1] = (1.0 + p) / 2.0; c[2] = 1.0;
  } else throw std::runtime_error("unrecognized Butcher tableau (not supported: explicit tableaux have a zero diagonal): " + name);
}
void bdf_weights(int order, std::vector<double>& w) {
  switch (order) {
    case 1: w = {1.0, -1.0}; break;
    case 2: w = {1.5, -2.0, 0.5}; break;
    case 3: w = {11.0 / 6.0, -3.0, 1.5, -1.0 / 3.0}; break;
    case 4: w = {25.0 / 12.0, -4.0, 3.0, -4.0 / 3.0, 0.25}; break;
    default: throw std::runtime_error("BDF order must be 1..4");
  }
}

// SolverManager::transientSolver, forward branch: fixed step count, no step cutting.  u: (nlocal) in / out.
int transient_solve(mrh_engine* h, double* u, double t0, double t1, double dt, const std::string& tableau, int bdf_order,
                    double* res, double* jac, const mrh_solver_opts& o, mrh_solver_info& info, cudaStream_t s) {
  EngineHost& H = h->host; DeviceImpl& D = *(DeviceImpl*)h->impl;
  if (!D.ready) throw std::runtime_error("mrh_finalize has not been called");
  CUDA_CK(cudaSetDevice(D.device));
  const int64_t n = H.crs.nrows;
  if (bdf_order > 1 && tableau != "BWE" && tableau != "DIRK-1,1") throw std::runtime_error("need to use BWE or DIRK-1,1 with higher order BDF");   // solverManager.cpp:175-190
  std::vector<double> A, b, c, bdf;
  butcher_tableau(tableau, A, b, c);
  const int nstages = (int)b.size();
  SolverWork& W = solver_work(D);
  W.ensure(n, o.linear_method == 1 ? 1 : std::max(1, std::min(o.krylov_vectors, 500)));
  Krylov K(D, W, s);
  W.ustage.alloc((size_t)n * nstages); W.uprev.alloc((size_t)n * std::max(bdf_order, 1)); W.ucur.alloc(n); W.nsteps_alloc = 0;
  mrh_solver_info total{}; total.converged = 1;
  double t = t0; int stepProg = 0, status = 0;
  const double timetol = t1 * 1.0e-6;
  while (t < t1 - timetol) {
    // BDF start-up: order k needs k previous steps; lower orders until they exist ("transient startup BDF order")
    const int order = std::min(bdf_order, stepProg + 1);
    bdf_weights(order, bdf);
    const int nsteps = order;
    // sol_prev shifts by one and takes sol; every stage starts from sol (solverManager.cpp:1293-1313).  The engine reads the
    // arrays interleaved as (nlocal, nsteps) / (nlocal, nstages); a change of nsteps re-packs the columns that exist.
    if (stepProg > 0 && nsteps != W.nsteps_alloc) {
      DevBuf<double> tmp; tmp.alloc((size_t)n * nsteps);
      for (int k = 1; k < nsteps; ++k) copy_col_kernel<<<K.grid, SOLVER_BLOCK, 0, s>>>(W.uprev.p, W.nsteps_alloc, k - 1, tmp.p, nsteps, k, n);
      CUDA_CK(cudaStreamSynchronize(s));
      CUDA_CK(cudaMemcpyAsync(W.uprev.p, tmp.p, (size_t)n * nsteps * sizeof(double), cudaMemcpyDeviceToDevice, s));
      CUDA_CK(cudaStreamSynchronize(s));
    } else {
      for (int k = nsteps - 1; k >= 1; --k) copy_col_kernel<<<K.grid, SOLVER_BLOCK, 0, s>>>(W.uprev.p, nsteps, k - 1, W.uprev.p, nsteps, k, n);
    }
    W.nsteps_alloc = nsteps;
    copy_col_kernel<<<K.grid, SOLVER_BLOCK, 0, s>>>(u, 1, 0, W.uprev.p, nsteps, 0, n);
    for (int k = 0; k < nstages; ++k) copy_col_kernel<<<K.grid, SOLVER_BLOCK, 0, s>>>(u, 1, 0, W.ustage.p, nstages, k, n);
    for (int stage = 0; stage < nstages; ++stage) {
      mrh_state st{};
      st.is_transient = 1; st.nstages = nstages; st.nsteps = nsteps; st.stage = stage;
      st.butcher_A = A.data(); st.butcher_b = b.data(); st.bdf_wts = bdf.data();
      st.time = t + c[stage] * dt; st.deltat = dt;                         // AssemblyManager::updateStage, assemblyManager.cpp:3276-3285
      st.u_prev = W.uprev.p; st.u_stage = W.ustage.p; st.stream = (void*)s;
      double* us = (nstages == 1) ? u : W.ucur.p;                            // sol_stage[stage] (aliases sol for one stage)
      if (nstages > 1) copy_col_kernel<<<K.grid, SOLVER_BLOCK, 0, s>>>(W.ustage.p, nstages, stage, us, 1, 0, n);
      mrh_solver_info si{};
      status += nonlinear_solve(h, st, us, res, jac, o, si);
      total.nonlinear_iters += si.nonlinear_iters; total.linear_iters += si.linear_iters; total.converged = total.converged && si.converged;
      total.resnorm_first = si.resnorm_first; total.resnorm_last = si.resnorm_last; total.last_linear_relres = si.last_linear_relres;
      if (nstages > 1) {
        copy_col_kernel<<<K.grid, SOLVER_BLOCK, 0, s>>>(us, 1, 0, W.ustage.p, nstages, stage, n);
        K.axpby(1.0, us, 1.0, u);                                            // sol += sol_stage[stage] - sol_prev[0]   (:1335-1338)
        sub_col_kernel<<<K.grid, SOLVER_BLOCK, 0, s>>>(W.uprev.p, nsteps, 0, u, n);
      }
    }
    t += dt; ++stepProg;
  }
  CUDA_CK(cudaStreamSynchronize(s));
  total.time_steps = stepProg; total.final_time = t;
  info = total;
  return status;
}

}  // namespace

namespace {
// SolverManager::projectDirichlet + setDirichlet (kind 0) / setInitial (kind 1): assemble the L2-projection system into the caller's
// work arrays, solve it with conjugate gradients (it is a mass matrix plus identity rows) and put the result into u - on the fixed
// DOFs only for Dirichlet data, everywhere for initial data.
int project_into(mrh_engine* h, int kind, double time, double* u, double* rhs, double* mass, const mrh_solver_opts& o_in, mrh_solver_info& info, cudaStream_t s) {
  EngineHost& H = h->host; DeviceImpl& D = *(DeviceImpl*)h->impl;
  const int64_t n = H.crs.nrows; const size_t nnz = H.crs.colind.size();
  CUDA_CK(cudaSetDevice(D.device));
  CUDA_CK(cudaMemsetAsync(rhs, 0, n * sizeof(double), s));
  CUDA_CK(cudaMemsetAsync(mass, 0, nnz * sizeof(double), s));
  device_projection(h, kind, time, rhs, mass, s);
  mrh_solver_opts o = o_in; o.linear_method = 1;
  SolverWork& W = solver_work(D);
  W.ensure(n, 1);
  double relres = 0.0;
  info = mrh_solver_info{};
  info.linear_iters = linear_solve(h, mass, rhs, W.du.p, o, s, &relres);
  info.last_linear_relres = relres; info.converged = relres <= o.linear_tol;
  if (kind == 0) copy_fixed_kernel<<<blocks_for(n, 256), 256, 0, s>>>(D.fixed.p, W.du.p, u, n);
  else CUDA_CK(cudaMemcpyAsync(u, W.du.p, n * sizeof(double), cudaMemcpyDeviceToDevice, s));
  CUDA_CK(cudaGetLastError());
  CUDA_CK(cudaStreamSynchronize(s));
  return info.converged ? 0 : 1;
}
}  // namespace

extern "C" {

int mrh_set_dirichlet(mrh_handle h, double* u, double time, double* rhs_work, double* mass_work, const mrh_solver_opts* opts, mrh_solver_info* info, void* stream) {
  if (!h || !u || !rhs_work || !mass_work) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h)
  mrh_solver_opts o; if (opts) o = *opts; else { mrh_default_solver_opts(&o); o.linear_tol = 1.0e-12; o.max_linear_iters = 1000; }
  mrh_solver_info si{};
  const int rc = project_into(h, 0, time, u, rhs_work, mass_work, o, si, (cudaStream_t)stream);
  if (info) *info = si;
  return rc;
  MRH_CATCH_CUDA(h)
}
int mrh_set_initial(mrh_handle h, double* u, double* rhs_work, double* mass_work, const mrh_solver_opts* opts, mrh_solver_info* info, void* stream) {
  if (!h || !u || !rhs_work || !mass_work) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h)
  mrh_solver_opts o; if (opts) o = *opts; else { mrh_default_solver_opts(&o); o.linear_tol = 1.0e-12; o.max_linear_iters = 1000; }
  mrh_solver_info si{};
  const int rc = project_into(h, 1, 0.0, u, rhs_work, mass_work, o, si, (cudaStream_t)stream);
  if (info) *info = si;
  return rc;
  MRH_CATCH_CUDA(h)
}

void mrh_default_solver_opts(mrh_solver_opts* o) {
  if (!o) return;
  memset(o, 0, sizeof(*o));
  o->max_nonlinear_iters = 10; o->nonlinear_tol = 1.0e-6; o->nonlinear_abs_tol = 1.0e-6;      // solverManager.cpp:61-67
  o->use_relative_tol = 1; o->use_absolute_tol = 0; o->allow_backtracking = 1;
  o->linear_method = 0; o->max_linear_iters = 100; o->krylov_vectors = 100; o->linear_tol = 1.0e-7;   // linearAlgebraInterface.cpp:47-50
  o->verbosity = 0;
}
int mrh_linear_solve(mrh_handle h, const double* jac, const double* rhs, double* x, const mrh_solver_opts* opts, mrh_solver_info* info, void* stream) {
  if (!h || !jac || !rhs || !x) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h)
  DeviceImpl& D = *(DeviceImpl*)h->impl;
  if (!D.ready) throw std::runtime_error("mrh_finalize has not been called");
  CUDA_CK(cudaSetDevice(D.device));
  mrh_solver_opts o; if (opts) o = *opts; else mrh_default_solver_opts(&o);
  double relres = 0.0;
  const int it = linear_solve(h, jac, rhs, x, o, (cudaStream_t)stream, &relres);
  CUDA_CK(cudaStreamSynchronize((cudaStream_t)stream));
  if (info) { memset(info, 0, sizeof(*info)); info->linear_iters = it; info->last_linear_relres = relres; info->converged = relres <= o.linear_tol; }
  return MRH_OK;
  MRH_CATCH_CUDA(h)
}
int mrh_nonlinear_solve(mrh_handle h, const mrh_state* st, double* u, double* res, double* jac, const mrh_solver_opts* opts, mrh_solver_info* info) {
  if (!h || !st || !u || !res || !jac) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h)
  mrh_solver_opts o; if (opts) o = *opts; else mrh_default_solver_opts(&o);
  mrh_solver_info si{};
  const int status = nonlinear_solve(h, *st, u, res, jac, o, si);
  if (info) *info = si;
  return status;          // 0 converged, 1 not converged within max_nonlinear_iters (the reference's status)
  MRH_CATCH_CUDA(h)
}
int mrh_transient_solve(mrh_handle h, double* u, double start_time, double end_time, double deltat, const char* butcher_tableau,
                        int32_t bdf_order, double* res, double* jac, const mrh_solver_opts* opts, mrh_solver_info* info, void* stream) {
  if (!h || !u || !res || !jac || !butcher_tableau || !(deltat > 0.0)) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h)
  mrh_solver_opts o; if (opts) o = *opts; else mrh_default_solver_opts(&o);
  mrh_solver_info si{};
  const int status = transient_solve(h, u, start_time, end_time, deltat, butcher_tableau, bdf_order, res, jac, o, si, (cudaStream_t)stream);
  if (info) *info = si;
  return status > 0 ? 1 : 0;
  MRH_CATCH_CUDA(h)
}

}  // extern "C"
