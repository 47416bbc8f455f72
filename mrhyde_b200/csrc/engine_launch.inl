// Part of engine.cu: device set-up, launch logic and the GPU-side C-ABI entry points.

// SMs the interior launch leaves free for the exchange kernels when it overlaps them (MRH_XCH_SMS overrides).  Measured on 2 B200s
// (256^3 per GPU): 4 reserved SMs 6.35 ms, 2: 6.27, 1: 6.27, 0: 6.20 (one GPU: 6.16) - every reserved SM costs 1/148 of the interior's
// throughput, while with none the pack / NCCL / unpack kernels slip in as the persistent CTAs of the interior launch retire.
static const int XCH_RESERVE_SMS = [] { const char* ev = getenv("MRH_XCH_SMS"); return ev ? std::max(0, atoi(ev)) : 0; }();
#if !defined(MRH_WSK_DEFAULT)
#define MRH_WSK_DEFAULT 0      // K-in-registers thermal kernel: MRH_WSK=1 enables it
#endif
#if !defined(MRH_TILED2D_DEFAULT)
#define MRH_TILED2D_DEFAULT 0  // tile-staged scatter for the Q1 quadrilateral kernel: measured slower than the direct scatter (1.72 vs 1.60 ms at 4096^2), MRH_TILED2D=1 enables it
#endif
#if !defined(MRH_WS8_DEFAULT)
#define MRH_WS8_DEFAULT 0      // lane-per-point thermal kernel off until it is measured faster (MRH_WS8=1 enables it)
#endif

namespace {

RefTab upload_tab(TabDev& D, const PointTables& T) {
  D.B.upload(T.B); D.G.upload(T.G); D.VB.upload(T.VB); D.VG.upload(T.VG); D.W.upload(T.W); D.FR.upload(T.FR);
  RefTab r{};
  r.FR = D.FR.p; r.B = D.B.p; r.G = D.G.p; r.VB = D.VB.p; r.VG = D.VG.p; r.W = D.W.p; r.nq = T.nq;
  for (int c = 0; c < 3; ++c) { r.tanU[c] = T.tanU[c]; r.tanV[c] = T.tanV[c]; }
  return r;
}

void build_patterns(EngineHost& H, DeviceImpl& D) {
  const int64_t ne = H.mesh.nelem; const int N = H.N;
  DevBuf<uint64_t> hash, skey; DevBuf<int> idx, bad;
  hash.alloc(ne); skey.alloc(ne); idx.alloc(ne); bad.alloc(1);
  CUDA_CK(cudaMemset(bad.p, 0, sizeof(int)));
  pattern_hash_kernel<<<blocks_for(ne, 128), 128>>>(D.lids.p, D.rowptr.p, D.colind.p, ne, N, hash.p, bad.p);
  CUDA_CK(cudaGetLastError());
  CUDA_CK(cudaMemcpy(skey.p, hash.p, ne * sizeof(uint64_t), cudaMemcpyDeviceToDevice));
  thrust::device_ptr<uint64_t> kp(skey.p); thrust::device_ptr<int> ip(idx.p);
  thrust::sequence(thrust::device, ip, ip + ne);
  thrust::sort_by_key(thrust::device, kp, kp + ne, ip);
  auto ends = thrust::unique_by_key(thrust::device, kp, kp + ne, ip);
  const int64_t npat = ends.first - kp;
  int hbad = 0; CUDA_CK(cudaMemcpy(&hbad, bad.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (hbad) throw std::runtime_error("CRS graph misses an element coupling (row/column of an element's LIDs not in the graph)");
  D.pat.alloc((size_t)npat * N * N); D.pat_id.alloc(ne);
  pattern_fill_kernel<<<blocks_for(npat * N, 128), 128>>>(D.lids.p, D.rowptr.p, D.colind.p, idx.p, npat, N, D.pat.p);
  CUDA_CK(cudaGetLastError());
  pattern_assign_kernel<<<blocks_for(ne, 128), 128>>>(D.lids.p, D.rowptr.p, D.colind.p, ne, N, hash.p, skey.p, npat, D.pat.p, D.pat_id.p, bad.p);
  CUDA_CK(cudaGetLastError());
  CUDA_CK(cudaMemcpy(&hbad, bad.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (hbad) throw std::runtime_error("slot-pattern dictionary verification failed (hash collision)");
  D.npatterns = npat;
}


void device_finalize(mrh_engine* h) {
  EngineHost& H = h->host; DeviceImpl& D = *(DeviceImpl*)h->impl;
  CUDA_CK(cudaSetDevice(D.device));
  H.finalizeHost();
  D.ready = false; D.chunk_ready = false;
  D.nodes.upload(H.mesh.nodes); D.e2n.upload(H.mesh.elem2node); D.lids.upload(H.dofs.lids); D.fixed.upload(H.isFixed);
  D.rowptr.upload(H.crs.rowptr); D.colind.upload(H.crs.colind);
  D.code.upload(H.ec.code); D.cpool.upload(H.ec.cpool);
  D.bp = BasePtrs();
  D.bp.vol_tab = upload_tab(D.vol_tab, H.ref.vol);
  for (int s = 0; s < H.ref.nsides; ++s) D.bp.side_tab[s] = upload_tab(D.side_tab[s], H.ref.side[s]);
  D.side_elems.clear();
  for (auto& S : H.mesh.sides) { D.side_elems.emplace_back(new DevBuf<int>()); D.side_elems.back()->upload(S.elems); }
  build_patterns(H, D);
  h->npatterns = D.npatterns;
  // tile plan of the tile-staged scatter (engine_tiles.inl)
  if (!D.tiles) D.tiles = new TileDev();
  D.tiles->release();
  if (!(H.cfg.flags & 5) && thermal_hex_q1_eligible(H.phys, H.dim, H.order, H.ref.vol.nq, H.volumeStateIndependent(), H.cfg.lump_mass)) {
    const std::string why = build_tile_plan<8>(H.tileOrder(&D.iface_tiles), H.mesh.nelem, D.lids.p, D.fixed.p, D.rowptr.p, D.colind.p,
                                               H.cfg.include_advection ? 64 : 36, *D.tiles);
    if (!D.tiles->ok) h->err = "tile-staged scatter unavailable (" + why + "), using direct scatter";
  }
  static const int tiled2d_env = [] { const char* ev = getenv("MRH_TILED2D"); return ev ? atoi(ev) : MRH_TILED2D_DEFAULT; }();
  D.jit.tiled2d = (tiled2d_env || (H.cfg.flags & 256)) != 0;
  if (D.jit.tiled2d && !(H.cfg.flags & 5) && thermal_quad_q1_eligible(H.phys, H.dim, H.order, H.ref.vol.nq, H.volumeStateIndependent(), H.cfg.lump_mass)) {
    const std::string why = build_tile_plan<4>(H.tileOrder(nullptr), H.mesh.nelem, D.lids.p, D.fixed.p, D.rowptr.p, D.colind.p,
                                               H.cfg.include_advection ? 16 : 10, *D.tiles);
    if (!D.tiles->ok) h->err = "tile-staged scatter unavailable (" + why + "), using direct scatter";
  }
  // NVRTC specialisation of the coefficient expressions for the specialised thermal kernel (jit.hpp)
  D.jit.clear();
  const bool quad2d = thermal_quad_q1_eligible(H.phys, H.dim, H.order, H.ref.vol.nq, H.volumeStateIndependent(), H.cfg.lump_mass);
  if (!(H.cfg.flags & 3) && (quad2d || thermal_hex_q1_eligible(H.phys, H.dim, H.order, H.ref.vol.nq, H.volumeStateIndependent(), H.cfg.lump_mass))) {
    try {
      std::string src;
      src += jit_function_source("source", H.vol_progs[TH_SOURCE], H.ec.code, H.ec.cpool);
      src += jit_function_source("diff", H.vol_progs[TH_DIFF], H.ec.code, H.ec.cpool);
      src += jit_function_source("rho_", H.vol_progs[TH_RHO], H.ec.code, H.ec.cpool);
      src += jit_function_source("cp_", H.vol_progs[TH_CP], H.ec.code, H.ec.cpool);
      src += "  static __device__ __forceinline__ double rhocp(double x, double y, double z, double t) { return rho_(x, y, z, t) * cp_(x, y, z, t); }\n";
      src += jit_function_source("bx", H.vol_progs[TH_BX], H.ec.code, H.ec.cpool);
      src += jit_function_source("by", H.vol_progs[TH_BY], H.ec.code, H.ec.cpool);
      src += jit_function_source("bz", H.vol_progs[TH_BZ], H.ec.code, H.ec.cpool);
      D.jit.setCoefficientSource(src, quad2d ? 2 : 3);
      const bool adv = H.cfg.include_advection != 0;
      if (!D.jit.thermal(false, adv, true) || !D.jit.thermal(false, adv, false)) h->err = "NVRTC specialisation unavailable, using the in-kernel interpreter: " + D.jit.last_error;
    } catch (const std::exception& ex) { D.jit.clear(); h->err = std::string("NVRTC specialisation skipped: ") + ex.what(); }
  }
  // NVRTC specialisation of the state-independent programs for the batched volume kernel (assemble_batched.h): pays off
  // on large meshes only (the compile takes seconds), flag bit 5 forces it (tests)
  {
    const bool thermal_fast = !(H.cfg.flags & 1) && (quad2d || thermal_hex_q1_eligible(H.phys, H.dim, H.order, H.ref.vol.nq, H.volumeStateIndependent(), H.cfg.lump_mass));
    bool any1 = false;
    for (int k = 0; k < MRH_NCOEF; ++k) any1 = any1 || H.vol_progs[k].kind == 1;
    if (!thermal_fast && any1 && !(H.cfg.flags & (2 | 16)) && (H.mesh.nelem >= 32768 || (H.cfg.flags & 32))) {
      try {
        // sin / cos calls hoisted into a lane-parallel atom table when they fit the kernel's scratch (the smallest D region of a
        // point is 33 doubles); MRH_BAT_ATOMS=0 keeps the plain printer
        int natoms = 0;
        // MRH_BAT_ATOMS: 0 off, 1 on, default automatic - the programs are evaluated by one lane per point, so hoisting pays when few
        // lanes are busy there: measured with the table-driven atoms, C4 (8 points, 18 atoms) 59.4 -> 53.8 ms, C3 (27 points, 6 atoms)
        // 45.4 -> 48.3 ms
        static const int atoms_env = [] { const char* ev = getenv("MRH_BAT_ATOMS"); return ev ? atoi(ev) : -1; }();
        const bool want_atoms = atoms_env > 0 || (atoms_env < 0 && H.ref.vol.nq <= 16);
        std::string preamble;
        std::string src = want_atoms ? jit_batched_atom_source(H.vol_progs, MRH_NCOEF, H.ec.code, H.ec.cpool, 32, natoms, &preamble) : std::string();
        if (atoms_env < 0 && natoms < 4) { src.clear(); natoms = 0; }           // too few distinct calls to be worth the extra pass
        if (src.empty() || natoms == 0) {
          natoms = 0; src.clear();
          for (int k = 0; k < MRH_NCOEF; ++k) {
            const std::string name = "p" + std::to_string(k);
            if (H.vol_progs[k].kind == 1) src += jit_function_source(name, H.vol_progs[k], H.ec.code, H.ec.cpool);
            else src += "  static __device__ __forceinline__ double " + name + "(double, double, double, double) { return 0.0; }\n";
          }
        }
        D.jit.setBatchedSource(src, H.phys, H.dim, H.order, 192, 0, natoms, natoms > 0 ? preamble : std::string());
        if (!D.jit.batched()) h->err = "NVRTC specialisation of the batched kernel unavailable, using the in-kernel interpreter: " + D.jit.last_error;
      } catch (const std::exception& ex) { h->err = std::string("NVRTC specialisation of the batched kernel skipped: ") + ex.what(); }
    }
  }
  D.dbc.upload(H.dbc_dofs); D.dbc_slots.alloc(H.dbc_dofs.size());
  if (!H.dbc_dofs.empty()) {
    dbc_slot_kernel<<<blocks_for((int64_t)H.dbc_dofs.size(), 128), 128>>>(D.dbc.p, (int64_t)H.dbc_dofs.size(), D.rowptr.p, D.colind.p, D.dbc_slots.p);
    CUDA_CK(cudaGetLastError());
  }
  CUDA_CK(cudaDeviceSynchronize());
  D.bp.nodes = D.nodes.p; D.bp.e2n = D.e2n.p; D.bp.lids = D.lids.p; D.bp.fixed = D.fixed.p; D.bp.rowptr = D.rowptr.p;
  D.bp.pat = D.pat.p; D.bp.pat_id = D.pat_id.p; D.bp.code = D.code.p; D.bp.cpool = D.cpool.p;
  if (!D.ev0) { CUDA_CK(cudaEventCreate(&D.ev0)); CUDA_CK(cudaEventCreate(&D.ev1)); }
  if (D.xch && H.plan.complete) exchange_upload(H, *D.xch);
  D.static_scratch = static_scratch_of(D.device);      // set-up time, this handle's device current (cudaSetDevice above)
  D.ready = true;
}

template <bool T, bool A, bool J>
void launch_thermal_fast(const DeviceImpl& D, const KParams& kp, cudaStream_t stream) {
  auto kern = thermal_hex_q1_kernel<T, A, J>;
  int per_sm = 1;
  CUDA_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, 0));
  if (per_sm < 1) per_sm = 1;
  const int64_t nblk = (kp.nwork + 127) / 128;
  const unsigned grid = (unsigned)std::min<int64_t>(nblk, (int64_t)D.num_sms * per_sm * 4);
  kern<<<grid, 128, 0, stream>>>(kp);
  CUDA_CK(cudaGetLastError());
}

template <bool T, bool A, bool J>
void launch_thermal_quad(const DeviceImpl& D, const KParams& kp, cudaStream_t stream) {
  auto kern = thermal_quad_q1_kernel<T, A, J>;
  int per_sm = 1;
  CUDA_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, 0));
  if (per_sm < 1) per_sm = 1;
  const int64_t nblk = (kp.nwork + 127) / 128;
  const unsigned grid = (unsigned)std::min<int64_t>(nblk, (int64_t)D.num_sms * per_sm * 4);
  kern<<<grid, 128, 0, stream>>>(kp);
  CUDA_CK(cudaGetLastError());
}

template <bool T, bool A>
void launch_thermal_quad_tiled(const DeviceImpl& D, const KParams& kp, size_t acc_bytes, cudaStream_t stream) {
  auto kern = thermal_quad_q1_tiled_kernel<T, A>;
  if (acc_bytes > 48 * 1024) CUDA_CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)acc_bytes));
  int per_sm = 1;
  CUDA_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, acc_bytes));
  if (per_sm < 1) per_sm = 1;
  const unsigned grid = (unsigned)std::min<int64_t>(kp.tile.ntiles, (int64_t)D.num_sms * per_sm * 4);
  kern<<<grid, 128, acc_bytes, stream>>>(kp);
  CUDA_CK(cudaGetLastError());
}

template <bool T, bool A, bool J>
void launch_thermal_tiled(const DeviceImpl& D, const KParams& kp, size_t acc_bytes, cudaStream_t stream) {
  if constexpr (!J) { (void)D; (void)kp; (void)acc_bytes; (void)stream; throw std::runtime_error("internal: tiled scatter requested for a residual-only pass"); } else {
  auto kern = thermal_hex_q1_tiled_kernel<T, A>;
  CUDA_CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)acc_bytes));
  int per_sm = 1;
  CUDA_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, acc_bytes));
  if (per_sm < 1) per_sm = 1;
  const unsigned grid = (unsigned)std::min<int64_t>(kp.tile.ntiles, (int64_t)D.num_sms * per_sm);
  kern<<<grid, 128, acc_bytes, stream>>>(kp);
  CUDA_CK(cudaGetLastError());
  }
}

// Chunks of consecutive tiles and, for each, the end of the CRS value prefix whose rows no later chunk touches
// (host, once per finalize).  On first-encounter numberings rows complete roughly in element order, so the prefix
// grows with every chunk; on other numberings the plan degenerates to one copy at the end.
void build_chunk_plan(EngineHost& H, DeviceImpl& D) {
  if (D.chunk_ready) return;
  D.chunk_ready = true;
  D.chunk_tile_end.clear(); D.chunk_val_end.clear();
  if (!D.tiles || !D.tiles->ok) return;
  const int64_t nt = D.tiles->ntiles;
  const int64_t nchunk = std::max<int64_t>(1, std::min<int64_t>(16, nt / (8 * (int64_t)D.num_sms)));
  if (nchunk < 2) return;
  const std::vector<int> perm = H.tileOrder();
  const int64_t nrows = H.crs.nrows, nelem = H.mesh.nelem;
  std::vector<int> last((size_t)nrows, 0);
  for (int64_t c = 0; c < nchunk; ++c) {
    const int64_t tb = nt * c / nchunk, te = nt * (c + 1) / nchunk;
    D.chunk_tile_end.push_back(te);
    for (int64_t k = tb * 128; k < std::min<int64_t>(te * 128, nelem); ++k) {
      const int* l = H.dofs.lids.data() + (size_t)perm[k] * 8;
      for (int i = 0; i < 8; ++i) last[l[i]] = (int)c;
    }
  }
  // rows below R_c are complete after chunk c: first row whose running maximum of `last` exceeds c
  int64_t r = 0; int run = 0;
  for (int64_t c = 0; c < nchunk; ++c) {
    while (r < nrows && std::max(run, last[r]) <= c) { run = std::max(run, last[r]); ++r; }
    D.chunk_val_end.push_back(c + 1 == nchunk ? (int64_t)H.crs.colind.size() : H.crs.rowptr[r]);
  }
  if (!D.copy_stream) CUDA_CK(cudaStreamCreateWithFlags(&D.copy_stream, cudaStreamNonBlocking));
  while ((int64_t)D.chunk_ev.size() < nchunk) { cudaEvent_t e; CUDA_CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); D.chunk_ev.push_back(e); }
}

// Copy-out plan of the host-buffer entry point: the volume launch is cut into chunks of consecutive tiles and the
// CRS values of all rows completed by chunk k are copied to the host while later chunks are computed.
struct HostCopy {
  double* jac_host = nullptr;
  bool used = false;
};

int device_assemble(mrh_engine* h, const mrh_state* st, double* res, double* jac, HostCopy* hc = nullptr) {
  EngineHost& H = h->host; DeviceImpl& D = *(DeviceImpl*)h->impl;
  if (!D.ready) throw std::runtime_error("mrh_finalize has not been called");
  CUDA_CK(cudaSetDevice(D.device));
  cudaStream_t stream = (cudaStream_t)st->stream;
  D.launches = 0;
  // multi-GPU: with a communicator attached the call also folds the rows of DOFs owned by another rank onto their owner
  // (exportVector/MatrixFromOverlapped); flag bit 7 leaves that to an explicit mrh_exchange
  const bool xfuse = D.xch && D.xch->ready && !(H.cfg.flags & 128) && !hc;
  bool exchanged = false, dbc_first = false, any_side_launch = false;
  for (auto& sl : H.side_launches) any_side_launch = any_side_launch || sl.active;
  if (xfuse && jac) {          // fixed rows are never scattered into: their unit diagonal can be written first, so that it travels with the rows
    const int64_t nd = (int64_t)H.dbc_dofs.size();
    if (nd > 0) { dbc_diag_kernel<<<blocks_for(nd, 256), 256, 0, stream>>>(D.dbc_slots.p, nd, jac); CUDA_CK(cudaGetLastError()); D.launches++; }
    dbc_first = true;
  }
  for (int side = -1; side < (int)H.side_launches.size(); ++side) {
    if (side >= 0 && !H.side_launches[side].active) continue;
    KParams kp;
    H.fillParams(kp, D.bp, *st, res, jac, side, side >= 0 ? D.side_elems[side]->p : nullptr);
    if (kp.tp.transient && (!st->u_prev || (kp.tp.stage > 0 && !st->u_stage))) throw std::runtime_error("transient assembly needs u_prev (and u_stage for stage > 0)");
    if (side < 0) CUDA_CK(cudaEventRecord(D.ev0, stream));
    const bool fast = side < 0 && !(H.cfg.flags & 1) &&
        thermal_hex_q1_eligible(H.phys, H.dim, H.order, H.ref.vol.nq, H.volumeStateIndependent(), H.cfg.lump_mass);
    const bool fast2d = side < 0 && !(H.cfg.flags & 1) &&
        thermal_quad_q1_eligible(H.phys, H.dim, H.order, H.ref.vol.nq, H.volumeStateIndependent(), H.cfg.lump_mass);
    if (fast2d) {
      const bool tr = kp.tp.transient != 0, adv = H.cfg.include_advection != 0, jc = jac != nullptr;
      JitKernel* jk = (H.cfg.flags & 2) ? nullptr : D.jit.thermal(tr, adv, jc);
      D.jit_used = jk != nullptr;
      const bool tiled = jc && D.tiles && D.tiles->ok && D.tiles->nn == 4 && D.tiles->nk == (adv ? 16 : 10);      // residual-only passes scatter directly
      if (tiled) fill_tile_params(*D.tiles, kp.tile);
      const size_t acc_bytes = !tiled ? 0 : adv ? ThermalQuadQ1<false, true, true>::tiled_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r)
                                                : ThermalQuadQ1<false, false, true>::tiled_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r);
      D.last_path = jk ? (tiled ? "nvrtc-specialised-2d+tiled" : "nvrtc-specialised-2d") : (tiled ? "specialised-2d+tiled" : "specialised-2d");
      if (jk) {
        const int64_t nblk = (kp.nwork + 127) / 128;
        unsigned grid = (unsigned)std::min<int64_t>(nblk, (int64_t)D.num_sms * jk->blocks_per_sm * 4);
        if (tiled) {
          if (jk->smem_set != acc_bytes) {
            JitApi& api = JitApi::get();
            if (acc_bytes > 48 * 1024 && api.FuncSetAttribute(jk->fn, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)acc_bytes) != CUDA_SUCCESS)
              throw std::runtime_error("CUDA error: cuFuncSetAttribute(shared memory) failed");
            if (api.OccupancyMaxActiveBlocks(&jk->tiled_blocks_per_sm, jk->fn, 128, acc_bytes) != CUDA_SUCCESS || jk->tiled_blocks_per_sm < 1) jk->tiled_blocks_per_sm = 1;
            jk->smem_set = acc_bytes;
          }
          grid = (unsigned)std::min<int64_t>(D.tiles->ntiles, (int64_t)D.num_sms * jk->tiled_blocks_per_sm * 4);
        }
        void* args[] = {(void*)&kp};
        if (JitApi::get().LaunchKernel(jk->fn, grid, 1, 1, 128, 1, 1, (unsigned)acc_bytes, (CUstream)stream, args, nullptr) != CUDA_SUCCESS)
          throw std::runtime_error("CUDA error: cuLaunchKernel failed for the NVRTC-specialised 2-D kernel");
      } else if (tiled) {
#define MRH_FAST2D(T, A) if (tr == T && adv == A) launch_thermal_quad_tiled<T, A>(D, kp, acc_bytes, stream);
        MRH_FAST2D(false, false) MRH_FAST2D(false, true) MRH_FAST2D(true, false) MRH_FAST2D(true, true)
#undef MRH_FAST2D
      } else {
#define MRH_FAST2D(T, A, Jc) if (tr == T && adv == A && jc == Jc) launch_thermal_quad<T, A, Jc>(D, kp, stream);
        MRH_FAST2D(false, false, false) MRH_FAST2D(false, false, true) MRH_FAST2D(false, true, false) MRH_FAST2D(false, true, true)
        MRH_FAST2D(true, false, false) MRH_FAST2D(true, false, true) MRH_FAST2D(true, true, false) MRH_FAST2D(true, true, true)
#undef MRH_FAST2D
      }
    } else if (fast) {
      const bool tr = kp.tp.transient != 0, adv = H.cfg.include_advection != 0, jc = jac != nullptr;
      JitKernel* jk = (H.cfg.flags & 2) ? nullptr : D.jit.thermal(tr, adv, jc);
      D.jit_used = jk != nullptr;
      const bool tiled = jc && D.tiles && D.tiles->ok && D.tiles->nk == (adv ? 64 : 36);       // residual-only passes scatter directly
      if (tiled) fill_tile_params(*D.tiles, kp.tile);
      const size_t acc_bytes = !tiled ? 0 : adv ? ThermalHexQ1<false, true, true>::tiled_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r)
                                                : ThermalHexQ1<false, false, true>::tiled_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r);
      D.last_path = jk ? (tiled ? "nvrtc-specialised+tiled" : "nvrtc-specialised") : (tiled ? "specialised+tiled" : "specialised");
      // warp-specialised variant of the tiled kernel (384 threads, one CTA per SM) when its shared memory fits
      const size_t ws_bytes = !tiled ? 0 : adv ? ThermalHexQ1<false, true, true>::ws_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r)
                                               : ThermalHexQ1<false, false, true>::ws_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r);
      const bool ws = tiled && jk && jk->fn_ws && !(H.cfg.flags & 8) && ws_bytes <= (size_t)D.max_smem_optin;
      // lane-per-point variant (tiled_ws8): symmetric element matrix only; flag bit 6 or MRH_WS8=0 switch it off
      const int ncg8 = D.jit.ws8_ncg;
      const size_t ws8_bytes = (!tiled || adv) ? 0 : tr ? ThermalHexQ1<true, false, true>::ws8_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r, ncg8)
                                                        : ThermalHexQ1<false, false, true>::ws8_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r, ncg8);
      static const int ws8_env = [] { const char* ev = getenv("MRH_WS8"); return ev ? atoi(ev) : MRH_WS8_DEFAULT; }();
      const bool ws8 = ws && !adv && jk->fn_ws8 && !(H.cfg.flags & 64) && ws8_env != 0 && ws8_bytes <= (size_t)D.max_smem_optin;
      // K-in-registers variant (tiled_wsk): MRH_WSK=1 / default per MRH_WSK_DEFAULT
      const size_t wsk_bytes = !tiled ? 0 : adv ? (tr ? ThermalHexQ1<true, true, true>::wsk_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r)
                                                      : ThermalHexQ1<false, true, true>::wsk_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r))
                                                : (tr ? ThermalHexQ1<true, false, true>::wsk_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r)
                                                      : ThermalHexQ1<false, false, true>::wsk_smem_bytes(D.tiles->max_acc, D.tiles->max_e, D.tiles->max_r));
      static const int wsk_env = [] { const char* ev = getenv("MRH_WSK"); return ev ? atoi(ev) : MRH_WSK_DEFAULT; }();
      const bool wsk = ws && !ws8 && jk->fn_wsk && wsk_env != 0 && wsk_bytes <= (size_t)D.max_smem_optin;
      if (ws) {
        JitApi& api = JitApi::get();
        const CUfunction fnk = ws8 ? jk->fn_ws8 : wsk ? jk->fn_wsk : jk->fn_ws;
        const size_t kbytes = ws8 ? ws8_bytes : wsk ? wsk_bytes : ws_bytes;
        const int ncg = ws8 ? ncg8 : 2, kthreads = ws8 ? 128 * (1 + ncg8) : 384;
        size_t& set_bytes = ws8 ? jk->ws8_smem_set : wsk ? jk->wsk_smem_set : jk->ws_smem_set;
        if (set_bytes != kbytes) {
          if (api.FuncSetAttribute(fnk, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)kbytes) != CUDA_SUCCESS)
            throw std::runtime_error("CUDA error: cuFuncSetAttribute(shared memory) failed");
          set_bytes = kbytes;
        }
        D.last_path = ws8 ? "nvrtc-specialised+tiled-ws8" : wsk ? "nvrtc-specialised+tiled-wsk" : "nvrtc-specialised+tiled-ws";
        const bool any_side = any_side_launch;
        const bool chunked = hc && !any_side && !H.cfg.fix_zero_rows && D.chunk_tile_end.size() > 1;
        // interface tiles first, their rows leave on the exchange stream while the interior tiles are assembled; the interior
        // launch leaves XCH_RESERVE_SMS SMs to the pack / NCCL / unpack kernels (a persistent CTA fills a whole SM)
        const bool overlap = xfuse && !any_side && D.iface_tiles > 0 && D.iface_tiles < D.tiles->ntiles;
        const size_t nchunk = chunked ? D.chunk_tile_end.size() : overlap ? 2 : 1;
        if (chunked) {            // fixed rows are never scattered into: their unit diagonal can be written first
          const int64_t nd = (int64_t)H.dbc_dofs.size();
          if (nd > 0) { dbc_diag_kernel<<<blocks_for(nd, 256), 256, 0, stream>>>(D.dbc_slots.p, nd, jac); CUDA_CK(cudaGetLastError()); D.launches++; }
          hc->used = true;
        }
        int64_t tb = 0, vb = 0;
        for (size_t c = 0; c < nchunk; ++c) {
          const int64_t te = chunked ? D.chunk_tile_end[c] : (overlap && c == 0) ? D.iface_tiles : D.tiles->ntiles;
          kp.tile.tile_begin = tb; kp.tile.tile_end = te;
          const int64_t sms = (overlap && c == 1) ? std::max(1, D.num_sms - XCH_RESERVE_SMS) : D.num_sms;
          const unsigned grid = (unsigned)std::min<int64_t>((te - tb + ncg - 1) / ncg, sms);
          void* args[] = {(void*)&kp};
          if (grid > 0 && api.LaunchKernel(fnk, grid, 1, 1, kthreads, 1, 1, (unsigned)kbytes, (CUstream)stream, args, nullptr) != CUDA_SUCCESS)
            throw std::runtime_error("CUDA error: cuLaunchKernel failed for the warp-specialised NVRTC kernel");
          if (chunked) {
            const int64_t ve = D.chunk_val_end[c];
            CUDA_CK(cudaEventRecord(D.chunk_ev[c], stream));
            CUDA_CK(cudaStreamWaitEvent(D.copy_stream, D.chunk_ev[c], 0));
            if (ve > vb) CUDA_CK(cudaMemcpyAsync(hc->jac_host + vb, jac + vb, (size_t)(ve - vb) * sizeof(double), cudaMemcpyDeviceToHost, D.copy_stream));
            vb = ve;
            if (c > 0) D.launches++;
          }
          if (overlap && c == 0) {
            ExchangeDev& X = *D.xch;
            CUDA_CK(cudaEventRecord(X.ev_iface, stream));
            CUDA_CK(cudaStreamWaitEvent(X.xstream, X.ev_iface, 0));
            exchange_run(D, X, res, jac, X.xstream);
            CUDA_CK(cudaEventRecord(X.ev_done, X.xstream));
            D.launches += X.launches + 1;
            exchanged = true;
          }
          tb = te;
        }
      } else if (jk) {
        const int64_t nblk = (kp.nwork + 127) / 128;
        unsigned grid = (unsigned)std::min<int64_t>(nblk, (int64_t)D.num_sms * jk->blocks_per_sm * 4);
        if (tiled) {
          if (jk->smem_set != acc_bytes) {
            JitApi& api = JitApi::get();
            if (acc_bytes > 48 * 1024 && api.FuncSetAttribute(jk->fn, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)acc_bytes) != CUDA_SUCCESS)
              throw std::runtime_error("CUDA error: cuFuncSetAttribute(shared memory) failed");
            if (api.OccupancyMaxActiveBlocks(&jk->tiled_blocks_per_sm, jk->fn, 128, acc_bytes) != CUDA_SUCCESS || jk->tiled_blocks_per_sm < 1) jk->tiled_blocks_per_sm = 1;
            jk->smem_set = acc_bytes;
          }
          grid = (unsigned)std::min<int64_t>(D.tiles->ntiles, (int64_t)D.num_sms * jk->tiled_blocks_per_sm);
        }
        void* args[] = {(void*)&kp};
        if (JitApi::get().LaunchKernel(jk->fn, grid, 1, 1, 128, 1, 1, (unsigned)acc_bytes, (CUstream)stream, args, nullptr) != CUDA_SUCCESS)
          throw std::runtime_error("CUDA error: cuLaunchKernel failed for the NVRTC-specialised kernel");
      } else {
#define MRH_FAST(T, A, Jc) if (tr == T && adv == A && jc == Jc) { if (tiled) launch_thermal_tiled<T, A, Jc>(D, kp, acc_bytes, stream); else launch_thermal_fast<T, A, Jc>(D, kp, stream); }
      MRH_FAST(false, false, false) MRH_FAST(false, false, true) MRH_FAST(false, true, false) MRH_FAST(false, true, true)
      MRH_FAST(true, false, false) MRH_FAST(true, false, true) MRH_FAST(true, true, false) MRH_FAST(true, true, true)
#undef MRH_FAST
      }
    } else {
      // volume and boundary terms (residual + Jacobian, or residual only): batched kernel (assemble_batched.h); element blocks
      // that do not fit its shared-memory layout go through the generic kernel
      bool done = false, jitted = false;
      if (!(H.cfg.flags & 16)) {          // side launches (boundary terms) take the ahead-of-time batched kernel as well
        JitKernel* jk = (side < 0 && D.jit.hasBatchedSource()) ? D.jit.batched() : nullptr;
        if (jk) {
          JitApi& api = JitApi::get(1);
          const bool tr = kp.tp.transient != 0;
          if (jk->smem_set != (size_t)kp.tab.nq * 2 + tr + 1) {     // launch shape of this module: once per (quadrature rule, transient)
            dispatch_cfg(H.phys, H.dim, H.order, [&](auto c) {
              using C = decltype(c);
              const size_t want = std::min<size_t>(batched_smem<C>(kp.tab.nq, tr, MRH_BATCHED_MAXW), (size_t)227 * 1024);
              if (api.FuncSetAttribute(jk->fn, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)want) != CUDA_SUCCESS)
                throw std::runtime_error("CUDA error: cuFuncSetAttribute(shared memory) failed");
              jk->tiled_blocks_per_sm = batched_pick_warps<C>(kp.tab.nq, tr, [&](int threads, size_t smem) {
                if (threads > D.jit.bat_max_threads) return 0;
                int b = 0; return api.OccupancyMaxActiveBlocks(&b, jk->fn, threads, smem) == CUDA_SUCCESS ? b : 0; }, jk->blocks_per_sm);
              jk->ws_smem_set = batched_smem<C>(kp.tab.nq, tr, jk->tiled_blocks_per_sm);
            });
            jk->smem_set = (size_t)kp.tab.nq * 2 + tr + 1;
          }
          const int nw = jk->tiled_blocks_per_sm;                  // (field reused: warps per CTA of the batched module)
          if (nw > 0) {
            kp.elems_per_block = nw;
            const unsigned grid = (unsigned)std::min<int64_t>((kp.nwork + nw - 1) / nw, (int64_t)D.num_sms * jk->blocks_per_sm);
            void* args[] = {(void*)&kp};
            if (api.LaunchKernel(jk->fn, grid, 1, 1, 32 * nw, 1, 1, (unsigned)jk->ws_smem_set, (CUstream)stream, args, nullptr) != CUDA_SUCCESS)
              throw std::runtime_error("CUDA error: cuLaunchKernel failed for the NVRTC-specialised batched kernel");
            done = jitted = true;
          }
        }
        if (!done) done = launch_batched_assembly(H.phys, H.dim, H.order, D.num_sms, kp, (void*)stream);
      }
      if (side < 0) D.last_path = jitted ? "nvrtc-batched" : done ? "batched" : "generic";
      if (!done) {
        const bool ok = launch_generic_assembly(H.phys, H.dim, H.order, jac != nullptr, D.num_sms, kp, (void*)stream);
        if (!ok) throw std::runtime_error("configuration not supported: no kernel for this physics/dimension/order");
      }
    }
    if (side < 0) CUDA_CK(cudaEventRecord(D.ev1, stream));
    D.launches++;
  }
  if (xfuse) {
    if (exchanged) CUDA_CK(cudaStreamWaitEvent(stream, D.xch->ev_done, 0));
    else { exchange_run(D, *D.xch, res, jac, stream); D.launches += D.xch->launches; }
  }
  if (jac && !(hc && hc->used)) {
    const int64_t nd = (int64_t)H.dbc_dofs.size();
    if (nd > 0 && !dbc_first) { dbc_diag_kernel<<<blocks_for(nd, 256), 256, 0, stream>>>(D.dbc_slots.p, nd, jac); CUDA_CK(cudaGetLastError()); D.launches++; }
    if (H.cfg.fix_zero_rows) {
      fix_zero_rows_kernel<<<blocks_for(H.crs.nrows, 256), 256, 0, stream>>>(D.rowptr.p, D.colind.p, H.crs.nrows, jac);
      CUDA_CK(cudaGetLastError()); D.launches++;
    }
  }
  return MRH_OK;
}

// L2-projection systems: kind 0 = strong Dirichlet data on the fixed DOFs (setDirichlet), 1 = initial data
// (setInitial).  rhs (nlocal) and mass (nnz) are accumulated into; the caller zeroes them.
int device_projection(mrh_engine* h, int kind, double time, double* rhs, double* mass, cudaStream_t stream) {
  EngineHost& H = h->host; DeviceImpl& D = *(DeviceImpl*)h->impl;
  if (!D.ready) throw std::runtime_error("mrh_finalize has not been called");
  if (!H.proj_unsupported.empty()) throw std::runtime_error(H.proj_unsupported + " (not supported)");
  CUDA_CK(cudaSetDevice(D.device));
  D.launches = 0;
  if (kind == 0) {
    for (size_t k = 0; k < H.dirichlet_launches.size(); ++k) {
      KParams kp;
      H.fillProjectionParams(kp, D.bp, time, rhs, mass, (int)k, D.side_elems[H.dirichlet_launches[k].side]->p);
      if (!launch_generic_projection(H.phys, H.dim, H.order, D.num_sms, kp, (void*)stream)) throw std::runtime_error("configuration not supported: no kernel for this physics/dimension/order");
      D.launches++;
    }
    if (mass) {
      projection_identity_kernel<<<blocks_for(H.crs.nrows, 256), 256, 0, stream>>>(D.rowptr.p, D.colind.p, D.fixed.p, H.crs.nrows, mass);
      CUDA_CK(cudaGetLastError()); D.launches++;
    }
  } else {
    KParams kp;
    H.fillProjectionParams(kp, D.bp, time, rhs, mass, -1, nullptr);
    if (!launch_generic_projection(H.phys, H.dim, H.order, D.num_sms, kp, (void*)stream)) throw std::runtime_error("configuration not supported: no kernel for this physics/dimension/order");
    D.launches++;
    if (mass) {            // setInitial always fixes zero rows (assemblyManager.cpp:1286-1306)
      fix_zero_rows_kernel<<<blocks_for(H.crs.nrows, 256), 256, 0, stream>>>(D.rowptr.p, D.colind.p, H.crs.nrows, mass);
      CUDA_CK(cudaGetLastError()); D.launches++;
    }
  }
  D.last_path = "projection";
  return MRH_OK;
}

// ---- micro-benchmarks (roofline denominators measured on the box itself) -------------------------------
__global__ void dfma_kernel(double* out, int iters) {
  double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double b = 1.0000001, c = 1e-9;
  for (int i = 0; i < iters; ++i) {
    a0 = a0 * b + c; a1 = a1 * b + c; a2 = a2 * b + c; a3 = a3 * b + c;
    a4 = a4 * b + c; a5 = a5 * b + c; a6 = a6 * b + c; a7 = a7 * b + c;
  }
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
// RED.F64 with the access pattern of the thermal scatter: consecutive threads hit consecutive 27-entry rows
__global__ void red_kernel(double* buf, int64_t nrows, int reps) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nrows) return;
  for (int r = 0; r < reps; ++r)
    for (int k = 0; k < 8; ++k) atomicAdd(buf + ((t + (k >> 1)) % nrows) * 27 + ((k * 5 + r) % 27), 1.0);
}
__global__ void copy_kernel(const double2* __restrict__ a, double2* __restrict__ b, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) b[i] = a[i];
}

}  // namespace

static int backend_create(mrh_engine* h) {
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    h->err = std::string("no CUDA device available (") + cudaGetErrorString(e) + "); this engine has no CPU fallback";
    cudaGetLastError();
    // the handle stays usable for host-only set-up/introspection; every GPU entry point fails with MRH_ECUDA
    h->impl = nullptr;
    return MRH_OK;
  }
  auto* D = new DeviceImpl();
  D->device = h->host.cfg.device;
  if (D->device < 0 || D->device >= ndev) { delete D; h->err = "CUDA device ordinal out of range"; return MRH_ECUDA; }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, D->device) != cudaSuccess) { delete D; h->err = "cudaGetDeviceProperties failed"; return MRH_ECUDA; }
  D->num_sms = prop.multiProcessorCount;
  D->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
  h->impl = D;
  return MRH_OK;
}
static void backend_destroy(mrh_engine* h) {
  auto* D = (DeviceImpl*)h->impl;
  if (!D) return;
  cudaSetDevice(D->device);
  if (D->ev0) { cudaEventDestroy(D->ev0); cudaEventDestroy(D->ev1); }
  for (cudaEvent_t e : D->chunk_ev) cudaEventDestroy(e);
  if (D->copy_stream) cudaStreamDestroy(D->copy_stream);
  delete D->tiles;
  delete D->xch;
  delete D;
}

static void backend_nodes_changed(mrh_engine* h) {
  auto* D = (DeviceImpl*)h->impl;
  if (!D || !D->ready) return;                      // not finalized yet: mrh_finalize uploads the new coordinates
  CUDA_CK(cudaSetDevice(D->device));
  CUDA_CK(cudaMemcpy(D->nodes.p, h->host.mesh.nodes.data(), h->host.mesh.nodes.size() * sizeof(double), cudaMemcpyHostToDevice));
}

#define MRH_NEED_GPU(h) if (!(h)->impl) { (h)->err = "no CUDA device available; this engine has no CPU fallback"; return MRH_ECUDA; }
#define MRH_CATCH_CUDA(h)                                                                             \
  } catch (const std::exception& e) {                                                                 \
    (h)->err = e.what();                                                                              \
    if ((h)->err.compare(0, 10, "CUDA error") == 0) return MRH_ECUDA;                                 \
    if ((h)->err.find("not able to decompose") != std::string::npos || (h)->err.find("could not evaluate") != std::string::npos) return MRH_EPARSE; \
    if ((h)->err.find("not supported") != std::string::npos) return MRH_EUNSUPPORTED;                 \
    return MRH_EINVAL; }

extern "C" {

int mrh_finalize(mrh_handle h) {
  if (!h) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h) device_finalize(h); return MRH_OK; MRH_CATCH_CUDA(h)
}
int mrh_assemble_res(mrh_handle h, const mrh_state* st, double* res) {
  if (!h || !st || !st->u || !res) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h) return device_assemble(h, st, res, nullptr); MRH_CATCH_CUDA(h)
}
int mrh_assemble_jacres(mrh_handle h, const mrh_state* st, double* res, double* jac) {
  if (!h || !st || !st->u || !res || !jac) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h) return device_assemble(h, st, res, jac); MRH_CATCH_CUDA(h)
}
int mrh_assemble_jacres_host(mrh_handle h, const mrh_state* st_host, double* res_host, double* jac_host) {
  if (!h || !st_host || !st_host->u || !res_host || !jac_host) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h)
  EngineHost& H = h->host; DeviceImpl& D = *(DeviceImpl*)h->impl;
  if (st_host->is_transient) throw std::runtime_error("host-buffer entry point is steady-state only; not supported for transient states");
  CUDA_CK(cudaSetDevice(D.device));
  const size_t nl = (size_t)H.dofs.nlocal, nnz = H.crs.colind.size();
  if (D.h_u.n != nl) { D.h_u.alloc(nl); D.h_res.alloc(nl); }
  if (D.h_jac.n != nnz) D.h_jac.alloc(nnz);
  cudaStream_t s = (cudaStream_t)st_host->stream;
  CUDA_CK(cudaMemcpyAsync(D.h_u.p, st_host->u, nl * sizeof(double), cudaMemcpyHostToDevice, s));
  CUDA_CK(cudaMemsetAsync(D.h_res.p, 0, nl * sizeof(double), s));
  CUDA_CK(cudaMemsetAsync(D.h_jac.p, 0, nnz * sizeof(double), s));
  mrh_state st = *st_host; st.u = D.h_u.p;
  build_chunk_plan(H, D);
  HostCopy hc; hc.jac_host = jac_host;
  // with a communicator attached the interface rows change after the kernels (exchange): no chunk may leave early, the whole
  // matrix is copied once the exchange is done
  const bool exchanging = D.xch && D.xch->ready && !(H.cfg.flags & 128);
  const int rc = device_assemble(h, &st, D.h_res.p, D.h_jac.p, exchanging ? nullptr : &hc);
  CUDA_CK(cudaMemcpyAsync(res_host, D.h_res.p, nl * sizeof(double), cudaMemcpyDeviceToHost, s));
  if (!hc.used) CUDA_CK(cudaMemcpyAsync(jac_host, D.h_jac.p, nnz * sizeof(double), cudaMemcpyDeviceToHost, s));
  CUDA_CK(cudaStreamSynchronize(s));
  if (hc.used) CUDA_CK(cudaStreamSynchronize(D.copy_stream));
  return rc;
  MRH_CATCH_CUDA(h)
}
// Allocation-free residual assembly with host buffers: the device copies of u and of the residual live in the
// file-scope static array g_static_scratch (engine.cu) instead of cudaMalloc'ed staging buffers.  Calls no cudaMalloc /
// cudaFree / allocating library routine: every kernel and table was built by mrh_finalize (steady-state variants are
// compiled there), so only copies, a memset and launches happen here.
int mrh_assemble_res_host_static(mrh_handle h, const mrh_state* st_host, double* res_host) {
  if (!h || !st_host || !st_host->u || !res_host) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h)
  EngineHost& H = h->host; DeviceImpl& D = *(DeviceImpl*)h->impl;
  if (!D.ready || !D.static_scratch) throw std::runtime_error("mrh_finalize has not been called");
  if (st_host->is_transient) throw std::runtime_error("allocation-free entry point is steady-state only; not supported for transient states");
  const int64_t nl = H.dofs.nlocal;
  if (nl > (int64_t)MRH_STATIC_SCRATCH_MAX_DOFS)
    throw std::runtime_error("mrh_assemble_res_host_static: " + std::to_string(nl) + " local DOFs exceed the static scratch limit of " +
                             std::to_string((int64_t)MRH_STATIC_SCRATCH_MAX_DOFS) + " (MRH_STATIC_SCRATCH_MAX_DOFS); use mrh_assemble_res with caller-owned device buffers");
  CUDA_CK(cudaSetDevice(D.device));
  cudaStream_t s = (cudaStream_t)st_host->stream;
  double* d_u = D.static_scratch;
  double* d_res = D.static_scratch + (size_t)MRH_STATIC_SCRATCH_MAX_DOFS;
  CUDA_CK(cudaMemcpyAsync(d_u, st_host->u, (size_t)nl * sizeof(double), cudaMemcpyHostToDevice, s));
  CUDA_CK(cudaMemsetAsync(d_res, 0, (size_t)nl * sizeof(double), s));      // the array keeps the last call's contents
  mrh_state st = *st_host; st.u = d_u;
  const int rc = device_assemble(h, &st, d_res, nullptr);
  CUDA_CK(cudaMemcpyAsync(res_host, d_res, (size_t)nl * sizeof(double), cudaMemcpyDeviceToHost, s));
  CUDA_CK(cudaStreamSynchronize(s));
  return rc;
  MRH_CATCH_CUDA(h)
}
int mrh_assemble_projection(mrh_handle h, int32_t kind, double time, double* rhs, double* mass, void* stream) {
  if (!h || !rhs || (kind != 0 && kind != 1)) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h) return device_projection(h, kind, time, rhs, mass, (cudaStream_t)stream); MRH_CATCH_CUDA(h)
}
// ---- multi-GPU (exchange.inl) ---------------------------------------------------------------------------------------------
int mrh_comm_unique_id(void* id128) {
  if (!id128) return MRH_EINVAL;
  NcclApi& N = NcclApi::get();
  if (!N.ok) return MRH_ECUDA;
  ncclUniqueId id;
  if (N.GetUniqueId(&id) != ncclSuccess) return MRH_ECUDA;
  static_assert(sizeof(ncclUniqueId) == 128, "NCCL unique id size");
  memcpy(id128, &id, 128);
  return MRH_OK;
}
int mrh_comm_init(mrh_handle h, const void* id128, int32_t rank, int32_t world) {
  if (!h || !id128 || world < 1 || rank < 0 || rank >= world) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h)
  EngineHost& H = h->host; DeviceImpl& D = *(DeviceImpl*)h->impl;
  NcclApi& N = NcclApi::get();
  if (!N.ok) throw std::runtime_error("CUDA error: NCCL unavailable: " + N.why);
  if (!H.part.active() || H.part.rank != rank || H.part.world != world)
    throw std::runtime_error("mrh_comm_init: set the partition first (mrh_build_brick_part or mrh_set_partition) with the same rank / world");
  CUDA_CK(cudaSetDevice(D.device));
  delete D.xch; D.xch = new ExchangeDev();
  D.xch->rank = rank; D.xch->world = world;
  ncclUniqueId id; memcpy(&id, id128, 128);
  NCCL_CK(N.CommInitRank(&D.xch->comm, world, id, rank));
  H.plan.build_send(H.part, H.crs);
  exchange_meta_over_nccl(H, *D.xch);
  if (D.ready) exchange_upload(H, *D.xch);          // otherwise mrh_finalize uploads the tables
  return MRH_OK;
  MRH_CATCH_CUDA(h)
}
int mrh_exchange(mrh_handle h, double* res, double* jac, void* stream) {
  if (!h || !res) return MRH_EINVAL;
  MRH_NEED_GPU(h)
  MRH_TRY(h)
  DeviceImpl& D = *(DeviceImpl*)h->impl;
  if (!D.xch || !D.xch->ready) throw std::runtime_error("mrh_exchange: no communicator attached (mrh_comm_init) or mrh_finalize not called");
  CUDA_CK(cudaSetDevice(D.device));
  exchange_run(D, *D.xch, res, jac, (cudaStream_t)stream);
  return MRH_OK;
  MRH_CATCH_CUDA(h)
}
int64_t mrh_get_ghost_block(mrh_handle h, int32_t* rows, int64_t* col_gids, double* values_host) {
  if (!h) return MRH_EINVAL;
  const auto& P = h->host.plan;
  const int64_t n = (int64_t)P.ghost_row.size();
  if (rows) memcpy(rows, P.ghost_row.data(), n * sizeof(int));
  if (col_gids) memcpy(col_gids, P.ghost_colgid.data(), n * sizeof(int64_t));
  if (values_host && n > 0) {
    auto* D = (DeviceImpl*)h->impl;
    if (!D || !D->xch || !D->xch->ready) return MRH_EINVAL;
    if (cudaMemcpy(values_host, D->xch->ghost.p, n * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) return MRH_ECUDA;
  }
  return n;
}
const char* mrh_last_kernel_path(mrh_handle h) {
  if (!h || !h->impl) return "none";
  auto* D = (DeviceImpl*)h->impl;
  return D->last_path;
}
int64_t mrh_last_launch_count(mrh_handle h) { return (h && h->impl) ? ((DeviceImpl*)h->impl)->launches : 0; }
double mrh_last_volume_kernel_ms(mrh_handle h) {
  if (!h || !h->impl) return -1.0;
  auto* D = (DeviceImpl*)h->impl;
  if (!D->ev1 || cudaEventSynchronize(D->ev1) != cudaSuccess) return -1.0;
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, D->ev0, D->ev1) != cudaSuccess) return -1.0;
  return (double)ms;
}

int mrh_microbench(int32_t device, double* dfma_tflops, double* red_gops, double* copy_gbs) {
  try {
    CUDA_CK(cudaSetDevice(device));
    cudaDeviceProp prop; CUDA_CK(cudaGetDeviceProperties(&prop, device));
    cudaEvent_t e0, e1; CUDA_CK(cudaEventCreate(&e0)); CUDA_CK(cudaEventCreate(&e1));
    float ms = 0.f;
    if (dfma_tflops) {
      const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 15;
      DevBuf<double> out; out.alloc((size_t)blocks * threads);
      dfma_kernel<<<blocks, threads>>>(out.p, 1024);
      double best = 0.0;
      for (int rep = 0; rep < 3; ++rep) {
        CUDA_CK(cudaEventRecord(e0)); dfma_kernel<<<blocks, threads>>>(out.p, iters); CUDA_CK(cudaEventRecord(e1));
        CUDA_CK(cudaEventSynchronize(e1)); CUDA_CK(cudaEventElapsedTime(&ms, e0, e1));
        best = std::max(best, 2.0 * 8.0 * iters * (double)blocks * threads / (ms * 1e-3) / 1e12);
      }
      *dfma_tflops = best;
    }
    if (red_gops) {
      const int64_t nrows = 16 * 1024 * 1024; const int reps = 4;
      DevBuf<double> buf; buf.alloc((size_t)nrows * 27);
      CUDA_CK(cudaMemset(buf.p, 0, (size_t)nrows * 27 * sizeof(double)));
      red_kernel<<<blocks_for(nrows, 128), 128>>>(buf.p, nrows, 1);
      double best = 0.0;
      for (int rep = 0; rep < 3; ++rep) {
        CUDA_CK(cudaEventRecord(e0)); red_kernel<<<blocks_for(nrows, 128), 128>>>(buf.p, nrows, reps); CUDA_CK(cudaEventRecord(e1));
        CUDA_CK(cudaEventSynchronize(e1)); CUDA_CK(cudaEventElapsedTime(&ms, e0, e1));
        best = std::max(best, 8.0 * reps * (double)nrows / (ms * 1e-3) / 1e9);
      }
      *red_gops = best;
    }
    if (copy_gbs) {
      const int64_t n = 256LL * 1024 * 1024;   // double2 elements: 4 GiB read + 4 GiB write
      DevBuf<double2> a, b; a.alloc(n); b.alloc(n);
      CUDA_CK(cudaMemset(a.p, 0, n * sizeof(double2)));
      copy_kernel<<<prop.multiProcessorCount * 16, 512>>>(a.p, b.p, n);
      double best = 0.0;
      for (int rep = 0; rep < 3; ++rep) {
        CUDA_CK(cudaEventRecord(e0)); copy_kernel<<<prop.multiProcessorCount * 16, 512>>>(a.p, b.p, n); CUDA_CK(cudaEventRecord(e1));
        CUDA_CK(cudaEventSynchronize(e1)); CUDA_CK(cudaEventElapsedTime(&ms, e0, e1));
        best = std::max(best, 2.0 * n * sizeof(double2) / (ms * 1e-3) / 1e9);
      }
      *copy_gbs = best;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return MRH_OK;
  } catch (const std::exception&) { return MRH_ECUDA; }
}

}  // extern "C"
