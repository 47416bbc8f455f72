// Part of engine.cu: construction of the tile plan used by the tile-staged scatter (params.h TilePlan).
// Elements are taken in tiles of 128 (host-chosen order, brick-blocked on structured meshes).  The 128 x 72
// contributions of a tile (64 element-matrix entries + 8 residual entries per element) are sorted by destination
// (CRS slot, then residual row; contributions of fixed rows last) with a block-wide radix sort.  The sorted order
// is the tile's contribution list: thread t of the assembly kernel sums positions [72 t, 72 t + 72).  The distinct
// destinations give the tile's slot / row lists for the flush.  Contribution lists are de-duplicated into a
// dictionary (structured meshes collapse to a few dozen lists).
// The same plan serves the Q1 quadrilateral kernel (assemble_fast2d.h): NN = 4 nodes, 16 + 4 = 20 contributions per element,
// padded to 24 per thread in the stored lists (pads read the zero cell and never start a run).
#include <cub/block/block_radix_sort.cuh>
#include <cub/block/block_scan.cuh>

namespace {

constexpr int TILE_ELEMS = 128;
template <int NN> constexpr int tile_nc() { return NN * NN + NN; }                 // contributions per element: 72 (Q1 hex), 20 (Q1 quad)
template <int NN> constexpr int tile_ncp() { return (tile_nc<NN>() + 7) & ~7; }       // list stride per thread: whole uint4 words
constexpr int SRC_BITS = 14;                               // entry * 128 + thread < 72 * 128 + 1
constexpr unsigned long long KEY_SENTINEL = ~0ull;
constexpr unsigned long long ROW_SPACE = 1ull << 45;       // destination keys: CRS slot | ROW_SPACE + row | masked
constexpr unsigned long long MASKED_DEST = KEY_SENTINEL >> SRC_BITS;

struct TileDev {
  DevBuf<int> perm, rows, cl_id;
  DevBuf<int64_t> eptr, rptr, base, ebeg;      // ebeg: start of each tile's slot list in `off` (lists are de-duplicated)
  DevBuf<uint32_t> off;
  DevBuf<uint16_t> clist, cbase;
  int64_t ntiles = 0, npatterns = 0, noffpatterns = 0;
  int max_acc = 0, nk = 0, max_e = 0, max_r = 0, nn = 8;
  bool ok = false;
  void release() { perm.release(); rows.release(); cl_id.release(); eptr.release(); rptr.release(); base.release(); ebeg.release(); off.release();
                   clist.release(); cbase.release(); ok = false; }
};

template <int NN>
struct TileSmem {
  typedef cub::BlockRadixSort<unsigned long long, TILE_ELEMS, tile_nc<NN>()> Sort;
  typedef cub::BlockScan<long long, TILE_ELEMS> Scan;
  union Temp { typename Sort::TempStorage sort; typename Scan::TempStorage scan; };
  Temp temp;
  unsigned long long last[TILE_ELEMS];
  unsigned long long hash[TILE_ELEMS];
  unsigned long long first_dest;
};

// destination keys of the NN * NN + NN contributions of element e held by thread `tid` (one variable, NN = 8: Q1 hex, 4: Q1 quad)
template <int NN>
__device__ __forceinline__ void element_keys(const int* lids, const uint8_t* fixed, const int64_t* rowptr, const int* colind,
                                             int64_t e, int tid, int nk, unsigned long long (&key)[NN * NN + NN], int* bad) {
  const int* l = lids + e * NN;
  for (int i = 0; i < NN; ++i) {
    const int li = l[i];
    const bool fx = fixed[li] != 0;
    for (int j = 0; j < NN; ++j) {
      const int s = fx ? -1 : rel_slot(rowptr, colind, li, l[j]);
      if (!fx && s < 0) atomicExch(bad, 1);
      const int a = i <= j ? i : j, b = i <= j ? j : i;
      const int entry = (nk == NN * NN) ? i * NN + j : a * NN - (a * (a - 1)) / 2 + (b - a);      // full or upper-triangular element store
      key[i * NN + j] = (s < 0) ? KEY_SENTINEL : (((unsigned long long)(rowptr[li] + s) << SRC_BITS) | (unsigned)(entry * TILE_ELEMS + tid));
    }
    key[NN * NN + i] = fx ? KEY_SENTINEL : (((ROW_SPACE + (unsigned long long)li) << SRC_BITS) | (unsigned)((nk + i) * TILE_ELEMS + tid));
  }
}

template <bool WRITE, int NN>
__global__ void __launch_bounds__(TILE_ELEMS) tile_plan_kernel(const int* perm, int64_t nelem, const int* lids, const uint8_t* fixed,
                                                               const int64_t* rowptr, const int* colind, int nk, int* E_count, int* R_count,
                                                               const int64_t* eptr, const int64_t* rptr, int64_t* tile_base,
                                                               uint32_t* tile_off, int* tile_rows, uint16_t* cl_all, uint16_t* cbase_all,
                                                               uint64_t* tile_hash, int* bad) {
  extern __shared__ __align__(16) unsigned char smraw[];
  typedef TileSmem<NN> TS;
  constexpr int TILE_NC = tile_nc<NN>(), TILE_NCP = tile_ncp<NN>();
  TS& S = *reinterpret_cast<TS*>(smraw);
  const int tid = threadIdx.x;
  const int64_t t = blockIdx.x, e0 = t * TILE_ELEMS + tid;
  unsigned long long key[TILE_NC];
  if (e0 < nelem) element_keys<NN>(lids, fixed, rowptr, colind, (int64_t)perm[e0], tid, nk, key, bad);
  else for (int k = 0; k < TILE_NC; ++k) key[k] = KEY_SENTINEL;

  typename TS::Sort(S.temp.sort).Sort(key);                // blocked arrangement: thread t holds positions [NC t, NC t + NC)
  S.last[tid] = key[TILE_NC - 1] >> SRC_BITS;
  if (tid == 0) S.first_dest = key[0] >> SRC_BITS;
  __syncthreads();
  // a contribution starts a run when its destination differs from its predecessor's
  unsigned long long prev = (tid == 0) ? 0ull : S.last[tid - 1];
  long long cnt = 0;                                        // packed counts: slots | rows << 20 | masked << 40
  bool starts0 = false;
  for (int k = 0; k < TILE_NC; ++k) {
    const unsigned long long d = key[k] >> SRC_BITS;
    const bool st = (tid == 0 && k == 0) || d != prev;
    if (k == 0) starts0 = st;
    if (st) cnt += (d < ROW_SPACE) ? 1ll : (d != MASKED_DEST ? (1ll << 20) : (1ll << 40));
    prev = d;
  }
  long long before = 0, total = 0;
  typename TS::Scan(S.temp.scan).ExclusiveSum(cnt, before, total);
  const int E = (int)(total & 0xfffff), R = (int)((total >> 20) & 0xfffff);
  if (!WRITE) { if (tid == 0) { E_count[t] = E; R_count[t] = R; } return; }

  int run = (int)((before & 0xfffff) + ((before >> 20) & 0xfffff) + (before >> 40));     // runs that start before my chunk
  const unsigned long long base = (E > 0) ? S.first_dest : 0ull;
  if (tid == 0) tile_base[t] = (int64_t)base;
  if (E + R > 32000) atomicExch(bad, 4);
  const int zero_src = (nk + NN) * TILE_ELEMS;
  uint16_t* out = cl_all + ((size_t)t * TILE_ELEMS + tid) * TILE_NCP;
  uint64_t h = 0x51ed27ull;
  const int loc0 = starts0 ? run : run - 1;
  cbase_all[(size_t)t * TILE_ELEMS + tid] = (uint16_t)loc0;
  h = mix64(h, (uint64_t)loc0);
  prev = (tid == 0) ? 0ull : S.last[tid - 1];
  for (int k = 0; k < TILE_NC; ++k) {
    const unsigned long long d = key[k] >> SRC_BITS;
    const bool st = (tid == 0 && k == 0) || d != prev;
    if (st) {
      if (d < ROW_SPACE) {
        const unsigned long long rel = d - base;
        if (rel > 0xffffffffull) atomicExch(bad, 3);
        tile_off[eptr[t] + run] = (uint32_t)rel;
      } else if (d != MASKED_DEST) tile_rows[rptr[t] + (run - E)] = (int)(d - ROW_SPACE);
      ++run;
    }
    const unsigned src = (d == MASKED_DEST) ? (unsigned)zero_src : (unsigned)(key[k] & ((1u << SRC_BITS) - 1));
    const unsigned c = src | (st ? 0x8000u : 0u);
    out[k] = (uint16_t)c; h = mix64(h, c);
    prev = d;
  }
  for (int k = TILE_NC; k < TILE_NCP; ++k) out[k] = (uint16_t)zero_src;      // pads: add 0.0 to the thread's last run
  S.hash[tid] = h;
  __syncthreads();
  if (tid == 0) { uint64_t g = 0x9d2cull; for (int k = 0; k < TILE_ELEMS; ++k) g = mix64(g, S.hash[k]); tile_hash[t] = g; }
}

__global__ void clist_fill_kernel(const uint16_t* cl_all, const uint16_t* cbase_all, const int* rep, int64_t npat, uint16_t* clist, uint16_t* cbase, int ncp) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t L = (int64_t)TILE_ELEMS * ncp;
  if (k < npat * L) clist[k] = cl_all[(int64_t)rep[k / L] * L + (k % L)];
  if (k < npat * TILE_ELEMS) cbase[k] = cbase_all[(int64_t)rep[k / TILE_ELEMS] * TILE_ELEMS + (k % TILE_ELEMS)];
}
// list id of every tile (binary search on the sorted unique hashes) and full verification against the dictionary
__global__ void clist_assign_kernel(const uint16_t* cl_all, const uint16_t* cbase_all, int64_t ntiles, const uint64_t* hash, const uint64_t* uhash,
                                    int64_t npat, const uint16_t* clist, const uint16_t* cbase, int* cl_id, int* bad, int ncp) {
  const int64_t t = blockIdx.x;
  const uint64_t h = hash[t];
  int64_t lo = 0, hi = npat;
  while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (uhash[mid] < h) lo = mid + 1; else hi = mid; }
  if (lo >= npat) { atomicExch(bad, 5); return; }
  if (threadIdx.x == 0) cl_id[t] = (int)lo;
  const int L = TILE_ELEMS * ncp;
  for (int k = threadIdx.x; k < L; k += blockDim.x) if (clist[lo * L + k] != cl_all[t * L + k]) atomicExch(bad, 5);
  for (int k = threadIdx.x; k < TILE_ELEMS; k += blockDim.x) if (cbase[lo * TILE_ELEMS + k] != cbase_all[t * TILE_ELEMS + k]) atomicExch(bad, 5);
}

// ---- dictionary of the slot lists: the lists hold CRS slots RELATIVE to the tile's first slot, which makes them identical for
// all tiles of the same shape on a mesh whose numbering is translation-invariant (the brick's interior: 131 072 lists of 17 KB
// collapse to a few dozen) - 136 B per element of DRAM traffic become L2 hits and 2.3 GB of device memory at 256^3 a megabyte
__global__ void off_hash_kernel(const uint32_t* off_all, const int64_t* eptr, uint64_t* hash) {
  __shared__ uint64_t part[128];
  const int64_t t = blockIdx.x, b = eptr[t]; const int E = (int)(eptr[t + 1] - b);
  uint64_t h = 0x2545f491ull + (uint64_t)threadIdx.x;
  for (int k = threadIdx.x; k < E; k += 128) h = mix64(h, ((uint64_t)k << 32) | off_all[b + k]);
  part[threadIdx.x] = h;
  __syncthreads();
  if (threadIdx.x == 0) { uint64_t g = mix64(0x9e37ull, (uint64_t)E); for (int k = 0; k < 128; ++k) g = mix64(g, part[k]); hash[t] = g; }
}
// pattern of every tile (binary search on the sorted unique hashes), its list start in the dictionary, and a full comparison
__global__ void off_assign_kernel(const uint32_t* off_all, const int64_t* eptr, const uint64_t* hash, const uint64_t* uhash, int64_t npat,
                                  const int* rep, const int64_t* dptr, int64_t* ebeg, int* bad) {
  const int64_t t = blockIdx.x;
  const uint64_t h = hash[t];
  int64_t lo = 0, hi = npat;
  while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (uhash[mid] < h) lo = mid + 1; else hi = mid; }
  if (lo >= npat) { atomicExch(bad, 6); return; }
  const int64_t r = rep[lo], b = eptr[t], br = eptr[r];
  const int E = (int)(eptr[t + 1] - b);
  if ((int)(eptr[r + 1] - br) != E) { atomicExch(bad, 6); return; }
  for (int k = threadIdx.x; k < E; k += blockDim.x) if (off_all[b + k] != off_all[br + k]) atomicExch(bad, 6);
  if (threadIdx.x == 0) ebeg[t] = dptr[lo];
}
__global__ void off_fill_kernel(const uint32_t* off_all, const int64_t* eptr, const int* rep, const int64_t* dptr, uint32_t* dict) {
  const int64_t p = blockIdx.x, r = rep[p], b = eptr[r];
  const int E = (int)(eptr[r + 1] - b);
  for (int k = threadIdx.x; k < E; k += blockDim.x) dict[dptr[p] + k] = off_all[b + k];
}

// throws on CUDA errors; leaves T.ok = false (with a reason) when the plan cannot be used
template <int NN>
std::string build_tile_plan(const std::vector<int>& perm_host, int64_t nelem, const int* d_lids, const uint8_t* d_fixed,
                            const int64_t* d_rowptr, const int* d_colind, int nk, TileDev& T) {
  T.release();
  T.nn = NN;
  const int64_t nt = (nelem + TILE_ELEMS - 1) / TILE_ELEMS;
  T.ntiles = nt; T.nk = nk; T.max_e = T.max_r = 0;
  T.perm.upload(perm_host);
  DevBuf<int> Ec, Rc, bad, idx; Ec.alloc(nt); Rc.alloc(nt); bad.alloc(1);
  CUDA_CK(cudaMemset(bad.p, 0, sizeof(int)));
  const size_t smem = sizeof(TileSmem<NN>);
  CUDA_CK(cudaFuncSetAttribute(tile_plan_kernel<false, NN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CUDA_CK(cudaFuncSetAttribute(tile_plan_kernel<true, NN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tile_plan_kernel<false, NN><<<(unsigned)nt, TILE_ELEMS, smem>>>(T.perm.p, nelem, d_lids, d_fixed, d_rowptr, d_colind, nk, Ec.p, Rc.p,
                                                             nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, bad.p);
  CUDA_CK(cudaGetLastError());
  std::vector<int> hE(nt), hR(nt);
  CUDA_CK(cudaMemcpy(hE.data(), Ec.p, nt * sizeof(int), cudaMemcpyDeviceToHost));
  CUDA_CK(cudaMemcpy(hR.data(), Rc.p, nt * sizeof(int), cudaMemcpyDeviceToHost));
  int hbad = 0; CUDA_CK(cudaMemcpy(&hbad, bad.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (hbad) { T.release(); return "an element entry has no CRS slot"; }
  std::vector<int64_t> eptr(nt + 1, 0), rptr(nt + 1, 0);
  int max_acc = 0;
  for (int64_t t = 0; t < nt; ++t) { eptr[t + 1] = eptr[t] + hE[t]; rptr[t + 1] = rptr[t] + hR[t]; max_acc = std::max(max_acc, hE[t] + hR[t]); T.max_e = std::max(T.max_e, hE[t]); T.max_r = std::max(T.max_r, hR[t]); }
  T.max_acc = max_acc;
  T.eptr.upload(eptr); T.rptr.upload(rptr); T.base.alloc(nt);
  DevBuf<uint32_t> off_all; off_all.alloc((size_t)eptr[nt] + 4);
  T.rows.alloc((size_t)rptr[nt] + 4);      // + 4: the bulk-copy staging reads whole 16-byte groups
  constexpr size_t L = (size_t)TILE_ELEMS * tile_ncp<NN>();
  DevBuf<uint16_t> cl_all, cbase_all; cl_all.alloc((size_t)nt * L); cbase_all.alloc((size_t)nt * TILE_ELEMS);
  DevBuf<uint64_t> hash, skey; hash.alloc(nt); skey.alloc(nt); idx.alloc(nt);
  tile_plan_kernel<true, NN><<<(unsigned)nt, TILE_ELEMS, smem>>>(T.perm.p, nelem, d_lids, d_fixed, d_rowptr, d_colind, nk, nullptr, nullptr,
                                                            T.eptr.p, T.rptr.p, T.base.p, off_all.p, T.rows.p, cl_all.p, cbase_all.p, hash.p, bad.p);
  CUDA_CK(cudaGetLastError());
  CUDA_CK(cudaMemcpy(&hbad, bad.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (hbad) { T.release(); return hbad == 3 ? "a tile spans more than 2^32 CRS slots" : "a tile has more than 32000 destinations"; }
  CUDA_CK(cudaMemcpy(skey.p, hash.p, nt * sizeof(uint64_t), cudaMemcpyDeviceToDevice));
  thrust::device_ptr<uint64_t> kp(skey.p); thrust::device_ptr<int> ip(idx.p);
  thrust::sequence(thrust::device, ip, ip + nt);
  thrust::sort_by_key(thrust::device, kp, kp + nt, ip);
  auto ends = thrust::unique_by_key(thrust::device, kp, kp + nt, ip);
  const int64_t npat = ends.first - kp;
  T.npatterns = npat;
  T.clist.alloc((size_t)npat * L); T.cbase.alloc((size_t)npat * TILE_ELEMS); T.cl_id.alloc(nt);
  clist_fill_kernel<<<blocks_for(npat * (int64_t)L, 256), 256>>>(cl_all.p, cbase_all.p, idx.p, npat, T.clist.p, T.cbase.p, tile_ncp<NN>());
  CUDA_CK(cudaGetLastError());
  clist_assign_kernel<<<(unsigned)nt, 256>>>(cl_all.p, cbase_all.p, nt, hash.p, skey.p, npat, T.clist.p, T.cbase.p, T.cl_id.p, bad.p, tile_ncp<NN>());
  CUDA_CK(cudaGetLastError());
  CUDA_CK(cudaMemcpy(&hbad, bad.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (hbad) { T.release(); return "contribution list dictionary verification failed (hash collision)"; }
  cl_all.release(); cbase_all.release();
  // slot-list dictionary (falls back to one list per tile when the lists do not repeat or a hash collides)
  {
    DevBuf<uint64_t> ohash, okey; ohash.alloc(nt); okey.alloc(nt);
    off_hash_kernel<<<(unsigned)nt, 128>>>(off_all.p, T.eptr.p, ohash.p);
    CUDA_CK(cudaGetLastError());
    CUDA_CK(cudaMemcpy(okey.p, ohash.p, nt * sizeof(uint64_t), cudaMemcpyDeviceToDevice));
    thrust::device_ptr<uint64_t> okp(okey.p);
    thrust::sequence(thrust::device, ip, ip + nt);
    thrust::sort_by_key(thrust::device, okp, okp + nt, ip);
    auto oends = thrust::unique_by_key(thrust::device, okp, okp + nt, ip);
    const int64_t nop = oends.first - okp;
    bool dedupe = nop <= nt / 2;
    if (dedupe) {
      std::vector<int> rep((size_t)nop);
      CUDA_CK(cudaMemcpy(rep.data(), idx.p, (size_t)nop * sizeof(int), cudaMemcpyDeviceToHost));
      std::vector<int64_t> dptr((size_t)nop + 1, 0);
      for (int64_t p = 0; p < nop; ++p) dptr[p + 1] = dptr[p] + ((hE[rep[p]] + 3) & ~3);       // every list starts on a 16-byte boundary
      DevBuf<int64_t> ddptr; ddptr.upload(dptr);
      T.ebeg.alloc(nt); T.off.alloc((size_t)dptr[nop] + 4);
      CUDA_CK(cudaMemset(T.off.p, 0, ((size_t)dptr[nop] + 4) * sizeof(uint32_t)));
      CUDA_CK(cudaMemset(bad.p, 0, sizeof(int)));
      off_assign_kernel<<<(unsigned)nt, 128>>>(off_all.p, T.eptr.p, ohash.p, okey.p, nop, idx.p, ddptr.p, T.ebeg.p, bad.p);
      CUDA_CK(cudaGetLastError());
      off_fill_kernel<<<(unsigned)nop, 128>>>(off_all.p, T.eptr.p, idx.p, ddptr.p, T.off.p);
      CUDA_CK(cudaGetLastError());
      CUDA_CK(cudaMemcpy(&hbad, bad.p, sizeof(int), cudaMemcpyDeviceToHost));
      if (hbad) { dedupe = false; T.ebeg.release(); T.off.release(); }
      else T.noffpatterns = nop;
    }
    if (!dedupe) {               // one list per tile: the lists as the plan kernel wrote them
      T.off.alloc((size_t)eptr[nt] + 4);
      CUDA_CK(cudaMemcpy(T.off.p, off_all.p, ((size_t)eptr[nt] + 4) * sizeof(uint32_t), cudaMemcpyDeviceToDevice));
      std::vector<int64_t> eb(eptr.begin(), eptr.begin() + nt);
      T.ebeg.upload(eb);
      T.noffpatterns = nt;
    }
  }
  CUDA_CK(cudaDeviceSynchronize());
  T.ok = true;
  return "";
}

void fill_tile_params(const TileDev& T, TilePlan& tp) {
  tp.perm = T.perm.p; tp.eptr = T.eptr.p; tp.ebeg = T.ebeg.p; tp.rptr = T.rptr.p; tp.base = T.base.p; tp.off = T.off.p; tp.rows = T.rows.p;
  tp.clist = T.clist.p; tp.cbase = T.cbase.p; tp.cl_id = T.cl_id.p; tp.ntiles = T.ntiles; tp.tile_begin = 0; tp.tile_end = T.ntiles; tp.max_acc = T.max_acc; tp.nk = T.nk; tp.max_e = T.max_e; tp.max_r = T.max_r;
}

}  // namespace
