// Part of engine.cu: construction of the tile plan used by the tile-staged scatter (params.h TilePlan).
// Elements are taken in tiles of 128 (host-chosen order, brick-blocked on structured meshes); for each tile the
// distinct CRS slots and residual rows its elements touch are found with a block-wide radix sort, and every element
// gets the position of its 64 + 8 contributions in those lists.  The per-element position vectors are
// de-duplicated into a dictionary like the slot patterns (structured meshes collapse to a few hundred vectors).
#include <cub/block/block_radix_sort.cuh>
#include <cub/block/block_scan.cuh>

namespace {

constexpr int TILE_ELEMS = 128;
constexpr unsigned long long SLOT_SENTINEL = ~0ull;
constexpr unsigned ROW_SENTINEL = ~0u;

struct TileDev {
  DevBuf<int> perm, rows, tpat_id;
  DevBuf<int64_t> eptr, rptr, base;
  DevBuf<uint32_t> off;
  DevBuf<uint16_t> tpat;
  int64_t ntiles = 0, npatterns = 0;
  int max_acc = 0;
  bool ok = false;
  void release() { perm.release(); rows.release(); tpat_id.release(); eptr.release(); rptr.release(); base.release(); off.release(); tpat.release(); ok = false; }
};

struct TileSmem {
  typedef cub::BlockRadixSort<unsigned long long, TILE_ELEMS, 64> Sort64;
  typedef cub::BlockRadixSort<unsigned, TILE_ELEMS, 8> Sort8;
  typedef cub::BlockScan<int, TILE_ELEMS> Scan;
  union Temp { typename Sort64::TempStorage s64; typename Sort8::TempStorage s8; typename Scan::TempStorage scan; };
  Temp temp;
  unsigned long long uniq[TILE_ELEMS * 64];
  unsigned urow[TILE_ELEMS * 8];
  unsigned long long last64[TILE_ELEMS];
  unsigned last8[TILE_ELEMS];
  int total;
};

// N = 8 (Q1 hex, one variable): slots of element e, fixed rows masked out
__device__ __forceinline__ void element_slots(const int* lids, const uint8_t* fixed, const int64_t* rowptr, const int* colind,
                                              int64_t e, unsigned long long (&key)[64], unsigned (&row)[8]) {
  const int* l = lids + e * 8;
  for (int i = 0; i < 8; ++i) {
    const int li = l[i];
    const bool fx = fixed[li] != 0;
    row[i] = fx ? ROW_SENTINEL : (unsigned)li;
    for (int j = 0; j < 8; ++j) {
      const int s = fx ? -1 : rel_slot(rowptr, colind, li, l[j]);
      key[i * 8 + j] = (s < 0) ? SLOT_SENTINEL : (unsigned long long)(rowptr[li] + s);
    }
  }
}

template <bool WRITE>
__global__ void __launch_bounds__(TILE_ELEMS) tile_plan_kernel(const int* perm, int64_t nelem, const int* lids, const uint8_t* fixed,
                                                               const int64_t* rowptr, const int* colind, int* E_count, int* R_count,
                                                               const int64_t* eptr, const int64_t* rptr, int64_t* tile_base,
                                                               uint32_t* tile_off, int* tile_rows, uint16_t* loc_all, uint64_t* hash, int* bad) {
  extern __shared__ __align__(16) unsigned char smraw[];
  TileSmem& S = *reinterpret_cast<TileSmem*>(smraw);
  const int tid = threadIdx.x;
  const int64_t t = blockIdx.x, e0 = t * TILE_ELEMS + tid;
  const int64_t e = (e0 < nelem) ? (int64_t)perm[e0] : -1;
  unsigned long long key[64]; unsigned row[8];
  if (e >= 0) element_slots(lids, fixed, rowptr, colind, e, key, row);
  else { for (int k = 0; k < 64; ++k) key[k] = SLOT_SENTINEL; for (int k = 0; k < 8; ++k) row[k] = ROW_SENTINEL; }

  // ---- distinct CRS slots of the tile ------------------------------------------------------------------
  TileSmem::Sort64(S.temp.s64).Sort(key);
  S.last64[tid] = key[63];
  __syncthreads();
  unsigned long long prev = (tid == 0) ? SLOT_SENTINEL : S.last64[tid - 1];
  int cnt = 0;
  for (int k = 0; k < 64; ++k) { if (key[k] != SLOT_SENTINEL && (key[k] != prev || (tid == 0 && k == 0))) ++cnt; prev = key[k]; }
  int offs = 0, total = 0;
  TileSmem::Scan(S.temp.scan).ExclusiveSum(cnt, offs, total);
  __syncthreads();
  const int E = total;
  if (WRITE) {
    prev = (tid == 0) ? SLOT_SENTINEL : S.last64[tid - 1];
    for (int k = 0; k < 64; ++k) { if (key[k] != SLOT_SENTINEL && (key[k] != prev || (tid == 0 && k == 0))) S.uniq[offs++] = key[k]; prev = key[k]; }
  } else if (tid == 0) E_count[t] = E;
  __syncthreads();

  // ---- distinct residual rows -----------------------------------------------------------------------------
  TileSmem::Sort8(S.temp.s8).Sort(row);
  S.last8[tid] = row[7];
  __syncthreads();
  unsigned prow = (tid == 0) ? ROW_SENTINEL : S.last8[tid - 1];
  cnt = 0;
  for (int k = 0; k < 8; ++k) { if (row[k] != ROW_SENTINEL && (row[k] != prow || (tid == 0 && k == 0))) ++cnt; prow = row[k]; }
  TileSmem::Scan(S.temp.scan).ExclusiveSum(cnt, offs, total);
  __syncthreads();
  const int R = total;
  if (!WRITE) { if (tid == 0) R_count[t] = R; return; }
  prow = (tid == 0) ? ROW_SENTINEL : S.last8[tid - 1];
  for (int k = 0; k < 8; ++k) { if (row[k] != ROW_SENTINEL && (row[k] != prow || (tid == 0 && k == 0))) S.urow[offs++] = row[k]; prow = row[k]; }
  __syncthreads();

  // ---- write the tile lists ---------------------------------------------------------------------------------
  const unsigned long long base = (E > 0) ? S.uniq[0] : 0ull;
  if (tid == 0) tile_base[t] = (int64_t)base;
  for (int k = tid; k < E; k += TILE_ELEMS) {
    const unsigned long long d = S.uniq[k] - base;
    if (d > 0xffffffffull) atomicExch(bad, 3);
    tile_off[eptr[t] + k] = (uint32_t)d;
  }
  for (int k = tid; k < R; k += TILE_ELEMS) tile_rows[rptr[t] + k] = (int)S.urow[k];
  if (E > 65534 || R > 65534) atomicExch(bad, 4);
  if (e < 0) return;

  // ---- position of this element's contributions in the lists --------------------------------------------------
  element_slots(lids, fixed, rowptr, colind, e, key, row);
  uint64_t h = 0x51ed27ull;
  uint16_t* out = loc_all + e * 72;
  for (int k = 0; k < 64; ++k) {
    unsigned loc = 0xffffu;
    if (key[k] != SLOT_SENTINEL) {
      int lo = 0, hi = E;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (S.uniq[mid] < key[k]) lo = mid + 1; else hi = mid; }
      loc = (unsigned)lo;
    }
    out[k] = (uint16_t)loc; h = mix64(h, loc);
  }
  for (int k = 0; k < 8; ++k) {
    unsigned loc = 0xffffu;
    if (row[k] != ROW_SENTINEL) {
      int lo = 0, hi = R;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (S.urow[mid] < row[k]) lo = mid + 1; else hi = mid; }
      loc = (unsigned)lo;
    }
    out[64 + k] = (uint16_t)loc; h = mix64(h, loc);
  }
  hash[e] = h;
}

__global__ void tpat_fill_kernel(const uint16_t* loc_all, const int* rep, int64_t npat, uint16_t* tpat) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= npat * 72) return;
  tpat[k] = loc_all[(int64_t)rep[k / 72] * 72 + (k % 72)];
}
__global__ void tpat_assign_kernel(const uint16_t* loc_all, int64_t nelem, const uint64_t* hash, const uint64_t* uhash, int64_t npat,
                                   const uint16_t* tpat, int* tpat_id, int* bad) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nelem) return;
  const uint64_t h = hash[e];
  int64_t lo = 0, hi = npat;
  while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (uhash[mid] < h) lo = mid + 1; else hi = mid; }
  tpat_id[e] = (int)lo;
  if (lo >= npat) { atomicExch(bad, 5); return; }
  for (int k = 0; k < 72; ++k) if (tpat[lo * 72 + k] != loc_all[e * 72 + k]) atomicExch(bad, 5);
}

// throws on CUDA errors; leaves T.ok = false (with a reason) when the plan cannot be used
std::string build_tile_plan(const std::vector<int>& perm_host, int64_t nelem, const int* d_lids, const uint8_t* d_fixed,
                            const int64_t* d_rowptr, const int* d_colind, TileDev& T) {
  T.release();
  const int64_t nt = (nelem + TILE_ELEMS - 1) / TILE_ELEMS;
  T.ntiles = nt;
  T.perm.upload(perm_host);
  DevBuf<int> Ec, Rc, bad, idx; Ec.alloc(nt); Rc.alloc(nt); bad.alloc(1);
  CUDA_CK(cudaMemset(bad.p, 0, sizeof(int)));
  const size_t smem = sizeof(TileSmem);
  CUDA_CK(cudaFuncSetAttribute(tile_plan_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CUDA_CK(cudaFuncSetAttribute(tile_plan_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tile_plan_kernel<false><<<(unsigned)nt, TILE_ELEMS, smem>>>(T.perm.p, nelem, d_lids, d_fixed, d_rowptr, d_colind, Ec.p, Rc.p,
                                                             nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, bad.p);
  CUDA_CK(cudaGetLastError());
  std::vector<int> hE(nt), hR(nt);
  CUDA_CK(cudaMemcpy(hE.data(), Ec.p, nt * sizeof(int), cudaMemcpyDeviceToHost));
  CUDA_CK(cudaMemcpy(hR.data(), Rc.p, nt * sizeof(int), cudaMemcpyDeviceToHost));
  std::vector<int64_t> eptr(nt + 1, 0), rptr(nt + 1, 0);
  int max_acc = 0;
  for (int64_t t = 0; t < nt; ++t) { eptr[t + 1] = eptr[t] + hE[t]; rptr[t + 1] = rptr[t] + hR[t]; max_acc = std::max(max_acc, hE[t] + hR[t]); }
  T.max_acc = max_acc;
  T.eptr.upload(eptr); T.rptr.upload(rptr); T.base.alloc(nt);
  T.off.alloc((size_t)std::max<int64_t>(eptr[nt], 1)); T.rows.alloc((size_t)std::max<int64_t>(rptr[nt], 1));
  DevBuf<uint16_t> loc_all; loc_all.alloc((size_t)nelem * 72);
  DevBuf<uint64_t> hash, skey; hash.alloc(nelem); skey.alloc(nelem); idx.alloc(nelem);
  tile_plan_kernel<true><<<(unsigned)nt, TILE_ELEMS, smem>>>(T.perm.p, nelem, d_lids, d_fixed, d_rowptr, d_colind, nullptr, nullptr,
                                                            T.eptr.p, T.rptr.p, T.base.p, T.off.p, T.rows.p, loc_all.p, hash.p, bad.p);
  CUDA_CK(cudaGetLastError());
  int hbad = 0; CUDA_CK(cudaMemcpy(&hbad, bad.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (hbad) { T.release(); return hbad == 3 ? "a tile spans more than 2^32 CRS slots" : "a tile has more than 65534 entries"; }
  CUDA_CK(cudaMemcpy(skey.p, hash.p, nelem * sizeof(uint64_t), cudaMemcpyDeviceToDevice));
  thrust::device_ptr<uint64_t> kp(skey.p); thrust::device_ptr<int> ip(idx.p);
  thrust::sequence(thrust::device, ip, ip + nelem);
  thrust::sort_by_key(thrust::device, kp, kp + nelem, ip);
  auto ends = thrust::unique_by_key(thrust::device, kp, kp + nelem, ip);
  const int64_t npat = ends.first - kp;
  T.npatterns = npat;
  T.tpat.alloc((size_t)npat * 72); T.tpat_id.alloc(nelem);
  tpat_fill_kernel<<<blocks_for(npat * 72, 256), 256>>>(loc_all.p, idx.p, npat, T.tpat.p);
  CUDA_CK(cudaGetLastError());
  tpat_assign_kernel<<<blocks_for(nelem, 128), 128>>>(loc_all.p, nelem, hash.p, skey.p, npat, T.tpat.p, T.tpat_id.p, bad.p);
  CUDA_CK(cudaGetLastError());
  CUDA_CK(cudaMemcpy(&hbad, bad.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (hbad) { T.release(); return "tile pattern dictionary verification failed (hash collision)"; }
  T.ok = true;
  return "";
}

void fill_tile_params(const TileDev& T, TilePlan& tp) {
  tp.perm = T.perm.p; tp.eptr = T.eptr.p; tp.rptr = T.rptr.p; tp.base = T.base.p; tp.off = T.off.p; tp.rows = T.rows.p;
  tp.tpat = T.tpat.p; tp.tpat_id = T.tpat_id.p; tp.ntiles = T.ntiles; tp.max_acc = T.max_acc; tp.pad_ = 0;
}

}  // namespace
