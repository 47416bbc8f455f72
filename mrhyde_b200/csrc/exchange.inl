// Part of engine.cu: device side of the multi-GPU path.  One process per GPU; NCCL carries one message per neighbour and step
// (ncclSend / ncclRecv inside one group), pack / unpack are kernels of this library, and for the tile-staged thermal kernels the
// interface tiles are assembled first so that the exchange overlaps the interior.  Replaces
// LinearAlgebraInterface::exportVectorFromOverlapped / exportMatrixFromOverlapped
// (/root/reference/src/interfaces/linearAlgebraInterface.hpp:296-337, Tpetra doExport with ADD).
// NCCL is resolved with dlopen at mrh_comm_init: a process that already loaded a libnccl.so.2 (PyTorch's) shares it.
#include <dlfcn.h>
#include <nccl.h>

namespace {

struct NcclApi {
  bool ok = false; std::string why;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  static NcclApi& get() {
    static NcclApi a = load();
    return a;
  }
  static NcclApi load() {
    NcclApi a;
    void* hn = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!hn) hn = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!hn) { a.why = "libnccl.so.2 not found"; return a; }
#define MRH_NSYM(f, n) a.f = (decltype(a.f))dlsym(hn, n); if (!a.f) { a.why = std::string("missing NCCL symbol ") + n; return a; }
    MRH_NSYM(GetUniqueId, "ncclGetUniqueId") MRH_NSYM(CommInitRank, "ncclCommInitRank") MRH_NSYM(CommDestroy, "ncclCommDestroy")
    MRH_NSYM(Send, "ncclSend") MRH_NSYM(Recv, "ncclRecv") MRH_NSYM(AllGather, "ncclAllGather")
    MRH_NSYM(GroupStart, "ncclGroupStart") MRH_NSYM(GroupEnd, "ncclGroupEnd") MRH_NSYM(GetErrorString, "ncclGetErrorString")
#undef MRH_NSYM
    a.ok = true;
    return a;
  }
};
#define NCCL_CK(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) \
    throw std::runtime_error(std::string("CUDA error: NCCL: ") + NcclApi::get().GetErrorString(r_) + " at " #call); } while (0)

// buf[j] = jac[desc[j]] for desc >= 0, res[-desc - 1] otherwise
__global__ void xch_pack_kernel(const int64_t* __restrict__ desc, int64_t n, const double* __restrict__ res, const double* __restrict__ jac, double* __restrict__ buf) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    const int64_t d = desc[j];
    buf[j] = d >= 0 ? jac[d] : res[-d - 1];
  }
}
// first nrows entries: res[rows] += buf; the rest: jac[slot] += buf, or ghost[-slot - 1] += buf.  Atomic: several neighbours may
// hold the same row, and the assembly kernels of this rank may still be adding into it.
__global__ void xch_unpack_kernel(const int64_t* __restrict__ desc, int64_t nrows, int64_t n, const double* __restrict__ buf, double* res, double* jac, double* ghost) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    const int64_t d = desc[j];
    if (j < nrows) atomicAdd(res + d, buf[j]);
    else if (d >= 0) atomicAdd(jac + d, buf[j]);
    else atomicAdd(ghost - d - 1, buf[j]);
  }
}

struct ExchangeDev {
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
  struct PeerDev { int rank; int64_t s_off, s_rows, s_n, r_off, r_rows, r_n; };     // offsets / counts in doubles within the buffers
  std::vector<PeerDev> peers;
  DevBuf<int64_t> s_desc, r_desc;
  DevBuf<double> sendbuf, recvbuf, ghost;
  int64_t nghost = 0;
  cudaStream_t xstream = nullptr;
  cudaEvent_t ev_iface = nullptr, ev_done = nullptr;
  bool ready = false;
  int64_t launches = 0;
  ~ExchangeDev() {
    if (comm && NcclApi::get().ok) NcclApi::get().CommDestroy(comm);
    if (xstream) cudaStreamDestroy(xstream);
    if (ev_iface) cudaEventDestroy(ev_iface);
    if (ev_done) cudaEventDestroy(ev_done);
  }
};

// builds the device tables from the (complete) host plan
void exchange_upload(EngineHost& H, ExchangeDev& X) {
  std::vector<int64_t> sd, rd;
  X.peers.clear();
  for (auto& p : H.plan.peers) {
    ExchangeDev::PeerDev pd{};
    pd.rank = p.rank; pd.s_off = (int64_t)sd.size(); pd.s_rows = (int64_t)p.send_rows.size();
    for (int r : p.send_rows) sd.push_back(-(int64_t)r - 1);
    for (int64_t s : p.send_slots) sd.push_back(s);
    pd.s_n = (int64_t)sd.size() - pd.s_off;
    pd.r_off = (int64_t)rd.size(); pd.r_rows = (int64_t)p.recv_rows.size();
    for (int r : p.recv_rows) rd.push_back((int64_t)r);
    for (int64_t s : p.recv_slots) rd.push_back(s);
    pd.r_n = (int64_t)rd.size() - pd.r_off;
    X.peers.push_back(pd);
  }
  X.s_desc.upload(sd); X.r_desc.upload(rd);
  X.sendbuf.alloc(std::max<size_t>(sd.size(), 1)); X.recvbuf.alloc(std::max<size_t>(rd.size(), 1));
  X.nghost = (int64_t)H.plan.ghost_row.size();
  X.ghost.alloc(std::max<size_t>((size_t)X.nghost, 1));
  if (!X.xstream) CUDA_CK(cudaStreamCreateWithFlags(&X.xstream, cudaStreamNonBlocking));
  if (!X.ev_iface) { CUDA_CK(cudaEventCreateWithFlags(&X.ev_iface, cudaEventDisableTiming)); CUDA_CK(cudaEventCreateWithFlags(&X.ev_done, cudaEventDisableTiming)); }
  X.ready = true;
}

// metadata of the plan moved over NCCL itself: counts by all-gather, then one send / receive per neighbour
void exchange_meta_over_nccl(EngineHost& H, ExchangeDev& X) {
  NcclApi& N = NcclApi::get();
  const int W = X.world, me = X.rank;
  std::vector<std::vector<int64_t>> out(W);
  std::vector<int64_t> mine(W, 0), all((size_t)W * W, 0);
  for (int p = 0; p < W; ++p) if (p != me) { out[p] = H.plan.meta_for(p, H.part, H.crs); mine[p] = (int64_t)out[p].size(); }
  DevBuf<int64_t> d_mine, d_all;
  d_mine.upload(mine); d_all.alloc((size_t)W * W);
  NCCL_CK(N.AllGather(d_mine.p, d_all.p, (size_t)W, ncclInt64, X.comm, 0));
  CUDA_CK(cudaStreamSynchronize(0));
  CUDA_CK(cudaMemcpy(all.data(), d_all.p, all.size() * sizeof(int64_t), cudaMemcpyDeviceToHost));
  std::vector<std::unique_ptr<DevBuf<int64_t>>> sbuf(W), rbuf(W);
  NCCL_CK(N.GroupStart());
  for (int p = 0; p < W; ++p) {
    if (p == me) continue;
    const int64_t ns = mine[p], nr = all[(size_t)p * W + me];
    if (ns > 0) { sbuf[p].reset(new DevBuf<int64_t>()); sbuf[p]->upload(out[p]); NCCL_CK(N.Send(sbuf[p]->p, (size_t)ns, ncclInt64, p, X.comm, 0)); }
    if (nr > 0) { rbuf[p].reset(new DevBuf<int64_t>()); rbuf[p]->alloc((size_t)nr); NCCL_CK(N.Recv(rbuf[p]->p, (size_t)nr, ncclInt64, p, X.comm, 0)); }
  }
  NCCL_CK(N.GroupEnd());
  CUDA_CK(cudaStreamSynchronize(0));
  for (int p = 0; p < W; ++p) {
    if (!rbuf[p]) continue;
    std::vector<int64_t> in(rbuf[p]->n);
    CUDA_CK(cudaMemcpy(in.data(), rbuf[p]->p, in.size() * sizeof(int64_t), cudaMemcpyDeviceToHost));
    H.plan.ingest(p, in.data(), (int64_t)in.size(), H.part, H.crs);
  }
  H.plan.complete = true;
}

// one additive exchange on stream s: rows of DOFs owned elsewhere go to their owner (jac == nullptr: residual only)
void exchange_run(DeviceImpl& D, ExchangeDev& X, double* res, double* jac, cudaStream_t s) {
  NcclApi& N = NcclApi::get();
  X.launches = 0;
  for (auto& p : X.peers) {
    const int64_t n = jac ? p.s_n : p.s_rows;
    if (n <= 0) continue;
    const unsigned grid = (unsigned)std::min<int64_t>((n + 255) / 256, (int64_t)D.num_sms * 4);
    xch_pack_kernel<<<grid, 256, 0, s>>>(X.s_desc.p + p.s_off, n, res, jac, X.sendbuf.p + p.s_off);
    CUDA_CK(cudaGetLastError()); ++X.launches;
  }
  NCCL_CK(N.GroupStart());
  for (auto& p : X.peers) {
    const int64_t ns = jac ? p.s_n : p.s_rows, nr = jac ? p.r_n : p.r_rows;
    if (ns > 0) NCCL_CK(N.Send(X.sendbuf.p + p.s_off, (size_t)ns, ncclDouble, p.rank, X.comm, s));
    if (nr > 0) NCCL_CK(N.Recv(X.recvbuf.p + p.r_off, (size_t)nr, ncclDouble, p.rank, X.comm, s));
  }
  NCCL_CK(N.GroupEnd());
  ++X.launches;
  if (jac && X.nghost > 0) CUDA_CK(cudaMemsetAsync(X.ghost.p, 0, (size_t)X.nghost * sizeof(double), s));
  for (auto& p : X.peers) {
    const int64_t n = jac ? p.r_n : p.r_rows;
    if (n <= 0) continue;
    const unsigned grid = (unsigned)std::min<int64_t>((n + 255) / 256, (int64_t)D.num_sms * 4);
    xch_unpack_kernel<<<grid, 256, 0, s>>>(X.r_desc.p + p.r_off, p.r_rows, n, X.recvbuf.p + p.r_off, res, jac, X.ghost.p);
    CUDA_CK(cudaGetLastError()); ++X.launches;
  }
}

}  // namespace
