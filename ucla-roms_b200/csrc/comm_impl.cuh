// comm_impl.cuh -- exchange_xxx (mpi_exchanges.F:34-40) implementation; included by romsb200.cu.
//  * one subdomain (np_xi*np_eta == 1): periodic wrap = one device-side halo-copy kernel per <=4 fields;
//  * several subdomains: one pack kernel fills the 8 neighbour messages of up to 4 fields (the reference
//    concatenates up to 4 fields per message too, mpi_exchanges.F:504-506) and one unpack kernel scatters the halos.
//    In between, by default (ROMS_HALO=p2p): the pack kernel has stored the messages DIRECTLY into the neighbours'
//    receive buffers over NVLink peer memory (CUDA IPC mapping of one buffer block per rank, double-buffered by the
//    parity of the exchange counter); its last block raises this rank's flag in every neighbour, and the blocks of the
//    unpack kernel wait for the flag of their direction: two launches per exchange, no NCCL kernel, no proxy thread.  ROMS_HALO=nccl (or no peer access):
//    one grouped ncclSend/ncclRecv over NVLink/NVSwitch.
//    Everything is enqueued on the library's stream: no host synchronisation inside a step.
#pragma once
#include "decomp.h"
#ifdef ROMS_WITH_NCCL
#include <nccl.h>
#endif

struct MsgGeom {
  DirBox box[8];
  double* buf[8];
  int active[8];
};
struct SyncGeom { unsigned long long* to[8]; unsigned long long* from[8]; };
// blockIdx.y = direction; blockIdx.z strides over the concatenated planes of the fields in hl.
// Peer-memory path (counter != nullptr / sg.from set): the PACK launch writes into the neighbours' receive buffers and its
// LAST block to finish raises this rank's flag in every neighbour (slot = the direction this rank is in, seen from
// there); every block of the UNPACK launch first waits for the flag of the direction it unpacks.  A block never waits
// for another block of its own GPU, so the pair cannot deadlock whatever the block scheduling order.
template <bool PACK>
__global__ void k_halo_msg(const __grid_constant__ KP p, const __grid_constant__ HaloList hl, const __grid_constant__ MsgGeom g, int nplanes,
                           const __grid_constant__ SyncGeom sg, unsigned long long epoch, unsigned* counter, int* fail, long long timeout_clk) {
  const int d = blockIdx.y;
  const bool act = g.active[d];
  if (!PACK) {
    if (!act) return;
    if (sg.from[d]) {
      __shared__ int lost;
      if (threadIdx.x == 0) {
        lost = 0;
        const long long t0 = clock64();
        unsigned long long v;
        do {
          asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(sg.from[d]) : "memory");
          // the neighbour never raised its flag within the (configurable, ROMS_HALO_TIMEOUT_S) bound: report and do NOT
          // unpack the stale buffer; the sticky flag is read by the host after the next stream synchronisation
          if (v < epoch && clock64() - t0 > timeout_clk) { *fail = 1; lost = 1; break; }
        } while (v < epoch);
      }
      __syncthreads();
      if (lost) return;
    }
  }
  if (act) {
    const DirBox b = g.box[d];
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell < b.ni * b.nj) {
      const int jj = cell / b.ni, ii = cell - jj * b.ni;
      const long c = CIDX(b.i0 + ii, b.j0 + jj);
      for (int pl = blockIdx.z; pl < nplanes; pl += gridDim.z) {
        int f = 0, k = pl;
        while (k >= hl.nk[f]) { k -= hl.nk[f]; ++f; }
        double* a = hl.a[f] + c + (long)k * p.SXY;
        double* m = g.buf[d] + (long)pl * (b.ni * b.nj) + cell;
        if (PACK) *m = *a; else *a = __ldcg(m);
      }
    }
  }
  if (PACK && counter) {
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned total = gridDim.x * gridDim.y * gridDim.z;
      if (atomicAdd(counter, 1u) == total - 1) {
        *counter = 0;
        __threadfence_system();
        for (int q = 0; q < 8; ++q)
          if (sg.to[q]) asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(sg.to[q]), "l"(epoch) : "memory");
      }
    }
  }
}

struct NcclState {
#ifdef ROMS_WITH_NCCL
  ncclComm_t comm = nullptr;
#endif
  int rank = 0, world = 1;
  double* sbuf[8] = {nullptr};
  double* rbuf[8] = {nullptr};
  double* red = nullptr;
  int nb[8];
  // ---- peer-memory halo path: one IPC-shared block per rank = 8 flags + 2 parities x 8 receive buffers
  int p2p = 0;
  unsigned long long epoch = 0;          // exchanges done so far (identical on every rank: same call sequence)
  char* shm = nullptr;                   // my block
  long long off[2][8];                   // byte offset of my receive buffer [parity][direction it arrives from]
  char* peer_shm[8] = {nullptr};         // block of the neighbour in direction d (my own block for a self-neighbour)
  long long peer_off[8][2][8];           // that neighbour's offsets
  std::vector<std::pair<int, void*>> opened;   // rank -> mapped base (cudaIpcOpenMemHandle)
  int* d_fail = nullptr;                 // pinned host flag: a flag wait timed out
  long long timeout_clk = 0;             // bound of one flag wait in SM clocks (ROMS_HALO_TIMEOUT_S seconds, default 600)
  unsigned* counter = nullptr;           // blocks of the current pack launch that have finished
};
struct P2PInfo { cudaIpcMemHandle_t handle; long long off[2][8]; };
int comm_init(Comm* cm, const roms_cfg* cfg) {
  cm->np_xi = cfg->np_xi > 0 ? cfg->np_xi : 1;
  cm->np_eta = cfg->np_eta > 0 ? cfg->np_eta : 1;
  cm->inode = cfg->inode; cm->jnode = cfg->jnode;
  cm->ew_periodic = cfg->ew_periodic; cm->ns_periodic = cfg->ns_periodic;
  return 0;
}
int comm_destroy(Comm* cm) {
  NcclState* ns = (NcclState*)cm->nccl;
  if (ns) {
    for (int d = 0; d < 8; ++d) { cudaFree(ns->sbuf[d]); cudaFree(ns->rbuf[d]); }
    cudaFree(ns->red);
    for (auto& o : ns->opened) cudaIpcCloseMemHandle(o.second);
    cudaFree(ns->shm);
    cudaFree(ns->counter);
    if (ns->d_fail) cudaFreeHost(ns->d_fail);
#ifdef ROMS_WITH_NCCL
    if (ns->comm) ncclCommDestroy(ns->comm);
#endif
    delete ns;
    cm->nccl = nullptr;
  }
  return 0;
}

#ifdef ROMS_WITH_NCCL
// Peer-memory halo path: every rank allocates ONE block (flags + receive buffers), the ranks all-gather its IPC handle
// and layout, and map the blocks of their neighbours.  Falls back to ncclSend/ncclRecv (on ALL ranks) if any mapping fails.
static int p2p_setup(NcclState* ns, const KP& p) {
  long long cur = 8 * sizeof(unsigned long long);
  cur = (cur + 255) / 256 * 256;
  for (int par = 0; par < 2; ++par)
    for (int d = 0; d < 8; ++d) {
      const DirBox b = decomp_recv_box(p.Lm, p.Mm, d, p.swall, p.nwall, p.wwall, p.ewall);
      ns->off[par][d] = cur;
      cur += ((long long)4 * (p.N + 1) * b.ni * b.nj * sizeof(double) + 255) / 256 * 256;
    }
  CK(cudaMalloc(&ns->shm, cur));
  CK(cudaMemset(ns->shm, 0, cur));
  CK(cudaMalloc(&ns->counter, sizeof(unsigned)));
  CK(cudaMemset(ns->counter, 0, sizeof(unsigned)));
  CK(cudaHostAlloc(&ns->d_fail, sizeof(int), cudaHostAllocMapped));
  *ns->d_fail = 0;
  {
    // all ranks must stay within this bound of each other (host I/O between steps included); MPI would block for ever
    double secs = 600.0;
    if (const char* e = getenv("ROMS_HALO_TIMEOUT_S")) secs = atof(e);
    int dev = 0, khz = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    ns->timeout_clk = (long long)(secs * 1e3 * (khz > 0 ? khz : 1965000));
  }
  P2PInfo mine;
  memset(&mine, 0, sizeof(mine));
  int ok = cudaIpcGetMemHandle(&mine.handle, ns->shm) == cudaSuccess ? 1 : 0;
  memcpy(mine.off, ns->off, sizeof(mine.off));
  std::vector<P2PInfo> all(ns->world);
  char* dbuf = nullptr;
  CK(cudaMalloc(&dbuf, sizeof(P2PInfo) * (ns->world + 1)));
  CK(cudaMemcpy(dbuf, &mine, sizeof(mine), cudaMemcpyHostToDevice));
  ncclResult_t r = ncclAllGather(dbuf, dbuf + sizeof(P2PInfo), sizeof(P2PInfo), ncclChar, ns->comm, 0);
  if (r != ncclSuccess) { g_err = std::string("ncclAllGather(p2p): ") + ncclGetErrorString(r); return -1000 - (int)r; }
  CK(cudaStreamSynchronize(0));
  CK(cudaMemcpy(all.data(), dbuf + sizeof(P2PInfo), sizeof(P2PInfo) * ns->world, cudaMemcpyDeviceToHost));
  for (int d = 0; d < 8 && ok; ++d) {
    const int nb = ns->nb[d];
    if (nb < 0) continue;
    memcpy(ns->peer_off[d], all[nb].off, sizeof(all[nb].off));
    if (nb == ns->rank) { ns->peer_shm[d] = ns->shm; continue; }
    void* base = nullptr;
    for (auto& o : ns->opened) if (o.first == nb) base = o.second;
    if (!base) {
      if (cudaIpcOpenMemHandle(&base, all[nb].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
      ns->opened.push_back({nb, base});
    }
    ns->peer_shm[d] = (char*)base;
  }
  // every rank must take the same path: min over ranks of `ok`
  int* dok = (int*)dbuf;
  CK(cudaMemcpy(dok, &ok, sizeof(int), cudaMemcpyHostToDevice));
  r = ncclAllReduce(dok, dok, 1, ncclInt, ncclMin, ns->comm, 0);
  if (r != ncclSuccess) { g_err = std::string("ncclAllReduce(p2p): ") + ncclGetErrorString(r); return -1000 - (int)r; }
  CK(cudaStreamSynchronize(0));
  CK(cudaMemcpy(&ok, dok, sizeof(int), cudaMemcpyDeviceToHost));
  cudaFree(dbuf);
  ns->p2p = ok;
  return 0;
}
#endif

static int comm_attach(Comm* cm, const KP& p, const char* id, int len, int rank, int world) {
#ifdef ROMS_WITH_NCCL
  if (world != cm->np_xi * cm->np_eta) { g_err = "roms_comm_attach: world size != np_xi*np_eta"; return -1; }
  if (rank != cm->jnode * cm->np_xi + cm->inode) { g_err = "roms_comm_attach: rank != jnode*np_xi+inode (mpi_setup.F:61-70)"; return -1; }
  if (len < (int)sizeof(ncclUniqueId)) { g_err = "roms_comm_attach: id too short"; return -1; }
  NcclState* ns = new NcclState();
  ns->rank = rank; ns->world = world;
  ncclUniqueId uid;
  memcpy(&uid, id, sizeof(uid));
  ncclResult_t r = ncclCommInitRank(&ns->comm, world, uid, rank);
  if (r != ncclSuccess) { g_err = std::string("ncclCommInitRank: ") + ncclGetErrorString(r); delete ns; return -1000 - (int)r; }
  for (int d = 0; d < 8; ++d) {
    ns->nb[d] = decomp_neighbor(cm->np_xi, cm->np_eta, cm->inode, cm->jnode, cm->ew_periodic, cm->ns_periodic, d);
    const DirBox b = decomp_send_box(p.Lm, p.Mm, d, p.swall, p.nwall, p.wwall, p.ewall);
    const size_t cap = (size_t)4 * (p.N + 1) * b.ni * b.nj * sizeof(double);
    CK(cudaMalloc(&ns->sbuf[d], cap));
    CK(cudaMalloc(&ns->rbuf[d], cap));
  }
  CK(cudaMalloc(&ns->red, sizeof(double) * 7 * (world + 1)));
  cm->nccl = ns;
  const char* mode = getenv("ROMS_HALO");
  if (!mode || strcmp(mode, "nccl") != 0) {
    int rc = p2p_setup(ns, p);
    if (rc) return rc;
  }
  return 0;
#else
  (void)cm; (void)p; (void)id; (void)len; (void)rank; (void)world;
  g_err = "library built without NCCL";
  return -1;
#endif
}

static int comm_exchange_local(Comm* cm, const KP& p, cudaStream_t st, const XF* f, int nf, long* launches) {
  if (!cm->ew_periodic && !cm->ns_periodic) return 0;
  for (int n0 = 0; n0 < nf; n0 += 4) {
    HaloList hl;
    hl.n = 0;
    int maxk = 1;
    for (int n = n0; n < nf && n < n0 + 4; ++n) { hl.a[hl.n] = f[n].a; hl.nk[hl.n] = f[n].nk; if (f[n].nk > maxk) maxk = f[n].nk; hl.n++; }
    const int ncell = 4 * (p.Lm + 4) + 4 * p.Mm;
    dim3 b(128, 1, 1), g((ncell + 127) / 128, hl.n, maxk);
    k_halo_periodic<<<g, b, 0, st>>>(p, hl, cm->ew_periodic, cm->ns_periodic);
    (*launches)++;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { g_err = std::string("halo kernel: ") + cudaGetErrorString(e); return -(int)e; }
  return 0;
}

int comm_exchange(Comm* cm, const KP& p, cudaStream_t st, const XF* f, int nf, long* launches) {
  if (cm->np_xi * cm->np_eta == 1) return comm_exchange_local(cm, p, st, f, nf, launches);
#ifdef ROMS_WITH_NCCL
  NcclState* ns = (NcclState*)cm->nccl;
  if (!ns) { g_err = "multi-rank context without communicator: call roms_comm_attach first"; return -1; }
  for (int n0 = 0; n0 < nf; n0 += 4) {
    HaloList hl;
    hl.n = 0;
    int nplanes = 0;
    for (int n = n0; n < nf && n < n0 + 4; ++n) { hl.a[hl.n] = f[n].a; hl.nk[hl.n] = f[n].nk; nplanes += f[n].nk; hl.n++; }
    for (int q = hl.n; q < 4; ++q) { hl.a[q] = nullptr; hl.nk[q] = 1 << 30; }
    MsgGeom gs, gr;
    int maxcell = 4;
    for (int d = 0; d < 8; ++d) {
      gs.box[d] = decomp_send_box(p.Lm, p.Mm, d, p.swall, p.nwall, p.wwall, p.ewall); gs.buf[d] = ns->sbuf[d]; gs.active[d] = ns->nb[d] >= 0;
      gr.box[d] = decomp_recv_box(p.Lm, p.Mm, d, p.swall, p.nwall, p.wwall, p.ewall); gr.buf[d] = ns->rbuf[d]; gr.active[d] = ns->nb[d] >= 0;
      if (gs.box[d].ni * gs.box[d].nj > maxcell) maxcell = gs.box[d].ni * gs.box[d].nj;
    }
    dim3 b(128, 1, 1), g((maxcell + 127) / 128, 8, nplanes < 32 ? nplanes : 32);
    if (ns->p2p) {
      // messages go straight into the neighbours' receive buffers (peer memory over NVLink), then flags
      const int par = (int)(ns->epoch & 1);
      ns->epoch++;
      SyncGeom sg;
      for (int d = 0; d < 8; ++d) {
        const int o = decomp_opposite(d);
        sg.to[d] = sg.from[d] = nullptr;
        if (ns->nb[d] >= 0) {
          gs.buf[d] = (double*)(ns->peer_shm[d] + ns->peer_off[d][par][o]);
          gr.buf[d] = (double*)(ns->shm + ns->off[par][d]);
          sg.to[d] = (unsigned long long*)ns->peer_shm[d] + o;
          sg.from[d] = (unsigned long long*)ns->shm + d;
        }
      }
      k_halo_msg<true><<<g, b, 0, st>>>(p, hl, gs, nplanes, sg, ns->epoch, ns->counter, ns->d_fail, ns->timeout_clk);
      k_halo_msg<false><<<g, b, 0, st>>>(p, hl, gr, nplanes, sg, ns->epoch, nullptr, ns->d_fail, ns->timeout_clk);
      (*launches) += 2;
      continue;
    }
    SyncGeom none;
    memset(&none, 0, sizeof(none));
    k_halo_msg<true><<<g, b, 0, st>>>(p, hl, gs, nplanes, none, 0ull, nullptr, nullptr, 0LL);
    (*launches)++;
    ncclGroupStart();
    for (int d = 0; d < 8; ++d) {
      const int o = decomp_opposite(d);
      const size_t cnt = (size_t)nplanes * gs.box[d].ni * gs.box[d].nj;
      if (ns->nb[d] >= 0) {
        if (ns->nb[d] == ns->rank) cudaMemcpyAsync(ns->rbuf[o], ns->sbuf[d], cnt * sizeof(double), cudaMemcpyDeviceToDevice, st);
        else ncclSend(ns->sbuf[d], cnt, ncclDouble, ns->nb[d], ns->comm, st);
      }
      if (ns->nb[o] >= 0 && ns->nb[o] != ns->rank) ncclRecv(ns->rbuf[o], cnt, ncclDouble, ns->nb[o], ns->comm, st);
    }
    ncclResult_t r = ncclGroupEnd();
    if (r != ncclSuccess) { g_err = std::string("nccl send/recv: ") + ncclGetErrorString(r); return -1000 - (int)r; }
    k_halo_msg<false><<<g, b, 0, st>>>(p, hl, gr, nplanes, none, 0ull, nullptr, nullptr, 0LL);
    (*launches)++;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { g_err = std::string("halo exchange: ") + cudaGetErrorString(e); return -(int)e; }
  return 0;
#else
  (void)p; (void)st; (void)f; (void)nf; (void)launches;
  g_err = "library built without NCCL";
  return -1;
#endif
}

int comm_check_failed(Comm* cm) {
  NcclState* ns = (NcclState*)cm->nccl;
  if (ns && ns->d_fail && *(volatile int*)ns->d_fail) { g_err = "halo exchange: a neighbour never raised its flag within ROMS_HALO_TIMEOUT_S (rank lost?); halos were left untouched"; return -2; }
  return 0;
}

// diag.F:434-504 (rank tree over 13 doubles) -> one ncclAllGather of 7 doubles, combined in rank order
int comm_diag_reduce(Comm* cm, double red[7]) {
  if (cm->np_xi * cm->np_eta == 1) return 0;
#ifdef ROMS_WITH_NCCL
  NcclState* ns = (NcclState*)cm->nccl;
  if (!ns) { g_err = "no communicator"; return -1; }
  cudaStream_t st = 0;
  CK(cudaMemcpy(ns->red, red, 7 * sizeof(double), cudaMemcpyHostToDevice));
  ncclResult_t r = ncclAllGather(ns->red, ns->red + 7, 7, ncclDouble, ns->comm, st);
  if (r != ncclSuccess) { g_err = std::string("ncclAllGather: ") + ncclGetErrorString(r); return -1000 - (int)r; }
  CK(cudaStreamSynchronize(st));
  std::vector<double> all(7 * ns->world);
  CK(cudaMemcpy(all.data(), ns->red + 7, all.size() * sizeof(double), cudaMemcpyDeviceToHost));
  for (int q = 0; q < 7; ++q) red[q] = 0.0;
  for (int rk = 0; rk < ns->world; ++rk) {
    for (int q = 0; q < 5; ++q) red[q] += all[7 * rk + q];
    if (all[7 * rk + 5] > red[5]) { red[5] = all[7 * rk + 5]; red[6] = all[7 * rk + 6]; }
  }
  return 0;
#else
  return -1;
#endif
}

extern "C" {
int roms_comm_unique_id(char* out, int len) {
#ifdef ROMS_WITH_NCCL
  if (len < (int)sizeof(ncclUniqueId)) { g_err = "roms_comm_unique_id: buffer too small"; return -1; }
  ncclUniqueId uid;
  ncclResult_t r = ncclGetUniqueId(&uid);
  if (r != ncclSuccess) { g_err = std::string("ncclGetUniqueId: ") + ncclGetErrorString(r); return -1000 - (int)r; }
  memset(out, 0, len);
  memcpy(out, &uid, sizeof(uid));
  return 0;
#else
  (void)out; (void)len; g_err = "library built without NCCL"; return -1;
#endif
}
int roms_comm_attach(roms_ctx* h, const char* id, int len, int rank, int world) {
  CK(cudaSetDevice(h->device));
  return comm_attach(&h->comm, h->kp, id, len, rank, world);
}
// host-only geometry, exported so that the decomposition logic is testable without a GPU
int roms_decomp_neighbor(int np_xi, int np_eta, int inode, int jnode, int ew_periodic, int ns_periodic, int d) {
  return decomp_neighbor(np_xi, np_eta, inode, jnode, ew_periodic, ns_periodic, d);
}
int roms_decomp_box(int Lm, int Mm, int d, int send, const int walls[4], int out[4]) {
  const int sw = walls ? walls[0] : 0, nw = walls ? walls[1] : 0, ww = walls ? walls[2] : 0, ew = walls ? walls[3] : 0;
  const DirBox b = send ? decomp_send_box(Lm, Mm, d, sw, nw, ww, ew) : decomp_recv_box(Lm, Mm, d, sw, nw, ww, ew);
  out[0] = b.i0; out[1] = b.j0; out[2] = b.ni; out[3] = b.nj;
  return 0;
}
int roms_decomp_opposite(int d) { return decomp_opposite(d); }
int roms_decomp_extent(int LL, int np, int node, int* L, int* corn) {
  if (LL < 1 || np < 1 || node < 0 || node >= np || !L || !corn) return 1;
  decomp_extent(LL, np, node, L, corn);
  return *L >= 2 ? 0 : 1;   // a block narrower than the halo cannot be exchanged
}
}
