// comm.h -- exchange_xxx (mpi_exchanges.F:34-40) for one subdomain per GPU.
// Single-rank subdomains: the periodic wrap of mpi_setup.F:231-250 is a device-side halo copy.
// Multi-rank (np_xi*np_eta > 1): see comm_impl.cuh (pack straight into the neighbours' buffers over NVLink peer memory + flags ->
// unpack; ROMS_HALO=nccl: pack kernel -> grouped ncclSend/ncclRecv -> unpack).
#pragma once
#include <cuda_runtime.h>
#include "../../include/roms_b200.h"
#include "kp.h"

struct XF { double* a; int nk; };

struct Comm {
  int np_xi = 1, np_eta = 1, inode = 0, jnode = 0;
  int ew_periodic = 0, ns_periodic = 0;
  void* nccl = nullptr;      // NcclState* when multi-rank
};

int comm_init(Comm* cm, const roms_cfg* cfg);
int comm_destroy(Comm* cm);
int comm_exchange(Comm* cm, const KP& p, cudaStream_t st, const XF* f, int nf, long* launches);
int comm_diag_reduce(Comm* cm, double red[7]);
// after a stream synchronisation: non-zero if a halo wait of any earlier exchange timed out (a neighbour rank was lost)
int comm_check_failed(Comm* cm);
