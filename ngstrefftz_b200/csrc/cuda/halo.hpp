// Inter-GPU exchange of state rows (halo.cu). One HaloComm per solver handle; NCCL over NVLink / NVSwitch.
#pragma once
#include <cuda_runtime.h>

#include <cstring>

namespace ngst {

// 128-byte NCCL unique id, created by one rank and distributed by the caller (e.g. torch.distributed broadcast)
void halo_unique_id(void* out128);

class HaloComm {
 public:
  HaloComm(int rank, int nranks, const void* id128);
  ~HaloComm();
  HaloComm(const HaloComm&) = delete;
  HaloComm& operator=(const HaloComm&) = delete;
  int rank() const { return rank_; }
  int nranks() const { return nranks_; }

  // send_cnt / recv_cnt: rows per peer (host arrays of nranks entries); d_send_rows / d_recv_rows: the local element ids,
  // peer after peer; buffers hold (sum of counts) * rowlen doubles. Everything is enqueued on `stream`.
  void Exchange(double* wf, int rowlen, const int* d_send_rows, const int* send_cnt, const int* d_recv_rows,
                const int* recv_cnt, double* d_sendbuf, double* d_recvbuf, cudaStream_t stream);
  // every rank contributes `bytes` bytes (device buffer) and receives all contributions in rank order (device buffer)
  void AllGatherBytes(const void* d_in, void* d_out, size_t bytes, cudaStream_t stream);
  // stream-ordered barrier: completes on `stream` once every rank has reached it on its stream (no host synchronisation)
  void Barrier(double* d_scalar, cudaStream_t stream);
  // FP64 sum over ranks of one scalar (Energy / Error); blocks until the result is on the host
  double AllReduceSum(double local, double* d_scalar, cudaStream_t stream);

 private:
  int rank_, nranks_;
  void* comm_ = nullptr;  // ncclComm_t
};

}  // namespace ngst
