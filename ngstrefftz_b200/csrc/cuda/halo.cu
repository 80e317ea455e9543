// Per-level halo exchange of state rows between the GPUs of one box: gather the rows a rank has
// just updated and that other ranks hold, grouped ncclSend/ncclRecv over NVLink, scatter the received rows.
// The reference has no distributed Propagate; a row is one element's wavefront row (src/twavetents.cpp:741).
//
// NCCL is resolved with dlopen at run time (the copy already loaded by the host process -- e.g. torch's -- or the
// system libnccl.so.2), so the library has no link-time NCCL dependency and single-GPU use never touches it.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <stdexcept>
#include <string>

#include "halo.hpp"

namespace ngst {

namespace {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

template <typename T>
void sym(void* lib, const char* name, T& fn) {
  fn = reinterpret_cast<T>(dlsym(lib, name));
  if (!fn) throw std::runtime_error(std::string("COMM: libnccl has no symbol ") + name);
}

NcclApi& api() {
  static NcclApi a;
  if (a.lib) return a;
  void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) throw std::runtime_error(std::string("COMM: cannot load libnccl.so.2: ") + dlerror());
  sym(lib, "ncclGetUniqueId", a.GetUniqueId);
  sym(lib, "ncclCommInitRank", a.CommInitRank);
  sym(lib, "ncclCommDestroy", a.CommDestroy);
  sym(lib, "ncclGroupStart", a.GroupStart);
  sym(lib, "ncclGroupEnd", a.GroupEnd);
  sym(lib, "ncclSend", a.Send);
  sym(lib, "ncclRecv", a.Recv);
  sym(lib, "ncclAllReduce", a.AllReduce);
  sym(lib, "ncclAllGather", a.AllGather);
  sym(lib, "ncclGetErrorString", a.GetErrorString);
  a.lib = lib;
  return a;
}

void nccl_check(ncclResult_t r, const char* what) {
  if (r != ncclSuccess) throw std::runtime_error(std::string("COMM: ") + what + ": " + api().GetErrorString(r));
}
#define NC(x) nccl_check((x), #x)

// one CTA per row; rows are rowlen doubles
__global__ void pack_rows_kernel(const double* __restrict__ wf, const int* __restrict__ rows, int rowlen,
                                 double* __restrict__ buf) {
  const double* src = wf + (size_t)rows[blockIdx.x] * rowlen;
  double* dst = buf + (size_t)blockIdx.x * rowlen;
  for (int i = threadIdx.x; i < rowlen; i += blockDim.x) dst[i] = src[i];
}

__global__ void unpack_rows_kernel(double* __restrict__ wf, const int* __restrict__ rows, int rowlen,
                                   const double* __restrict__ buf) {
  double* dst = wf + (size_t)rows[blockIdx.x] * rowlen;
  const double* src = buf + (size_t)blockIdx.x * rowlen;
  for (int i = threadIdx.x; i < rowlen; i += blockDim.x) dst[i] = src[i];
}

}  // namespace

void halo_unique_id(void* out128) {
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  NC(api().GetUniqueId(&id));
  std::memcpy(out128, &id, sizeof(id));
}

HaloComm::HaloComm(int rank, int nranks, const void* id128) : rank_(rank), nranks_(nranks) {
  ncclUniqueId id;
  std::memcpy(&id, id128, sizeof(id));
  ncclComm_t c = nullptr;
  NC(api().CommInitRank(&c, nranks, id, rank));
  comm_ = c;
}

HaloComm::~HaloComm() {
  if (comm_) api().CommDestroy(static_cast<ncclComm_t>(comm_));
}

void HaloComm::Exchange(double* wf, int rowlen, const int* d_send_rows, const int* send_cnt, const int* d_recv_rows,
                        const int* recv_cnt, double* d_sendbuf, double* d_recvbuf, cudaStream_t stream) {
  int ns = 0, nr = 0;
  for (int q = 0; q < nranks_; q++) { ns += send_cnt[q]; nr += recv_cnt[q]; }
  if (ns == 0 && nr == 0) return;
  if (ns > 0) pack_rows_kernel<<<ns, 128, 0, stream>>>(wf, d_send_rows, rowlen, d_sendbuf);
  ncclComm_t c = static_cast<ncclComm_t>(comm_);
  NC(api().GroupStart());
  size_t so = 0, ro = 0;
  for (int q = 0; q < nranks_; q++) {
    if (send_cnt[q] > 0) {
      NC(api().Send(d_sendbuf + so * rowlen, (size_t)send_cnt[q] * rowlen, ncclDouble, q, c, stream));
      so += send_cnt[q];
    }
    if (recv_cnt[q] > 0) {
      NC(api().Recv(d_recvbuf + ro * rowlen, (size_t)recv_cnt[q] * rowlen, ncclDouble, q, c, stream));
      ro += recv_cnt[q];
    }
  }
  NC(api().GroupEnd());
  if (nr > 0) unpack_rows_kernel<<<nr, 128, 0, stream>>>(wf, d_recv_rows, rowlen, d_recvbuf);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) throw std::runtime_error(std::string("CUDA: halo pack/unpack: ") + cudaGetErrorString(e));
}

void HaloComm::AllGatherBytes(const void* d_in, void* d_out, size_t bytes, cudaStream_t stream) {
  NC(api().AllGather(d_in, d_out, bytes, ncclChar, static_cast<ncclComm_t>(comm_), stream));
}

void HaloComm::Barrier(double* d_scalar, cudaStream_t stream) {
  NC(api().AllReduce(d_scalar, d_scalar, 1, ncclDouble, ncclSum, static_cast<ncclComm_t>(comm_), stream));
}

double HaloComm::AllReduceSum(double local, double* d_scalar, cudaStream_t stream) {
  cudaError_t e = cudaMemcpyAsync(d_scalar, &local, sizeof(double), cudaMemcpyHostToDevice, stream);
  if (e != cudaSuccess) throw std::runtime_error(std::string("CUDA: ") + cudaGetErrorString(e));
  NC(api().AllReduce(d_scalar, d_scalar, 1, ncclDouble, ncclSum, static_cast<ncclComm_t>(comm_), stream));
  double out = 0;
  e = cudaMemcpyAsync(&out, d_scalar, sizeof(double), cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  if (e != cudaSuccess) throw std::runtime_error(std::string("CUDA: ") + cudaGetErrorString(e));
  return out;
}

}  // namespace ngst
