// vq_nccl.h -- NCCL bound at run time (dlopen "libnccl.so.2"), so the library
// loads on hosts without NCCL and shares the copy torch already mapped when the
// host process is a torchrun rank.  Only the handful of entry points the VQ path
// needs: one allreduce(sum, f64) per k-means pass over NVLink / NVSwitch.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace vqb200 {

struct NcclUniqueId { char internal[128]; };  // == ncclUniqueId (nccl.h:37-38)

struct NcclApi {
    int (*GetUniqueId)(NcclUniqueId *);
    int (*CommInitRank)(void **comm, int nranks, NcclUniqueId id, int rank);
    int (*CommInitAll)(void **comms, int ndev, const int *devlist);
    int (*CommDestroy)(void *comm);
    int (*AllReduce)(const void *send, void *recv, size_t count, int dtype, int op, void *comm,
                     cudaStream_t stream);
    int (*GroupStart)();
    int (*GroupEnd)();
    const char *(*GetErrorString)(int);
};
constexpr int kNcclFloat64 = 8;  // nccl.h:286
constexpr int kNcclSum = 0;      // nccl.h:260

// nullptr (and last_error set) if libnccl cannot be loaded
const NcclApi *nccl_api();

}  // namespace vqb200
