// TEST INFRASTRUCTURE -- see cuda_runtime.h in this directory.  Declarations only: comm.cu binds NCCL with dlopen
// and only takes decltype() of these; the emulation runs one rank, so none of them is ever called.
#pragma once
#include "cuda_runtime.h"
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef enum { ncclSuccess = 0, ncclUnhandledCudaError = 1 } ncclResult_t;
typedef enum { ncclChar = 0, ncclInt = 2, ncclDouble = 8 } ncclDataType_t;
typedef enum { ncclSum = 0, ncclMax = 2 } ncclRedOp_t;
extern "C" {
ncclResult_t ncclGetUniqueId(ncclUniqueId *id);
ncclResult_t ncclCommInitRank(ncclComm_t *comm, int nranks, ncclUniqueId id, int rank);
ncclResult_t ncclCommDestroy(ncclComm_t comm);
const char *ncclGetErrorString(ncclResult_t r);
ncclResult_t ncclGroupStart();
ncclResult_t ncclGroupEnd();
ncclResult_t ncclSend(const void *buf, size_t count, ncclDataType_t t, int peer, ncclComm_t comm, cudaStream_t s);
ncclResult_t ncclRecv(void *buf, size_t count, ncclDataType_t t, int peer, ncclComm_t comm, cudaStream_t s);
ncclResult_t ncclAllReduce(const void *s, void *r, size_t count, ncclDataType_t t, ncclRedOp_t op, ncclComm_t comm, cudaStream_t st);
ncclResult_t ncclAllGather(const void *s, void *r, size_t count, ncclDataType_t t, ncclComm_t comm, cudaStream_t st);
}
