// comm.cu -- multi-rank transport (replaces the MPI calls of CommBrick, src/comm_brick.cpp:538-1053).
// NCCL is bound at run time with dlopen so that a process which already carries torch's NCCL uses that
// one copy.  Single-rank contexts never touch this file's NCCL paths.
#include <dlfcn.h>

#include "ctx.cuh"

int comm_init(b200md_ctx *c, const void *nccl_id)
{
  (void) nccl_id;
  return b200md_fail(c, B200MD_ERR_ARG, "multi-rank contexts are not available in this build yet");
}

void comm_destroy(b200md_ctx *c) { (void) c; }

int comm_allreduce_sum(b200md_ctx *c, double *dev, int n)
{
  (void) dev; (void) n;
  return b200md_fail(c, B200MD_ERR_ARG, "multi-rank contexts are not available in this build yet");
}

extern "C" int b200md_nccl_unique_id(void *id_out)
{
  (void) id_out;
  return b200md_fail(nullptr, B200MD_ERR_ARG, "multi-rank contexts are not available in this build yet");
}
