// TEST INFRASTRUCTURE -- see cuda_runtime.h in this directory.  The few NCCL calls comm.cu binds with dlopen, for
// emulated ranks that are threads of one process: point-to-point groups, all-reduce (double sum, int max) and
// all-gather through a shared mailbox with two rendezvous per call.  A rank that does not show up within two minutes
// (an error on another rank) aborts the process with a message instead of hanging the test run.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <condition_variable>
#include <deque>
#include <map>
#include <mutex>
#include <vector>

typedef struct emu_stream *cudaStream_t;
typedef enum { ncclSuccess = 0, ncclUnhandledCudaError = 1, ncclInvalidArgument = 4 } ncclResult_t;
typedef enum { ncclChar = 0, ncclInt = 2, ncclDouble = 8 } ncclDataType_t;
typedef enum { ncclSum = 0, ncclMax = 2 } ncclRedOp_t;
typedef struct { char internal[128]; } ncclUniqueId;

namespace {
struct Shared {
  int nranks = 0, joined = 0;
  std::mutex m;
  std::condition_variable cv;
  int arrived = 0;
  unsigned gen = 0;
  std::map<std::pair<int, int>, std::deque<std::vector<char>>> mail;   // (src, dst) -> messages in posting order
  std::vector<std::vector<char>> stage;
  bool broken = false;
  // false = the communicator is broken (a rank failed, or did not show up): every caller gets an error, nobody hangs
  bool barrier()
  {
    std::unique_lock<std::mutex> l(m);
    if (broken) return false;
    const unsigned g = gen;
    if (++arrived == nranks) {
      arrived = 0;
      gen++;
      cv.notify_all();
      return true;
    }
    if (!cv.wait_for(l, std::chrono::seconds(120), [&] { return gen != g || broken; })) {
      fprintf(stderr, "fake nccl: %d of %d ranks reached a collective within 120 s -- another rank failed or took a different path\n", arrived, nranks);
      broken = true;
      cv.notify_all();
    }
    return !broken;
  }
};
struct Comm { Shared *s; int rank; };
struct Op { bool send; Comm *c; void *buf; size_t bytes; int peer; };
std::mutex g_m;
std::map<uint64_t, Shared *> g_comms;
uint64_t g_next = 1;
thread_local std::vector<Op> t_ops;
thread_local int t_depth = 0;
size_t width(ncclDataType_t t) { return t == ncclChar ? 1 : t == ncclInt ? 4 : 8; }

ncclResult_t flush()
{
  if (t_ops.empty()) return ncclSuccess;
  Shared *s = t_ops[0].c->s;
  const int me = t_ops[0].c->rank;
  {
    std::lock_guard<std::mutex> l(s->m);
    for (const Op &o : t_ops)
      if (o.send) s->mail[{me, o.peer}].emplace_back((const char *) o.buf, (const char *) o.buf + o.bytes);
  }
  if (!s->barrier()) return ncclUnhandledCudaError;
  {
    std::lock_guard<std::mutex> l(s->m);
    for (const Op &o : t_ops)
      if (!o.send) {
        auto &q = s->mail[{o.peer, me}];
        if (q.empty() || q.front().size() != o.bytes) {
          fprintf(stderr, "fake nccl: rank %d expects %zu bytes from rank %d, %s\n", me, o.bytes, o.peer,
                  q.empty() ? "nothing was sent" : "a message of another size was sent");
          s->broken = true;
          s->cv.notify_all();
          t_ops.clear();
          return ncclUnhandledCudaError;
        }
        memcpy(o.buf, q.front().data(), o.bytes);
        q.pop_front();
      }
  }
  if (!s->barrier()) return ncclUnhandledCudaError;
  t_ops.clear();
  return ncclSuccess;
}
}    // namespace

extern "C" {
ncclResult_t ncclGetUniqueId(ncclUniqueId *id)
{
  std::lock_guard<std::mutex> l(g_m);
  memset(id, 0, sizeof *id);
  const uint64_t v = g_next++;
  memcpy(id->internal, &v, 8);
  return ncclSuccess;
}
ncclResult_t ncclCommInitRank(Comm **comm, int nranks, ncclUniqueId id, int rank)
{
  uint64_t key;
  memcpy(&key, id.internal, 8);
  Shared *s;
  {
    std::lock_guard<std::mutex> l(g_m);
    Shared *&slot = g_comms[key];
    if (!slot) {
      slot = new Shared();
      slot->nranks = nranks;
      slot->stage.resize(nranks);
    }
    s = slot;
  }
  if (s->nranks != nranks || rank < 0 || rank >= nranks) return ncclInvalidArgument;
  *comm = new Comm{s, rank};
  if (!s->barrier()) return ncclUnhandledCudaError;
  return ncclSuccess;
}
ncclResult_t ncclCommDestroy(Comm *c) { delete c; return ncclSuccess; }
const char *ncclGetErrorString(ncclResult_t r) { return r == ncclSuccess ? "no error" : "fake nccl: communicator broken (a rank failed or did not arrive)"; }
// test harness: a rank died outside the library; release everybody who waits for it
void fake_nccl_break_all()
{
  std::lock_guard<std::mutex> l(g_m);
  for (auto &kv : g_comms) {
    std::lock_guard<std::mutex> l2(kv.second->m);
    kv.second->broken = true;
    kv.second->cv.notify_all();
  }
}
ncclResult_t ncclGroupStart() { t_depth++; return ncclSuccess; }
ncclResult_t ncclGroupEnd() { return --t_depth == 0 ? flush() : ncclSuccess; }
ncclResult_t ncclSend(const void *buf, size_t count, ncclDataType_t t, int peer, Comm *c, cudaStream_t)
{
  t_ops.push_back({true, c, (void *) buf, count * width(t), peer});
  return t_depth ? ncclSuccess : flush();
}
ncclResult_t ncclRecv(void *buf, size_t count, ncclDataType_t t, int peer, Comm *c, cudaStream_t)
{
  t_ops.push_back({false, c, buf, count * width(t), peer});
  return t_depth ? ncclSuccess : flush();
}
ncclResult_t ncclAllGather(const void *sbuf, void *rbuf, size_t count, ncclDataType_t t, Comm *c, cudaStream_t)
{
  Shared *s = c->s;
  const size_t bytes = count * width(t);
  {
    std::lock_guard<std::mutex> l(s->m);
    s->stage[c->rank].assign((const char *) sbuf, (const char *) sbuf + bytes);
  }
  if (!s->barrier()) return ncclUnhandledCudaError;
  for (int r = 0; r < s->nranks; r++) {
    if (s->stage[r].size() != bytes) { fprintf(stderr, "fake nccl: all-gather sizes differ between ranks\n"); return ncclUnhandledCudaError; }
    memcpy((char *) rbuf + (size_t) r * bytes, s->stage[r].data(), bytes);
  }
  if (!s->barrier()) return ncclUnhandledCudaError;
  return ncclSuccess;
}
ncclResult_t ncclAllReduce(const void *sbuf, void *rbuf, size_t count, ncclDataType_t t, ncclRedOp_t op, Comm *c, cudaStream_t)
{
  Shared *s = c->s;
  const size_t bytes = count * width(t);
  if (!((t == ncclDouble && op == ncclSum) || (t == ncclInt && op == ncclMax))) return ncclInvalidArgument;
  {
    std::lock_guard<std::mutex> l(s->m);
    s->stage[c->rank].assign((const char *) sbuf, (const char *) sbuf + bytes);
  }
  if (!s->barrier()) return ncclUnhandledCudaError;
  for (size_t k = 0; k < count; k++) {
    if (t == ncclDouble) {
      double acc = 0.0;   // rank order, the same on every rank
      for (int r = 0; r < s->nranks; r++) acc += ((const double *) s->stage[r].data())[k];
      ((double *) rbuf)[k] = acc;
    } else {
      int acc = ((const int *) s->stage[0].data())[k];
      for (int r = 1; r < s->nranks; r++) acc = std::max(acc, ((const int *) s->stage[r].data())[k]);
      ((int *) rbuf)[k] = acc;
    }
  }
  if (!s->barrier()) return ncclUnhandledCudaError;
  return ncclSuccess;
}
}
