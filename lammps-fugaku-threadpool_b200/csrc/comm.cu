// comm.cu -- ghost atoms and atom migration: what Comm/CommBrick does for the hot path
//   CommBrick::setup        (src/comm_brick.cpp:191-483)   -> comm_plan         (peers, shifts)
//   CommBrick::borders      (src/comm_brick.cpp:771-979)   -> comm_borders      (send lists, ghost creation)
//   CommBrick::forward_comm (src/comm_brick.cpp:538-591)   -> halo_forward_pos
//   Comm::forward_comm(Pair*) (src/comm_brick.cpp:1063-1112, PairEAM::pack/unpack_forward_comm
//                            src/MANYBODY/pair_eam.cpp:855-876) -> halo_forward_scalar
//   CommBrick::exchange     (src/comm_brick.cpp:652-759)   -> comm_exchange     (atoms that left the sub-domain)
// re-designed for one GPU per rank on an NVSwitch node: a SINGLE-stage exchange with the (up to) 26 neighbor
// directions instead of 6 staged swaps.  An atom within the ghost cutoff of a face/edge/corner is sent straight
// to the rank across that face/edge/corner (periodic shift applied by the sender, AtomVec::pack_border
// src/atom_vec.cpp:808-826), which yields the same ghost set as the staged scheme whenever every sub-domain is at
// least one ghost cutoff wide, without the dependency chain between dimensions.  Directions whose peer is this
// rank itself (periodic dimension with one rank) are plain device copies; the others are ncclSend/ncclRecv pairs
// in one group on the context's stream, received directly into the ghost slots of the position array.
// NCCL is bound at run time with dlopen, so a host process that already carries an NCCL (torch) shares it.
#include <dlfcn.h>
#include <nccl.h>

#include <utility>
#include <vector>

#include "ctx.cuh"

// ------------------------------------------------------------------------------------------------
// NCCL binding
// ------------------------------------------------------------------------------------------------
namespace {
struct NcclApi {
  void *h = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclSend) Send = nullptr;
  decltype(&ncclRecv) Recv = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
} g_nccl;

const char *nccl_load()
{
  if (g_nccl.h) return nullptr;
  const char *names[] = {"libnccl.so.2", "libnccl.so"};
  void *h = nullptr;
  for (const char *n : names) {
    h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) return "cannot dlopen libnccl.so.2";
#define BIND(name)                                                        \
  g_nccl.name = (decltype(g_nccl.name)) dlsym(h, "nccl" #name);           \
  if (!g_nccl.name) return "libnccl lacks nccl" #name
  BIND(GetUniqueId); BIND(CommInitRank); BIND(CommDestroy); BIND(GetErrorString); BIND(GroupStart); BIND(GroupEnd);
  BIND(Send); BIND(Recv); BIND(AllReduce); BIND(AllGather);
#undef BIND
  g_nccl.h = h;
  return nullptr;
}
}  // namespace

#define NC(call)                                                                                          \
  do {                                                                                                    \
    ncclResult_t r__ = (call);                                                                            \
    if (r__ != ncclSuccess)                                                                               \
      return b200md_fail(c, B200MD_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r__)); \
  } while (0)

extern "C" int b200md_nccl_unique_id(void *id_out)
{
  b200md_ctx *c = nullptr;
  static_assert(sizeof(ncclUniqueId) <= B200MD_NCCL_ID_BYTES, "id size");
  if (!id_out) return b200md_fail(c, B200MD_ERR_ARG, "id_out == NULL");
  const char *why = nccl_load();
  if (why) return b200md_fail(c, B200MD_ERR_CUDA, "%s", why);
  ncclUniqueId id;
  NC(g_nccl.GetUniqueId(&id));
  memset(id_out, 0, B200MD_NCCL_ID_BYTES);
  memcpy(id_out, &id, sizeof(id));
  return B200MD_OK;
}

int comm_init(b200md_ctx *c, const void *nccl_id)
{
  if (!nccl_id) return b200md_fail(c, B200MD_ERR_ARG, "nranks > 1 needs the id from b200md_nccl_unique_id()");
  const char *why = nccl_load();
  if (why) return b200md_fail(c, B200MD_ERR_CUDA, "%s", why);
  ncclUniqueId id;
  memcpy(&id, nccl_id, sizeof(id));
  ncclComm_t comm = nullptr;
  NC(g_nccl.CommInitRank(&comm, c->nranks, id, c->rank));
  c->nccl = (void *) comm;
  return B200MD_OK;
}

void comm_destroy(b200md_ctx *c)
{
  for (int r = 0; r < 64; r++)
    if (c->put.peer_base[r]) { cudaIpcCloseMemHandle(c->put.peer_base[r]); c->put.peer_base[r] = nullptr; }
  if (c->put.base) cudaFree(c->put.base);
  if (c->put.done_d) cudaFree(c->put.done_d);
  c->put.base = nullptr;
  c->put.done_d = nullptr;
  c->put.imported = c->put.epoch_ok = false;
  if (c->nccl) g_nccl.CommDestroy((ncclComm_t) c->nccl);
  c->nccl = nullptr;
}

int comm_allreduce_sum(b200md_ctx *c, double *dev, int n)
{
  if (c->nranks == 1) return B200MD_OK;
  NC(g_nccl.AllReduce(dev, dev, (size_t) n, ncclDouble, ncclSum, (ncclComm_t) c->nccl, c->stream));
  return B200MD_OK;
}

int comm_allreduce_max_int(b200md_ctx *c, int *dev, int n)
{
  if (c->nranks == 1) return B200MD_OK;
  NC(g_nccl.AllReduce(dev, dev, (size_t) n, ncclInt, ncclMax, (ncclComm_t) c->nccl, c->stream));
  return B200MD_OK;
}

// ------------------------------------------------------------------------------------------------
// plan: direction d in 0..25 <-> (dx,dy,dz) in {-1,0,1}^3 \ {0}; peers on the rank grid
// ------------------------------------------------------------------------------------------------
static void dir_vec(int d, int v[3])
{
  int full = d < 13 ? d : d + 1;
  v[0] = full % 3 - 1;
  v[1] = (full / 3) % 3 - 1;
  v[2] = full / 9 - 1;
}

// rank at grid location loc + sign*dv, wrapped through periodic dimensions; -1 if outside a non-periodic box
static int grid_peer(const int procgrid[3], const int periodic[3], const int loc[3], const int dv[3], int sign)
{
  int l[3];
  for (int k = 0; k < 3; k++) {
    l[k] = loc[k] + sign * dv[k];
    if (l[k] < 0 || l[k] >= procgrid[k]) {
      if (!periodic[k]) return -1;
      l[k] = (l[k] + procgrid[k]) % procgrid[k];
    }
  }
  return (l[2] * procgrid[1] + l[1]) * procgrid[0] + l[0];
}

// host-only, no device needed: who receives what this rank sends in direction d (peer_send[d]) and whose
// direction-d sends this rank receives (peer_recv[d]); -1 = nobody.  Exposed for the CPU-side tests.
extern "C" int b200md_plan_peers(const int procgrid[3], const int periodic[3], int rank, int *peer_send, int *peer_recv)
{
  const int np = procgrid[0] * procgrid[1] * procgrid[2];
  if (np < 1 || rank < 0 || rank >= np) return B200MD_ERR_ARG;
  int loc[3] = {rank % procgrid[0], (rank / procgrid[0]) % procgrid[1], rank / (procgrid[0] * procgrid[1])};
  for (int d = 0; d < NDIR; d++) {
    int dv[3];
    dir_vec(d, dv);
    peer_send[d] = grid_peer(procgrid, periodic, loc, dv, +1);
    peer_recv[d] = grid_peer(procgrid, periodic, loc, dv, -1);
  }
  return B200MD_OK;
}

int comm_plan(b200md_ctx *c)
{
  b200md_plan_peers(c->procgrid, c->periodic, c->rank, c->peer_send, c->peer_recv);
  for (int d = 0; d < NDIR; d++) {
    int dv[3];
    dir_vec(d, dv);
    for (int k = 0; k < 3; k++) {
      double s = 0.0;  // shift applied by the SENDER when the message crosses the periodic boundary
      if (dv[k] < 0 && c->myloc[k] == 0) s = c->prd[k];
      if (dv[k] > 0 && c->myloc[k] == c->procgrid[k] - 1) s = -c->prd[k];
      c->hp.shift[d][k] = s;
    }
    c->hp.self[d] = (c->peer_send[d] == c->rank) ? 1 : 0;
  }
  return B200MD_OK;
}

// ------------------------------------------------------------------------------------------------
// per-atom direction masks -> per-direction, order-preserving index lists (count / scan / fill)
//   MODE 0 (borders):  bit d set if the atom lies within the ghost cutoff of every face direction d crosses;
//                      slab tests inclusive like the reference (src/comm_brick.cpp:836-841)
//   MODE 1 (exchange): exactly one bit: d if the atom left through direction d, bit 26 if it stays
//                      (x < sublo -> lower rank, x >= subhi -> upper rank; src/comm_brick.cpp:700-707)
// ------------------------------------------------------------------------------------------------
#define NBITS (NDIR + 1)
struct MaskGeom {
  double lo_cut[3], hi_cut[3];  // MODE 0: sublo + cutghost, subhi - cutghost.  MODE 1: sublo, subhi
  int send_lo[3], send_hi[3];   // a neighbor exists on that side
  // MODE 1, more than two ranks in a periodic dimension: an atom that Domain::pbc wrapped to the far end of the
  // box belongs to the neighbor on the OTHER side (x >= wrap_to_lo -> lower rank, x < wrap_to_hi -> upper rank)
  double wrap_to_lo[3], wrap_to_hi[3];
};

__device__ __forceinline__ unsigned dirs_to_mask(int lo0, int hi0, int lo1, int hi1, int lo2, int hi2)
{
  unsigned m = 0;
  int c = 0;
#pragma unroll
  for (int dz = -1; dz <= 1; dz++)
#pragma unroll
    for (int dy = -1; dy <= 1; dy++)
#pragma unroll
      for (int dx = -1; dx <= 1; dx++) {
        if (dx == 0 && dy == 0 && dz == 0) continue;
        int ok = (dx == 0 || (dx < 0 ? lo0 : hi0)) && (dy == 0 || (dy < 0 ? lo1 : hi1)) && (dz == 0 || (dz < 0 ? lo2 : hi2));
        m |= (unsigned) ok << c;
        c++;
      }
  return m;
}

template <int MODE> __device__ __forceinline__ unsigned atom_mask(const MaskGeom &b, const double4 &p)
{
  if (MODE == 0) {
    int lo0 = b.send_lo[0] && p.x <= b.lo_cut[0], hi0 = b.send_hi[0] && p.x >= b.hi_cut[0];
    int lo1 = b.send_lo[1] && p.y <= b.lo_cut[1], hi1 = b.send_hi[1] && p.y >= b.hi_cut[1];
    int lo2 = b.send_lo[2] && p.z <= b.lo_cut[2], hi2 = b.send_hi[2] && p.z >= b.hi_cut[2];
    if (!(lo0 | hi0 | lo1 | hi1 | lo2 | hi2)) return 0u;
    return dirs_to_mask(lo0, hi0, lo1, hi1, lo2, hi2);
  } else {
    int lo0 = p.x < b.lo_cut[0], hi0 = p.x >= b.hi_cut[0];
    int lo1 = p.y < b.lo_cut[1], hi1 = p.y >= b.hi_cut[1];
    int lo2 = p.z < b.lo_cut[2], hi2 = p.z >= b.hi_cut[2];
    if (!(lo0 | hi0 | lo1 | hi1 | lo2 | hi2)) return 1u << NDIR;  // stays
    if (hi0 && p.x >= b.wrap_to_lo[0]) { hi0 = 0; lo0 = 1; } else if (lo0 && p.x < b.wrap_to_hi[0]) { lo0 = 0; hi0 = 1; }
    if (hi1 && p.y >= b.wrap_to_lo[1]) { hi1 = 0; lo1 = 1; } else if (lo1 && p.y < b.wrap_to_hi[1]) { lo1 = 0; hi1 = 1; }
    if (hi2 && p.z >= b.wrap_to_lo[2]) { hi2 = 0; lo2 = 1; } else if (lo2 && p.z < b.wrap_to_hi[2]) { lo2 = 0; hi2 = 1; }
    // the one direction with dv_k = -1 / 0 / +1 per dimension
    int full = (lo0 ? 0 : hi0 ? 2 : 1) + 3 * (lo1 ? 0 : hi1 ? 2 : 1) + 9 * (lo2 ? 0 : hi2 ? 2 : 1);
    int d = full < 13 ? full : full - 1;
    return 1u << d;
  }
}

#define MASK_BLOCK 256
template <int MODE>
__global__ void k_mask_count(const double4 *__restrict__ pos, int n, MaskGeom mg, int nblk, int *__restrict__ blockcnt)
{
  __shared__ int cnt[NBITS];
  if (threadIdx.x < NBITS) cnt[threadIdx.x] = 0;
  __syncthreads();
  int i = blockIdx.x * MASK_BLOCK + threadIdx.x;
  unsigned m = 0;
  if (i < n) m = atom_mask<MODE>(mg, pos[i]);
  if (__any_sync(0xffffffffu, m != 0)) {
    for (int d = 0; d < NBITS; d++) {
      unsigned b = __ballot_sync(0xffffffffu, (m >> d) & 1u);
      if (b && (threadIdx.x & 31) == 0) atomicAdd(&cnt[d], __popc(b));
    }
  }
  __syncthreads();
  if (threadIdx.x < NBITS) blockcnt[threadIdx.x * nblk + blockIdx.x] = cnt[threadIdx.x];
}

// one block per direction: exclusive scan of that direction's per-block counts, total to totals[d]
__global__ void k_mask_scan(const int *__restrict__ blockcnt, int *__restrict__ blockoff, int nblk, int *__restrict__ totals)
{
  __shared__ int wsum[32];
  __shared__ int carry_s;
  int d = blockIdx.x;
  const int *in = blockcnt + (size_t) d * nblk;
  int *out = blockoff + (size_t) d * nblk;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int start = 0; start < nblk; start += blockDim.x) {
    int i = start + threadIdx.x;
    int t = (i < nblk) ? in[i] : 0;
    int incl = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) wsum[w] = incl;
    __syncthreads();
    int woff = 0;
    for (int k = 0; k < w; k++) woff += wsum[k];
    int carry = carry_s;
    if (i < nblk) out[i] = carry + woff + incl - t;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry_s = carry + woff + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) totals[d] = carry_s;
}

struct SendOff { int off[NBITS + 1]; };

template <int MODE>
__global__ void k_mask_fill(const double4 *__restrict__ pos, int n, MaskGeom mg, int nblk, const int *__restrict__ blockcnt,
                            const int *__restrict__ blockoff, SendOff so, int *__restrict__ send_idx,
                            signed char *__restrict__ send_dir)
{
  __shared__ int wcnt[MASK_BLOCK / 32];
  __shared__ int s_cnt[NBITS];   // this block's count per direction: most blocks (interior atoms) have none at all
  __shared__ int s_any;
  if (threadIdx.x == 0) s_any = 0;
  __syncthreads();
  if (threadIdx.x < NBITS) {
    const int cn = blockcnt[threadIdx.x * nblk + blockIdx.x];
    s_cnt[threadIdx.x] = cn;
    if (cn) s_any = 1;
  }
  __syncthreads();
  if (!s_any) return;   // block-uniform
  int i = blockIdx.x * MASK_BLOCK + threadIdx.x;
  unsigned m = 0;
  if (i < n) m = atom_mask<MODE>(mg, pos[i]);
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int d = 0; d < NBITS; d++) {
    if (s_cnt[d] == 0) continue;  // block-uniform
    unsigned b = __ballot_sync(0xffffffffu, (m >> d) & 1u);
    if (lane == 0) wcnt[w] = __popc(b);
    __syncthreads();
    if ((m >> d) & 1u) {
      int base = 0;
      for (int k = 0; k < w; k++) base += wcnt[k];
      int r = base + __popc(b & ((1u << lane) - 1u));
      int e = so.off[d] + blockoff[d * nblk + blockIdx.x] + r;
      send_idx[e] = i;
      send_dir[e] = (signed char) d;
    }
    __syncthreads();
  }
}

// counts per direction of MODE's masks over the owned atoms -> host (flags_h[8..8+NBITS))
template <int MODE> static int mask_count(b200md_ctx *c, const MaskGeom &mg, int n, int *counts)
{
  int nblk = cdiv(n, MASK_BLOCK);
  if (nblk < 1) nblk = 1;
  TRY(dev_ensure(c, c->border_blockcnt, (size_t) 2 * NBITS * nblk));
  TRY(dev_ensure(c, c->border_total, NBITS));
  int *cnt = c->border_blockcnt.p, *off = cnt + (size_t) NBITS * nblk;
  k_mask_count<MODE><<<nblk, MASK_BLOCK, 0, c->stream>>>(c->pos.p, n, mg, nblk, cnt);
  LAUNCH_CHECK();
  k_mask_scan<<<NBITS, 1024, 0, c->stream>>>(cnt, off, nblk, c->border_total.p);
  LAUNCH_CHECK();
  CU(cudaMemcpyAsync(c->flags_h + 8, c->border_total.p, NBITS * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  for (int d = 0; d < NBITS; d++) counts[d] = c->flags_h[8 + d];
  return B200MD_OK;
}

template <int MODE> static int mask_fill(b200md_ctx *c, const MaskGeom &mg, int n, const SendOff &so, int total)
{
  int nblk = cdiv(n, MASK_BLOCK);
  if (nblk < 1) nblk = 1;
  TRY(dev_ensure(c, c->send_idx, (size_t) total + 1));
  TRY(dev_ensure(c, c->send_dir, (size_t) total + 1));
  int *cnt = c->border_blockcnt.p, *off = cnt + (size_t) NBITS * nblk;
  k_mask_fill<MODE><<<nblk, MASK_BLOCK, 0, c->stream>>>(c->pos.p, n, mg, nblk, cnt, off, so, c->send_idx.p, c->send_dir.p);
  LAUNCH_CHECK();
  return B200MD_OK;
}

// every rank learns every rank's per-direction counts; mine[] in, theirs[d] = count rank peer_recv[d] sends in d
static int exchange_counts(b200md_ctx *c, const int *mine, int *theirs)
{
  if (c->nranks == 1) {
    for (int d = 0; d < NDIR; d++) theirs[d] = c->peer_recv[d] == c->rank ? mine[d] : 0;
    return B200MD_OK;
  }
  TRY(dev_ensure(c, c->allcounts, (size_t) 32 * (c->nranks + 1)));
  int *mine_d = c->allcounts.p, *all_d = c->allcounts.p + 32;
  CU(cudaMemcpyAsync(mine_d, mine, NDIR * sizeof(int), cudaMemcpyHostToDevice, c->stream));
  NC(g_nccl.AllGather(mine_d, all_d, 32, ncclInt, (ncclComm_t) c->nccl, c->stream));
  CU(cudaMemcpyAsync(c->flags_h + 64, all_d, (size_t) 32 * c->nranks * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  for (int d = 0; d < NDIR; d++) theirs[d] = c->peer_recv[d] >= 0 ? c->flags_h[64 + 32 * c->peer_recv[d] + d] : 0;
  return B200MD_OK;
}

// one NCCL group: direction-d segment of `sbuf` (send offsets) to peer_send[d], direction-d segment of `rbuf`
// (receive offsets) from peer_recv[d]; `bytes` per atom.  Messages between one pair of ranks are issued in the
// same (ascending d) order on both sides.
static int sendrecv_dirs(b200md_ctx *c, const char *sbuf, const int *soff, const int *scount, char *rbuf, const int *roff,
                         const int *rcount, size_t bytes)
{
  if (c->nranks == 1) return B200MD_OK;
  cudaStream_t st = c->hstream ? c->hstream : c->stream;  // overlap mode redirects the per-step halo to s_comm
  NC(g_nccl.GroupStart());
  for (int d = 0; d < NDIR; d++) {
    if (c->peer_send[d] >= 0 && c->peer_send[d] != c->rank && scount[d] > 0)
      NC(g_nccl.Send(sbuf + (size_t) soff[d] * bytes, (size_t) scount[d] * bytes, ncclChar, c->peer_send[d], (ncclComm_t) c->nccl, st));
    if (c->peer_recv[d] >= 0 && c->peer_recv[d] != c->rank && rcount[d] > 0)
      NC(g_nccl.Recv(rbuf + (size_t) roff[d] * bytes, (size_t) rcount[d] * bytes, ncclChar, c->peer_recv[d], (ncclComm_t) c->nccl, st));
  }
  NC(g_nccl.GroupEnd());
  return B200MD_OK;
}

// ------------------------------------------------------------------------------------------------
// pack kernels: gather the send list, apply the sender-side periodic shift, and either write the ghost slot
// directly (peer == self) or fill the contiguous send buffer that NCCL ships
// ------------------------------------------------------------------------------------------------
template <int WITH_TAG>
__global__ void k_halo_pack_pos(double4 *__restrict__ pos, int *__restrict__ tag, const int *__restrict__ send_idx,
                                const signed char *__restrict__ send_dir, int nsend, int nlocal, HaloPlan hp,
                                double4 *__restrict__ sendbuf, int *__restrict__ sendtag)
{
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nsend) return;
  const int d = send_dir[e], s = send_idx[e];
  double4 p = pos[s];
  p.x += hp.shift[d][0];
  p.y += hp.shift[d][1];
  p.z += hp.shift[d][2];
  if (hp.self[d]) {
    const int g = nlocal + hp.goff[d] + (e - hp.soff[d]);
    pos[g] = p;
    if (WITH_TAG) tag[g] = tag[s];
  } else {
    sendbuf[e] = p;
    if (WITH_TAG) sendtag[e] = tag[s];
  }
}

__global__ void k_halo_pack_scalar(double *__restrict__ a, const int *__restrict__ send_idx,
                                   const signed char *__restrict__ send_dir, int nsend, int nlocal, HaloPlan hp,
                                   double *__restrict__ sendbuf)
{
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nsend) return;
  const int d = send_dir[e];
  const double v = a[send_idx[e]];
  if (hp.self[d]) a[nlocal + hp.goff[d] + (e - hp.soff[d])] = v;
  else sendbuf[e] = v;
}

static bool any_remote(const b200md_ctx *c)
{
  if (c->nranks == 1) return false;
  for (int d = 0; d < NDIR; d++)
    if (c->peer_send[d] >= 0 && c->peer_send[d] != c->rank) return true;
  return false;
}

static int put_plan_epoch(b200md_ctx *c);
int halo_put_check(b200md_ctx *c);

// CommBrick::borders
int comm_borders(b200md_ctx *c)
{
  const int n = c->nlocal;
  TRY(comm_plan(c));
  MaskGeom mg = {};
  for (int k = 0; k < 3; k++) {
    mg.lo_cut[k] = c->sublo[k] + c->cutneighmax;  // cutghost = cutneighmax (src/comm.cpp:Comm::init, comm_brick.cpp:425-455)
    mg.hi_cut[k] = c->subhi[k] - c->cutneighmax;
    mg.send_lo[k] = c->periodic[k] || c->myloc[k] > 0;
    mg.send_hi[k] = c->periodic[k] || c->myloc[k] < c->procgrid[k] - 1;
  }
  int counts[NBITS];
  TRY(mask_count<0>(c, mg, n, counts));
  counts[NDIR] = 0;
  TRY(exchange_counts(c, counts, c->recvcount));
  SendOff so;
  int st = 0, gt = 0;
  for (int d = 0; d < NDIR; d++) {
    c->sendcount[d] = counts[d];
    so.off[d] = c->hp.soff[d] = st;
    c->hp.goff[d] = gt;
    st += counts[d];
    gt += c->recvcount[d];
  }
  so.off[NDIR] = so.off[NBITS] = c->hp.soff[NDIR] = st;
  c->hp.goff[NDIR] = gt;
  c->nsend = st;
  c->nghost = gt;
  const size_t nall = (size_t) n + gt;
  TRY(dev_ensure(c, c->pos, nall, true));
  TRY(dev_ensure(c, c->tag, nall, true));
  const bool remote = any_remote(c);
  if (remote) {
    TRY(dev_ensure(c, c->sendbuf4, (size_t) st + 1));
    TRY(dev_ensure(c, c->sendbuf_i, (size_t) st + 1));
  }
  if (st > 0) {
    TRY(mask_fill<0>(c, mg, n, so, st));
    KProf kp(c, KC_HALO);
    k_halo_pack_pos<1><<<cdiv(st, 256), 256, 0, c->stream>>>(c->pos.p, c->tag.p, c->send_idx.p, c->send_dir.p, st, n, c->hp,
                                                              c->sendbuf4.p, c->sendbuf_i.p);
    LAUNCH_CHECK();
  }
  if (remote) {
    TRY(sendrecv_dirs(c, (const char *) c->sendbuf4.p, c->hp.soff, c->sendcount, (char *) (c->pos.p + n), c->hp.goff,
                      c->recvcount, sizeof(double4)));
    TRY(sendrecv_dirs(c, (const char *) c->sendbuf_i.p, c->hp.soff, c->sendcount, (char *) (c->tag.p + n), c->hp.goff,
                      c->recvcount, sizeof(int)));
  }
  TRY(halo_put_check(c));
  return put_plan_epoch(c);
}

// ------------------------------------------------------------------------------------------------
// halo as ONE fused pack-and-put kernel over peer memory (NVLink P2P through cudaIpc mappings) instead of
// pack -> send buffer -> 26 ncclSend/ncclRecv: CommBrick::forward_comm (src/comm_brick.cpp:538-591) without a staging
// copy on the sender and without a communication kernel of a library.
//   sender   k_halo_pack_put: record e of direction d goes straight to landing[parity][goff_peer[d] + e - soff[d]] in the
//            receiver's memory; every block fences (system scope) and counts itself done; the last block to finish
//            stores the halo's sequence number into the receiver's flag of every direction it served
//   receiver k_halo_wait (one warp) spins on its own flags until every direction it expects carries this sequence number,
//            then k_halo_unpack copies the landed records into the ghost slots
// Two landing buffers alternate by halo parity: a rank can be at most one halo ahead of the peer that still reads the
// other buffer (to put halo k+2 it needs halo k+1 from that peer, which the peer packs after unpacking halo k).
// Rebuild-time exchanges (counts, borders with tags, migration) stay on NCCL; they also order the epochs.
// ------------------------------------------------------------------------------------------------
#define PUT_FLAG_BYTES 256
struct PutPlan {
  double4 *dst[NDIR];   // first landing record of direction d in the receiver's memory (nullptr: self or nothing to send)
  int *flag[NDIR];      // the receiver's flag of direction d for this parity
  unsigned seq;
};

__global__ void k_halo_pack_put(double4 *__restrict__ arr, const int *__restrict__ send_idx, const signed char *__restrict__ send_dir,
                                int nsend, int nlocal, HaloPlan hp, PutPlan pp, int *__restrict__ done)
{
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < nsend) {
    const int d = send_dir[e], s = send_idx[e];
    double4 p = arr[s];
    p.x += hp.shift[d][0];
    p.y += hp.shift[d][1];
    p.z += hp.shift[d][2];
    if (hp.self[d]) arr[nlocal + hp.goff[d] + (e - hp.soff[d])] = p;
    else pp.dst[d][e - hp.soff[d]] = p;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int t = atomicAdd(done, 1);
    if (t == (int) gridDim.x - 1) {   // every block's records are visible system-wide: raise the receivers' flags
      *done = 0;
      __threadfence_system();
      for (int d = 0; d < NDIR; d++)
        if (pp.flag[d]) *(volatile int *) pp.flag[d] = (int) pp.seq;
      __threadfence_system();
    }
  }
}

// lane d waits for direction d; gives up after about two seconds of GPU time (a peer that died must not hang the box)
__global__ void k_halo_wait(const int *__restrict__ flags, unsigned mask, unsigned seq, int *__restrict__ timed_out)
{
  const int d = threadIdx.x;
  if (d < NDIR && ((mask >> d) & 1u)) {
    const volatile int *f = flags + d;
    const long long t0 = clock64();
    while ((int) ((unsigned) *f - seq) < 0) {
      if (clock64() - t0 > 4000000000LL) { *timed_out = 1; break; }
    }
  }
  __threadfence_system();
}

__global__ void k_halo_unpack(double4 *__restrict__ arr, const double4 *__restrict__ landing, int nlocal, int nghost, HaloPlan hp,
                              unsigned remote_mask)
{
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nghost) return;
  int d = 0;
  while (d < NDIR - 1 && g >= hp.goff[d + 1]) d++;
  if ((remote_mask >> d) & 1u) {   // (landed through L2 by a peer: read around L1)
    const double2 *q = reinterpret_cast<const double2 *>(landing + g);
    const double2 a = __ldcg(q), b = __ldcg(q + 1);
    arr[nlocal + g] = make_double4(a.x, a.y, b.x, b.y);
  }
}

// after borders(): does every rank's ghost set fit its landing buffer, and where does each of my directions land?
static int put_plan_epoch(b200md_ctx *c)
{
  c->put.epoch_ok = false;
  if (c->nranks == 1 || !c->put.enabled || !c->put.imported) return B200MD_OK;
  TRY(dev_ensure(c, c->allcounts, (size_t) 32 * (c->nranks + 1)));
  int mine[32] = {0};
  for (int d = 0; d < NDIR; d++) mine[d] = c->hp.goff[d];
  mine[31] = (size_t) c->nghost <= c->put.cap ? 1 : 0;
  int *mine_d = c->allcounts.p, *all_d = c->allcounts.p + 32;
  CU(cudaMemcpyAsync(mine_d, mine, 32 * sizeof(int), cudaMemcpyHostToDevice, c->stream));
  NC(g_nccl.AllGather(mine_d, all_d, 32, ncclInt, (ncclComm_t) c->nccl, c->stream));
  std::vector<int> all((size_t) 32 * c->nranks);
  CU(cudaMemcpyAsync(all.data(), all_d, all.size() * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  bool ok = true;
  for (int r = 0; r < c->nranks; r++) ok = ok && all[(size_t) 32 * r + 31] == 1;
  for (int d = 0; d < NDIR; d++) {
    const int peer = c->peer_send[d];
    c->put.peer_goff[d] = (peer >= 0 && peer != c->rank) ? all[(size_t) 32 * peer + d] : 0;
    if (peer >= 0 && peer != c->rank && !c->put.peer_base[peer]) ok = false;
  }
  c->put.epoch_ok = ok;
  return B200MD_OK;
}

static int halo_forward_put(b200md_ctx *c, double4 *arr)
{
  cudaStream_t st = c->hstream ? c->hstream : c->stream;
  const unsigned seq = ++c->put.seq, par = seq & 1u;
  PutPlan pp = {};
  pp.seq = seq;
  unsigned wait_mask = 0;
  for (int d = 0; d < NDIR; d++) {
    const int ps = c->peer_send[d], pr = c->peer_recv[d];
    if (ps >= 0 && ps != c->rank && c->sendcount[d] > 0) {
      char *pb = (char *) c->put.peer_base[ps];
      pp.dst[d] = (double4 *) (pb + PUT_FLAG_BYTES) + (size_t) par * c->put.peer_cap[ps] + c->put.peer_goff[d];
      pp.flag[d] = (int *) pb + par * 32 + d;
    }
    if (pr >= 0 && pr != c->rank && c->recvcount[d] > 0) wait_mask |= 1u << d;
  }
  KProf kp(c, KC_HALO);
  const int blocks = cdiv(c->nsend > 0 ? c->nsend : 1, 256);
  k_halo_pack_put<<<blocks, 256, 0, st>>>(arr, c->send_idx.p, c->send_dir.p, c->nsend, c->nlocal, c->hp, pp, c->put.done_d);
  LAUNCH_CHECK();
  if (wait_mask) {
    k_halo_wait<<<1, 32, 0, st>>>((const int *) c->put.base + par * 32, wait_mask, seq, c->put.done_d + 1);
    LAUNCH_CHECK();
    const double4 *landing = (const double4 *) (c->put.base + PUT_FLAG_BYTES) + (size_t) par * c->put.cap;
    k_halo_unpack<<<cdiv(c->nghost, 256), 256, 0, st>>>(arr, landing, c->nlocal, c->nghost, c->hp, wait_mask);
    LAUNCH_CHECK();
  }
  return B200MD_OK;
}

// a wait of the put halo gave up (read where the step synchronises anyway: at every rebuild)
int halo_put_check(b200md_ctx *c)
{
  if (!c->put.done_d || !c->put.epoch_ok) return B200MD_OK;
  int t = 0;
  CU(cudaMemcpyAsync(&t, c->put.done_d + 1, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  if (t) return b200md_fail(c, B200MD_ERR_CUDA, "halo put: a peer's records did not arrive within the wait limit");
  return B200MD_OK;
}

// CommBrick::forward_comm: ghost records {x,y,z,w} from their owners; the sender adds the periodic shift to x,y,z
// and passes w through (type for the position array, fp for the EAM {x,y,z,fp} array)
int halo_forward_vec4(b200md_ctx *c, double4 *arr)
{
  if (c->put.epoch_ok) return halo_forward_put(c, arr);
  if (c->nsend > 0) {
    KProf kp(c, KC_HALO);
    k_halo_pack_pos<0><<<cdiv(c->nsend, 256), 256, 0, c->hstream ? c->hstream : c->stream>>>(arr, c->tag.p, c->send_idx.p, c->send_dir.p, c->nsend,
                                                                    c->nlocal, c->hp, c->sendbuf4.p, c->sendbuf_i.p);
    LAUNCH_CHECK();
  }
  if (any_remote(c)) {
    KProf kp(c, KC_HALO);
    TRY(sendrecv_dirs(c, (const char *) c->sendbuf4.p, c->hp.soff, c->sendcount, (char *) (arr + c->nlocal), c->hp.goff,
                      c->recvcount, sizeof(double4)));
  }
  return B200MD_OK;
}

int halo_forward_pos(b200md_ctx *c) { return halo_forward_vec4(c, c->pos.p); }

// Comm::forward_comm(Pair*) with comm_forward = 1 (EAM fp): a[nlocal + g] from the owner's a[]
int halo_forward_scalar(b200md_ctx *c, double *a)
{
  const bool remote = any_remote(c);
  if (remote) TRY(dev_ensure(c, c->sendbuf1, (size_t) c->nsend + 1));
  if (c->nsend > 0) {
    KProf kp(c, KC_HALO);
    k_halo_pack_scalar<<<cdiv(c->nsend, 256), 256, 0, c->stream>>>(a, c->send_idx.p, c->send_dir.p, c->nsend, c->nlocal, c->hp,
                                                                   c->sendbuf1.p);
    LAUNCH_CHECK();
  }
  if (remote) {
    KProf kp(c, KC_HALO);
    TRY(sendrecv_dirs(c, (const char *) c->sendbuf1.p, c->hp.soff, c->sendcount, (char *) (a + c->nlocal), c->hp.goff,
                      c->recvcount, sizeof(double)));
  }
  return B200MD_OK;
}

// ------------------------------------------------------------------------------------------------
// exchange: owned atoms that left the sub-domain move to the rank that now owns them (one hop per dimension,
// like the reference); everything an atom carries travels as two double4 records:
//   {x, y, z, type}  and  {vx, vy, vz, (tag, image)}     (AtomVecAtomic::pack_exchange, src/atom_vec.cpp:1153-1200)
// ------------------------------------------------------------------------------------------------
__global__ void k_exch_pack(const double4 *__restrict__ pos, const double *__restrict__ vx, const double *__restrict__ vy,
                            const double *__restrict__ vz, const int *__restrict__ tag, const int *__restrict__ image,
                            const int *__restrict__ idx, int nmove, int nstay, const int *__restrict__ stay_idx,
                            double4 *__restrict__ sbuf_x, double4 *__restrict__ sbuf_v, double4 *__restrict__ pos_o,
                            double *__restrict__ vx_o, double *__restrict__ vy_o, double *__restrict__ vz_o,
                            int *__restrict__ tag_o, int *__restrict__ image_o)
{
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < nmove) {
    const int i = idx[e];
    sbuf_x[e] = pos[i];
    sbuf_v[e] = make_double4(vx[i], vy[i], vz[i], __hiloint2double(tag[i], image[i]));
  } else if (e < nmove + nstay) {
    const int k = e - nmove, i = stay_idx[k];
    pos_o[k] = pos[i];
    vx_o[k] = vx[i];
    vy_o[k] = vy[i];
    vz_o[k] = vz[i];
    tag_o[k] = tag[i];
    image_o[k] = image[i];
  }
}

__global__ void k_exch_unpack(const double4 *__restrict__ rbuf_v, int nrecv, int base, double *__restrict__ vx,
                              double *__restrict__ vy, double *__restrict__ vz, int *__restrict__ tag, int *__restrict__ image)
{
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nrecv) return;
  const double4 r = rbuf_v[e];
  vx[base + e] = r.x;
  vy[base + e] = r.y;
  vz[base + e] = r.z;
  tag[base + e] = __double2hiint(r.w);
  image[base + e] = __double2loint(r.w);
}

int comm_exchange(b200md_ctx *c)
{
  if (c->nranks == 1) return B200MD_OK;  // Domain::pbc already put every atom back into the only sub-domain
  TRY(comm_plan(c));
  const int n = c->nlocal;
  MaskGeom mg = {};
  for (int k = 0; k < 3; k++) {
    const bool split = c->procgrid[k] > 1;
    mg.lo_cut[k] = split ? c->sublo[k] : -1e300;
    mg.hi_cut[k] = split ? c->subhi[k] : 1e300;
    mg.wrap_to_lo[k] = 1e300;
    mg.wrap_to_hi[k] = -1e300;
    if (c->procgrid[k] > 2 && c->periodic[k]) {
      const double w = c->prd[k] / c->procgrid[k];
      if (c->myloc[k] == 0) mg.wrap_to_lo[k] = c->boxhi[k] - w;
      if (c->myloc[k] == c->procgrid[k] - 1) mg.wrap_to_hi[k] = c->boxlo[k] + w;
    }
  }
  int counts[NBITS], rcount[NDIR];
  TRY(mask_count<1>(c, mg, n, counts));
  TRY(exchange_counts(c, counts, rcount));
  SendOff so;
  int soff[NDIR + 1], roff[NDIR + 1], st = 0, rt = 0;
  for (int d = 0; d < NDIR; d++) {
    if (counts[d] > 0 && c->peer_send[d] < 0)
      return b200md_fail(c, B200MD_ERR_LOST_ATOMS, "Lost atoms: %d left the non-periodic box through direction %d", counts[d], d);
    so.off[d] = soff[d] = st;
    roff[d] = rt;
    st += counts[d];
    rt += rcount[d];
  }
  soff[NDIR] = st;
  roff[NDIR] = rt;
  so.off[NDIR] = st;  // the "stay" list follows the movers
  const int nstay = counts[NDIR];
  so.off[NBITS] = st + nstay;
  int anymove[2] = {st, rt};
  // nothing moves anywhere on this rank and nothing arrives: keep the arrays as they are
  if (anymove[0] == 0 && anymove[1] == 0) return B200MD_OK;
  const int nnew = nstay + rt;
  TRY(mask_fill<1>(c, mg, n, so, st + nstay));
  TRY(dev_ensure(c, c->sendbuf4, (size_t) 2 * st + 2));
  TRY(dev_ensure(c, c->recvbuf4, (size_t) rt + 1));
  const size_t cap = (size_t) nnew + nnew / 16 + 1024;
  if (c->pos_tmp.n < cap) TRY(dev_ensure(c, c->pos_tmp, cap));
  if (c->tag_tmp.n < cap) TRY(dev_ensure(c, c->tag_tmp, cap));
  if (c->image_tmp.n < (size_t) nnew) TRY(dev_ensure(c, c->image_tmp, cap));
  for (int k = 0; k < 3; k++)
    if (c->v_tmp[k].n < (size_t) nnew) TRY(dev_ensure(c, c->v_tmp[k], cap));
  double4 *sx = c->sendbuf4.p, *sv = c->sendbuf4.p + st + 1;
  k_exch_pack<<<cdiv(st + nstay, 256), 256, 0, c->stream>>>(c->pos.p, c->v[0].p, c->v[1].p, c->v[2].p, c->tag.p, c->image.p,
                                                             c->send_idx.p, st, nstay, c->send_idx.p + st, sx, sv, c->pos_tmp.p,
                                                             c->v_tmp[0].p, c->v_tmp[1].p, c->v_tmp[2].p, c->tag_tmp.p,
                                                             c->image_tmp.p);
  LAUNCH_CHECK();
  TRY(sendrecv_dirs(c, (const char *) sx, soff, counts, (char *) (c->pos_tmp.p + nstay), roff, rcount, sizeof(double4)));
  TRY(sendrecv_dirs(c, (const char *) sv, soff, counts, (char *) c->recvbuf4.p, roff, rcount, sizeof(double4)));
  if (rt > 0) {
    k_exch_unpack<<<cdiv(rt, 256), 256, 0, c->stream>>>(c->recvbuf4.p, rt, nstay, c->v_tmp[0].p, c->v_tmp[1].p, c->v_tmp[2].p,
                                                        c->tag_tmp.p, c->image_tmp.p);
    LAUNCH_CHECK();
  }
  std::swap(c->pos, c->pos_tmp);
  std::swap(c->tag, c->tag_tmp);
  std::swap(c->image, c->image_tmp);
  for (int k = 0; k < 3; k++) std::swap(c->v[k], c->v_tmp[k]);
  c->nlocal = nnew;
  c->nmigrated += st;
  return B200MD_OK;
}

// ---- halo put: the landing buffer of this rank and the mappings of the peers' -------------------------------------
extern "C" int b200md_halo_put_export(b200md_ctx *c, int64_t cap_records, void *handle_out)
{
  if (!c) return b200md_fail(nullptr, B200MD_ERR_ARG, "NULL context");
  if (!handle_out || cap_records <= 0) return b200md_fail(c, B200MD_ERR_ARG, "halo_put_export: cap_records > 0 and handle_out[64] are required");
  CU(cudaSetDevice(c->device));
  if (c->put.base) return b200md_fail(c, B200MD_ERR_ARG, "halo_put_export: already exported");
  const size_t bytes = PUT_FLAG_BYTES + (size_t) 2 * (size_t) cap_records * sizeof(double4);
  CU(cudaMalloc((void **) &c->put.base, bytes));
  CU(cudaMemset(c->put.base, 0, PUT_FLAG_BYTES));
  CU(cudaMalloc((void **) &c->put.done_d, 2 * sizeof(int)));
  CU(cudaMemset(c->put.done_d, 0, 2 * sizeof(int)));
  c->put.cap = (size_t) cap_records;
  c->bytes_held += (double) bytes;
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, c->put.base));
  static_assert(sizeof(h) == B200MD_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t size");
  memcpy(handle_out, &h, sizeof(h));
  return B200MD_OK;
}

extern "C" int b200md_halo_put_import(b200md_ctx *c, int nranks, const void *handles, const int64_t *caps)
{
  if (!c) return b200md_fail(nullptr, B200MD_ERR_ARG, "NULL context");
  if (nranks != c->nranks || nranks > 64 || !handles || !caps) return b200md_fail(c, B200MD_ERR_ARG, "halo_put_import: handles[nranks][64] and caps[nranks] of this communicator's ranks (at most 64)");
  if (!c->put.base) return b200md_fail(c, B200MD_ERR_ARG, "halo_put_import before halo_put_export");
  CU(cudaSetDevice(c->device));
  for (int r = 0; r < nranks; r++) {   // every other rank of the box (at most 7): any of them may become a neighbor
    c->put.peer_cap[r] = (size_t) caps[r];
    if (r == c->rank || c->put.peer_base[r]) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char *) handles + (size_t) r * B200MD_IPC_HANDLE_BYTES, sizeof(h));
    CU(cudaIpcOpenMemHandle(&c->put.peer_base[r], h, cudaIpcMemLazyEnablePeerAccess));
  }
  c->put.imported = true;
  return B200MD_OK;
}

// halos that went through the put path so far (0: every halo used NCCL)
extern "C" int64_t b200md_halo_put_count(const b200md_ctx *c) { return c ? (int64_t) c->put.seq : 0; }
