// pair_lj.cu -- pair_style lj/cut on the device: PairLJCut::compute (src/pair_lj_cut.cpp:69-139).
// Full neighbor list, one thread per owned atom, no atomics and no ghost-force reverse_comm; the
// force of atom i is written once (this also replaces Verlet::force_clear, src/verlet.cpp:376-421).
// Energy / virial (EVFLAG) follow Pair::ev_tally_full (src/pair.cpp:1115-1160): half of each pair term
// per visit, reduced per block and then in a fixed order (reproducible run to run).
#include <stdlib.h>

#include "ctx.cuh"

struct LJOne { double cutsq, lj1, lj2, lj3, lj4, offset; };

__device__ __forceinline__ double4 ld_pos(const double4 *__restrict__ pos, int j)
{
  return ld_pos256(pos + j);
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}

// block-level reduction of NC components into partials[blockIdx.x*8 + c]; blockDim.x multiple of 32
template <int NC> __device__ __forceinline__ void block_partials(double (&acc)[NC], double *__restrict__ partials)
{
  __shared__ double s_red[32][NC];
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int c = 0; c < NC; c++) {
    double v = warp_sum(acc[c]);
    if (lane == 0) s_red[w][c] = v;
  }
  __syncthreads();
  if (threadIdx.x < NC) {
    double t = 0.0;
    for (int k = 0; k < nw; k++) t += s_red[k][threadIdx.x];
    partials[(size_t) blockIdx.x * 8 + threadIdx.x] = t;
  }
}

// One thread per cluster of CL consecutive owned atoms (neigh.cu): x_j is gathered once per list entry (one
// 256-bit load) and tested against all CL atoms of the cluster.  A pair contributes iff its mask bit is set (it is
// in the reference's list, decided at build time) and `rsq < cutsq` now -- exactly the reference's two tests.
template <int CL, int EVFLAG, int ONETYPE>
__global__ void __launch_bounds__(128)
k_lj_cluster(const double4 *__restrict__ pos, const int *__restrict__ numneigh, const int *__restrict__ neigh, int maxneigh,
             int nlocal, LJOne one, const double *__restrict__ tables, int ntypes, double *__restrict__ fx,
             double *__restrict__ fy, double *__restrict__ fz, double *__restrict__ partials)
{
  extern __shared__ double s_tab[];  // 6 tables of (ntypes+1)^2 when !ONETYPE
  const int nt1 = ntypes + 1, nt2 = nt1 * nt1;
  if (!ONETYPE) {
    for (int k = threadIdx.x; k < 6 * nt2; k += blockDim.x) s_tab[k] = tables[k];
    __syncthreads();
  }
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = q * CL;
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  if (i0 < nlocal) {
    double xi[CL], yi[CL], zi[CL], fxi[CL], fyi[CL], fzi[CL];
    int trow[CL];
#pragma unroll
    for (int a = 0; a < CL; a++) {
      fxi[a] = fyi[a] = fzi[a] = 0.0;
      if (i0 + a < nlocal) {
        const double4 p = pos[i0 + a];
        xi[a] = p.x; yi[a] = p.y; zi[a] = p.z;
        trow[a] = ONETYPE ? 0 : pos_type(p) * nt1;
      } else {
        xi[a] = yi[a] = zi[a] = 1e150;
        trow[a] = ONETYPE ? 0 : nt1;
      }
    }
    const int n = numneigh[q];
    const int *nb = neigh + (size_t) (q >> 5) * TILE * maxneigh + (q & 31);
#pragma unroll 2
    for (int k = 0; k < n; k++) {
      const unsigned ent = (unsigned) __ldcs(nb + (size_t) k * TILE);
      const int j = (int) (ent & CLJMASK);
      const double4 pj = ld_pos256(pos + j);
      const int jt = ONETYPE ? 0 : pos_type(pj);
#pragma unroll
      for (int a = 0; a < CL; a++) {
        const double dx = xi[a] - pj.x, dy = yi[a] - pj.y, dz = zi[a] - pj.z;
        const double rsq = dx * dx + dy * dy + dz * dz;
        double cutsq, lj1, lj2;
        if (ONETYPE) {
          cutsq = one.cutsq; lj1 = one.lj1; lj2 = one.lj2;
        } else {
          const int t = trow[a] + jt;
          cutsq = s_tab[t]; lj1 = s_tab[nt2 + t]; lj2 = s_tab[2 * nt2 + t];
        }
        // branch-free: the CL chains are independent and interleave (a divergent body per atom would serialise
        // four long FP64 dependency chains); out-of-range pairs are computed and discarded by the select
        const bool in = rsq < cutsq && ((ent >> (CLBITS + a)) & 1u);
        const double r2inv = rcp_pos(rsq);
        const double r6inv = r2inv * r2inv * r2inv;
        const double forcelj = r6inv * (lj1 * r6inv - lj2);
        const double fpair = in ? forcelj * r2inv : 0.0;
        fxi[a] += dx * fpair;
        fyi[a] += dy * fpair;
        fzi[a] += dz * fpair;
        if (EVFLAG) {
          double lj3, lj4, off;
          if (ONETYPE) {
            lj3 = one.lj3; lj4 = one.lj4; off = one.offset;
          } else {
            const int t = trow[a] + jt;
            lj3 = s_tab[3 * nt2 + t]; lj4 = s_tab[4 * nt2 + t]; off = s_tab[5 * nt2 + t];
          }
          acc[0] += in ? 0.5 * (r6inv * (lj3 * r6inv - lj4) - off) : 0.0;
          const double hf = 0.5 * fpair;
          acc[1] += dx * dx * hf;
          acc[2] += dy * dy * hf;
          acc[3] += dz * dz * hf;
          acc[4] += dx * dy * hf;
          acc[5] += dx * dz * hf;
          acc[6] += dy * dz * hf;
        }
      }
    }
#pragma unroll
    for (int a = 0; a < CL; a++)
      if (i0 + a < nlocal) {
        fx[i0 + a] = fxi[a];
        fy[i0 + a] = fyi[a];
        fz[i0 + a] = fzi[a];
      }
  }
  if (EVFLAG) block_partials<7>(acc, partials);
}

// Lane-cluster form: CL consecutive LANES (= CL consecutive, spatially compact atoms) share one list, so a warp
// step gathers 32/CL distinct x_j -- the CL lanes of a cluster read the same sector, which the L1 data pipe serves
// as one wavefront.  ncu on the per-atom kernel: LSU data pipe at 88 % of peak with 26 wavefronts per warp step
// (one per lane: 32 distinct sectors), FP64 pipe at 36 %: trading 1/efficiency more masked-out FP64 work for
// 1/CL of the gather wavefronts moves the kernel towards the FP64 pipe.
// SPLIT (overlap mode): only the warps (= 32-atom tiles) whose ghost flag equals `part` work, so the ghost-free
// tiles can run while the halo is still in flight
template <int CL, int EVFLAG, int ONETYPE, int SPLIT>
__global__ void __launch_bounds__(128)
k_lj_lanes(const double4 *__restrict__ pos, const int *__restrict__ numneigh, const int *__restrict__ neigh, int maxneigh,
           int nlocal, LJOne one, const double *__restrict__ tables, int ntypes, double *__restrict__ fx,
           double *__restrict__ fy, double *__restrict__ fz, double *__restrict__ partials,
           const int *__restrict__ tile_ghost, int part)
{
  extern __shared__ double s_tab[];
  const int nt1 = ntypes + 1, nt2 = nt1 * nt1;
  if (!ONETYPE) {
    for (int k = threadIdx.x; k < 6 * nt2; k += blockDim.x) s_tab[k] = tables[k];
    __syncthreads();
  }
  // SPLIT == 2 (host-buffer pipeline): the atom range [part, nlocal), part a multiple of the block size
  const int i = (SPLIT == 2 ? part : 0) + blockIdx.x * blockDim.x + threadIdx.x;
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  bool active = i < nlocal;
  if (SPLIT == 1 && active) active = (tile_ghost[i >> 5] != 0) == (part != 0);
  if (active) {
    const int q = i / CL, a = i % CL;
    const double4 pi = pos[i];
    const int trow = ONETYPE ? 0 : pos_type(pi) * nt1;
    const int n = numneigh[q];
    const int *nb = neigh + (size_t) (q >> 5) * TILE * maxneigh + (q & 31);
    double fxi = 0.0, fyi = 0.0, fzi = 0.0;
    // software pipeline, depth PF: the list entries of the NEXT batch are in flight while the PF gathers of this
    // batch are issued back to back and then consumed (the per-step chain index -> gather -> FP64 otherwise
    // serialises two DRAM/L2 latencies per neighbor)
    constexpr int PF = 4;
    const int nlast = max(n - 1, 0);
    unsigned enx[PF];
#pragma unroll
    for (int u = 0; u < PF; u++) enx[u] = (unsigned) __ldcs(nb + (size_t) min(u, nlast) * TILE);
    for (int k0 = 0; k0 < n; k0 += PF) {
      unsigned ent[PF];
      double4 pj[PF];
#pragma unroll
      for (int u = 0; u < PF; u++) {
        ent[u] = enx[u];
        pj[u] = ld_pos256(pos + (ent[u] & CLJMASK));
      }
#pragma unroll
      for (int u = 0; u < PF; u++) enx[u] = (unsigned) __ldcs(nb + (size_t) min(k0 + PF + u, nlast) * TILE);
#pragma unroll
      for (int u = 0; u < PF; u++) {
        const double dx = pi.x - pj[u].x, dy = pi.y - pj[u].y, dz = pi.z - pj[u].z;
        const double rsq = dx * dx + dy * dy + dz * dz;
        double cutsq, lj1, lj2;
        if (ONETYPE) {
          cutsq = one.cutsq; lj1 = one.lj1; lj2 = one.lj2;
        } else {
          const int t = trow + pos_type(pj[u]);
          cutsq = s_tab[t]; lj1 = s_tab[nt2 + t]; lj2 = s_tab[2 * nt2 + t];
        }
        const bool in = rsq < cutsq && ((ent[u] >> (CLBITS + a)) & 1u) && (k0 + u < n);
        const double r2inv = rcp_pos(rsq);
        const double r6inv = r2inv * r2inv * r2inv;
        const double forcelj = r6inv * (lj1 * r6inv - lj2);
        const double fpair = in ? forcelj * r2inv : 0.0;
        fxi += dx * fpair;
        fyi += dy * fpair;
        fzi += dz * fpair;
        if (EVFLAG) {
          double lj3, lj4, off;
          if (ONETYPE) {
            lj3 = one.lj3; lj4 = one.lj4; off = one.offset;
          } else {
            const int t = trow + pos_type(pj[u]);
            lj3 = s_tab[3 * nt2 + t]; lj4 = s_tab[4 * nt2 + t]; off = s_tab[5 * nt2 + t];
          }
          acc[0] += in ? 0.5 * (r6inv * (lj3 * r6inv - lj4) - off) : 0.0;
          const double hf = 0.5 * fpair;
          acc[1] += dx * dx * hf;
          acc[2] += dy * dy * hf;
          acc[3] += dz * dz * hf;
          acc[4] += dx * dy * hf;
          acc[5] += dx * dz * hf;
          acc[6] += dy * dz * hf;
        }
      }
    }
    fx[i] = fxi;
    fy[i] = fyi;
    fz[i] = fzi;
  }
  if (EVFLAG) block_partials<7>(acc, partials);
}

template <int CL> static void launch_lj_lanes(b200md_ctx *c, int blocks, int threads, size_t smem, int ev, int onetype, const LJOne &one, int n)
{
#define LJ_ARGS c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, one, c->lj.d, c->lj.ntypes, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p, nullptr, -1
  if (ev) {
    if (onetype) k_lj_lanes<CL, 1, 1, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    else k_lj_lanes<CL, 1, 0, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
  } else {
    if (onetype) k_lj_lanes<CL, 0, 1, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    else k_lj_lanes<CL, 0, 0, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
  }
#undef LJ_ARGS
}

// fixed-order reduction of per-block partials (ncomp <= 8 components) into red[dst + c]
__global__ void k_reduce_partials(const double *__restrict__ partials, int nblocks, int ncomp, double *__restrict__ red, int dst)
{
  __shared__ double s[1024];
  for (int comp = 0; comp < ncomp; comp++) {
    double t = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x) t += partials[(size_t) b * 8 + comp];
    s[threadIdx.x] = t;
    __syncthreads();
    for (int o = blockDim.x >> 1; o > 0; o >>= 1) {
      if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
      __syncthreads();
    }
    if (threadIdx.x == 0) red[dst + comp] = s[0];
    __syncthreads();
  }
}

int reduce_partials(b200md_ctx *c, int nblocks, int ncomp, int dst)
{
  k_reduce_partials<<<1, 1024, 0, c->stream>>>(c->partials.p, nblocks, ncomp, c->red_d, dst);
  LAUNCH_CHECK();
  return B200MD_OK;
}

template <int CL> static void launch_lj(b200md_ctx *c, int blocks, int threads, size_t smem, int ev, int onetype, const LJOne &one, int n)
{
#define LJ_ARGS c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, one, c->lj.d, c->lj.ntypes, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p
  if (ev) {
    if (onetype) k_lj_cluster<CL, 1, 1><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    else k_lj_cluster<CL, 1, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
  } else {
    if (onetype) k_lj_cluster<CL, 0, 1><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    else k_lj_cluster<CL, 0, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
  }
#undef LJ_ARGS
}

static int pair_lj_compute_tiled(b200md_ctx *c, int ev);

int pair_lj_compute(b200md_ctx *c, int eflag, int vflag)
{
  if (c->tiled_valid) return pair_lj_compute_tiled(c, (eflag || vflag) ? 1 : 0);
  const int n = c->nlocal;
  const bool lanes = c->cl == 1 ? getenv("B200MD_LJ_PLAIN") == nullptr : getenv("B200MD_LANES") != nullptr;
  const int nclus = lanes ? n : cdiv(n, c->cl);
  const int threads = 128, blocks = cdiv(nclus, threads);
  const int ev = (eflag || vflag) ? 1 : 0;
  if (ev) TRY(dev_ensure(c, c->partials, (size_t) blocks * 8));
  LJOne one = {c->lj.h_onetype[0], c->lj.h_onetype[1], c->lj.h_onetype[2], c->lj.h_onetype[3], c->lj.h_onetype[4], c->lj.h_onetype[5]};
  const int onetype = (c->lj.ntypes == 1);
  size_t smem = onetype ? 0 : (size_t) 6 * (c->lj.ntypes + 1) * (c->lj.ntypes + 1) * sizeof(double);
  {
    KProf kp(c, KC_PAIR);
    if (lanes && c->cl == 1) launch_lj_lanes<1>(c, blocks, threads, smem, ev, onetype, one, n);
    else if (lanes && c->cl == 2) launch_lj_lanes<2>(c, blocks, threads, smem, ev, onetype, one, n);
    else if (lanes) launch_lj_lanes<4>(c, blocks, threads, smem, ev, onetype, one, n);
    else if (c->cl == 1) launch_lj<1>(c, blocks, threads, smem, ev, onetype, one, n);
    else if (c->cl == 2) launch_lj<2>(c, blocks, threads, smem, ev, onetype, one, n);
    else launch_lj<4>(c, blocks, threads, smem, ev, onetype, one, n);
    LAUNCH_CHECK();
  }
  if (ev) TRY(reduce_partials(c, blocks, 7, 0));
  return B200MD_OK;
}

// overlap mode (capi.cu): forces of the tiles whose ghost flag equals `part`, no energy/virial tally
int pair_lj_compute_part(b200md_ctx *c, int part)
{
  const int n = c->nlocal;
  const int threads = 128, blocks = cdiv(n, threads);
  LJOne one = {c->lj.h_onetype[0], c->lj.h_onetype[1], c->lj.h_onetype[2], c->lj.h_onetype[3], c->lj.h_onetype[4], c->lj.h_onetype[5]};
  const int onetype = (c->lj.ntypes == 1);
  size_t smem = onetype ? 0 : (size_t) 6 * (c->lj.ntypes + 1) * (c->lj.ntypes + 1) * sizeof(double);
  KProf kp(c, KC_PAIR);
#define LJ_ARGS c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, one, c->lj.d, c->lj.ntypes, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p, c->tile_ghost.p, part
  if (onetype) k_lj_lanes<1, 0, 1, 1><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
  else k_lj_lanes<1, 0, 0, 1><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
#undef LJ_ARGS
  LAUNCH_CHECK();
  return B200MD_OK;
}

// host-buffer pipeline (capi.cu): forces of the owned atoms [first, first + count) only, no energy/virial tally, so
// the download of one range's results runs under the force kernel of the next
bool pair_lj_range_usable(const b200md_ctx *c) { return c->cl == 1 && !c->tiled_valid && getenv("B200MD_LJ_PLAIN") == nullptr; }

int pair_lj_compute_range(b200md_ctx *c, int first, int count)
{
  const int threads = 128, blocks = cdiv(count, threads);
  if (first % threads) return b200md_fail(c, B200MD_ERR_ARG, "pair range must start at a multiple of %d", threads);
  LJOne one = {c->lj.h_onetype[0], c->lj.h_onetype[1], c->lj.h_onetype[2], c->lj.h_onetype[3], c->lj.h_onetype[4], c->lj.h_onetype[5]};
  const int onetype = (c->lj.ntypes == 1);
  size_t smem = onetype ? 0 : (size_t) 6 * (c->lj.ntypes + 1) * (c->lj.ntypes + 1) * sizeof(double);
  KProf kp(c, KC_PAIR);
#define LJ_ARGS c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, first + count, one, c->lj.d, c->lj.ntypes, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p, nullptr, first
  if (onetype) k_lj_lanes<1, 0, 1, 2><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
  else k_lj_lanes<1, 0, 0, 2><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
#undef LJ_ARGS
  LAUNCH_CHECK();
  return B200MD_OK;
}

// ---- per-atom tallies (eflag_atom / vflag_atom: Pair::ev_tally, src/pair.cpp:1037-1097) ----------------------
// A pass of its own over the same list, so the timestep kernels above carry no extra stores or registers: per-atom
// energy and virial are wanted on dump / thermo steps only (compute pe/atom, compute stress/atom).  With the full
// list every atom takes half of each of its pairs -- exactly what the reference's half list gives both partners.
// out = [7][stride]: energy, then the virial components xx yy zz xy xz yz.
__global__ void __launch_bounds__(128)
k_lj_peratom(const double4 *__restrict__ pos, const int *__restrict__ numneigh, const int *__restrict__ neigh, int maxneigh,
             int nlocal, const double *__restrict__ tables, int ntypes, double *__restrict__ out, size_t stride)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const int nt1 = ntypes + 1, nt2 = nt1 * nt1;
  const double4 pi = pos[i];
  const int trow = pos_type(pi) * nt1;
  const int n = numneigh[i];
  const int *nb = neigh + (size_t) (i >> 5) * TILE * maxneigh + (i & 31);
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  for (int k = 0; k < n; k++) {
    const int j = __ldg(nb + (size_t) k * TILE) & CLJMASK;
    const double4 pj = pos[j];
    const double dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
    const double rsq = dx * dx + dy * dy + dz * dz;
    const int t = trow + pos_type(pj);
    if (rsq >= __ldg(tables + t)) continue;
    const double r2inv = 1.0 / rsq, r6inv = r2inv * r2inv * r2inv;
    const double fpair = r6inv * (__ldg(tables + nt2 + t) * r6inv - __ldg(tables + 2 * nt2 + t)) * r2inv;
    acc[0] += 0.5 * (r6inv * (__ldg(tables + 3 * nt2 + t) * r6inv - __ldg(tables + 4 * nt2 + t)) - __ldg(tables + 5 * nt2 + t));
    const double hf = 0.5 * fpair;
    acc[1] += dx * dx * hf;
    acc[2] += dy * dy * hf;
    acc[3] += dz * dz * hf;
    acc[4] += dx * dy * hf;
    acc[5] += dx * dz * hf;
    acc[6] += dy * dz * hf;
  }
#pragma unroll
  for (int c = 0; c < 7; c++) out[(size_t) c * stride + i] = acc[c];
}

int pair_lj_peratom(b200md_ctx *c, double *out, size_t stride)
{
  const int n = c->nlocal;
  k_lj_peratom<<<cdiv(n, 128), 128, 0, c->stream>>>(c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, c->lj.d, c->lj.ntypes, out, stride);
  LAUNCH_CHECK();
  return B200MD_OK;
}

// ------------------------------------------------------------------------------------------------------------
// Shared-memory tiled lj/cut (opt-in: B200MD_LJ_TILED=1; written after round 1's GPU budget was spent, validated under
// the CPU emulation only).  ncu on k_lj_lanes: the L1 LSU data pipe is the limiter -- every gathered neighbor is a
// distinct 32-byte sector, 26 wavefronts per warp step -- while DRAM sits at 19 % and the FP64 pipe at 36 %.
// Here a block owns TG consecutive bins of one x-row.  Bins are >= the neighbor cutoff wide, so every neighbor of its
// atoms lies in the 3 x 3 x (TG + 2) bins around it, and those are 9 CONTIGUOUS ranges of cell_list (bins of an
// x-row are consecutive).  The block gathers that neighbourhood once per step into shared memory (each atom once
// per block instead of once per pair: about 6 x fewer global gathers for lj at rc 2.5) and the inner loop reads
// neighbors from shared memory through 16-bit slot numbers (which also halves the list traffic, the largest DRAM
// term of the step).  List order, arithmetic and summation order are those of k_lj_lanes: forces are bit-identical.
// ------------------------------------------------------------------------------------------------------------
template <int EVFLAG, int ONETYPE>
__global__ void __launch_bounds__(128)
k_lj_tiled(TileGeom T, const int *__restrict__ bin_start, const int *__restrict__ cell_list, const double4 *__restrict__ pos,
           const int *__restrict__ numneigh, const unsigned short *__restrict__ neigh16, int maxneigh, int nlocal, int max_staged,
           LJOne one, const double *__restrict__ tables, int ntypes, double *__restrict__ fx, double *__restrict__ fy,
           double *__restrict__ fz, double *__restrict__ partials)
{
  // staged atoms as four arrays x[], y[], z[], w[] of max_staged doubles (the lists are sorted by slot, so the lanes
  // of a warp read NEARBY slots at the same step: 8-byte elements then fall into different banks, where 32-byte
  // records would use only a quarter of them), then 6 tables of (ntypes+1)^2 when !ONETYPE
  extern __shared__ double4 s_tile[];
  double *s_x = reinterpret_cast<double *>(s_tile), *s_y = s_x + max_staged, *s_z = s_y + max_staged, *s_w = s_z + max_staged;
  double *s_tab = s_w + max_staged;
  const int nt1 = ntypes + 1, nt2 = nt1 * nt1;
  int ks[9], off[10], own_s, own_e, by, bz;
  tile_rows(T, bin_start, blockIdx.x, ks, off, own_s, own_e, by, bz);
  if (!ONETYPE)
    for (int k = threadIdx.x; k < 6 * nt2; k += blockDim.x) s_tab[k] = tables[k];
#pragma unroll
  for (int r = 0; r < 9; r++) {
    const int len = off[r + 1] - off[r];
    for (int k = threadIdx.x; k < len; k += blockDim.x) {
      const double4 q = ld_pos256(pos + cell_list[ks[r] + k]);
      s_x[off[r] + k] = q.x; s_y[off[r] + k] = q.y; s_z[off[r] + k] = q.z;
      if (!ONETYPE) s_w[off[r] + k] = q.w;
    }
  }
  __syncthreads();
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  for (int slot = own_s + threadIdx.x; slot < own_e; slot += blockDim.x) {
    const int i = cell_list[slot];
    if (i >= nlocal) continue;
    const double4 pi = pos[i];
    const int trow = ONETYPE ? 0 : pos_type(pi) * nt1;
    const int n = numneigh[i];
    const unsigned short *nb = neigh16 + (size_t) (i >> 5) * TILE * maxneigh + (i & 31);
    double fxi = 0.0, fyi = 0.0, fzi = 0.0;
    // the list entries of the NEXT batch are in flight (global memory) while this batch is computed from shared memory
    constexpr int PF = 4;
    const int nlast = max(n - 1, 0);
    unsigned short enx[PF];
#pragma unroll
    for (int u = 0; u < PF; u++) enx[u] = __ldcs(nb + (size_t) min(u, nlast) * TILE);
    for (int k0 = 0; k0 < n; k0 += PF) {
      unsigned short ent[PF];
#pragma unroll
      for (int u = 0; u < PF; u++) ent[u] = enx[u];
#pragma unroll
      for (int u = 0; u < PF; u++) enx[u] = __ldcs(nb + (size_t) min(k0 + PF + u, nlast) * TILE);
#pragma unroll
      for (int u = 0; u < PF; u++) {
      double4 pj;
      pj.x = s_x[ent[u]]; pj.y = s_y[ent[u]]; pj.z = s_z[ent[u]];
      pj.w = ONETYPE ? 0.0 : s_w[ent[u]];
      const double dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
      const double rsq = dx * dx + dy * dy + dz * dz;
      double cutsq, lj1, lj2;
      if (ONETYPE) {
        cutsq = one.cutsq; lj1 = one.lj1; lj2 = one.lj2;
      } else {
        const int t = trow + pos_type(pj);
        cutsq = s_tab[t]; lj1 = s_tab[nt2 + t]; lj2 = s_tab[2 * nt2 + t];
      }
      const bool in = rsq < cutsq && (k0 + u < n);
      const double r2inv = rcp_pos(rsq);
      const double r6inv = r2inv * r2inv * r2inv;
      const double forcelj = r6inv * (lj1 * r6inv - lj2);
      const double fpair = in ? forcelj * r2inv : 0.0;
      fxi += dx * fpair;
      fyi += dy * fpair;
      fzi += dz * fpair;
      if (EVFLAG) {
        double lj3, lj4, offs;
        if (ONETYPE) {
          lj3 = one.lj3; lj4 = one.lj4; offs = one.offset;
        } else {
          const int t = trow + pos_type(pj);
          lj3 = s_tab[3 * nt2 + t]; lj4 = s_tab[4 * nt2 + t]; offs = s_tab[5 * nt2 + t];
        }
        acc[0] += in ? 0.5 * (r6inv * (lj3 * r6inv - lj4) - offs) : 0.0;
        const double hf = 0.5 * fpair;
        acc[1] += dx * dx * hf;
        acc[2] += dy * dy * hf;
        acc[3] += dz * dz * hf;
        acc[4] += dx * dy * hf;
        acc[5] += dx * dz * hf;
        acc[6] += dy * dz * hf;
      }
      }
    }
    fx[i] = fxi;
    fy[i] = fyi;
    fz[i] = fzi;
  }
  if (EVFLAG) block_partials<7>(acc, partials);
}

// does this context want (and qualify for) the tiled kernels at the coming list build?  Also fixes the segment length.
bool pair_tiled_wanted(b200md_ctx *c)
{
  if (!c->tiled || c->cl != 1 || c->host_ghosts) return false;
  if (c->pair_style == 2 && !pair_eam_can_split(c)) return false;   // eam: the single-type kernels only
  if (c->pair_style != 1 && c->pair_style != 2) return false;
  if (c->tiled_g_user <= 0) {
    // bins per block: about 112 atoms per 128-thread block (lj melt: 18.5 atoms per bin -> 6; chain: 3 per bin -> 37)
    const double per_bin = (double) c->nlocal / ((double) c->bg.nbin[0] * c->bg.nbin[1] * c->bg.nbin[2]);
    int g = (int) (112.0 / (per_bin > 0.1 ? per_bin : 0.1) + 0.5);
    c->tiled_g = g < 1 ? 1 : (g > c->bg.nbin[0] ? c->bg.nbin[0] : g);
  } else
    c->tiled_g = c->tiled_g_user;
  return true;
}

// after a list build that emitted the 16-bit slot lists (k_build_bins<..., TILED16>; flags_h[6..7] are its verdict)
int pair_lj_tiled_prepare(b200md_ctx *c)
{
  const bool emitted = c->tiled_emitted;
  c->tiled_emitted = false;
  c->tiled_valid = false;
  if (!emitted) return B200MD_OK;
  const TileGeom T = tile_geom(c);
  c->tiled_nseg = T.nsegx * c->bg.nbin[1] * c->bg.nbin[2];
  c->tiled_max_staged = c->flags_h[6];
  const size_t nt2 = (size_t) (c->lj.ntypes + 1) * (c->lj.ntypes + 1);
  const size_t extra = c->pair_style == 2 ? (size_t) 7 * (c->eam.nr + 1) * sizeof(double) : (c->lj.ntypes == 1 ? 0 : 6 * nt2 * sizeof(double));
  const size_t smem = (size_t) c->tiled_max_staged * sizeof(double4) + extra;
  // usable only when every slot fits 16 bits and the staging area fits shared memory; else the default kernels run
  c->tiled_valid = !c->flags_h[7] && c->tiled_max_staged > 0 && c->tiled_max_staged < 0xFFFF && smem <= 200 * 1024;
  return B200MD_OK;
}

static int pair_lj_compute_tiled(b200md_ctx *c, int ev)
{
  const TileGeom T = tile_geom(c);
  const int nseg = c->tiled_nseg, onetype = (c->lj.ntypes == 1);
  const size_t nt2 = (size_t) (c->lj.ntypes + 1) * (c->lj.ntypes + 1);
  // one type: no w[] array and no tables (24 bytes per staged atom: 7 blocks per SM instead of 5 on the lj melt)
  const size_t smem = onetype ? (size_t) c->tiled_max_staged * 3 * sizeof(double)
                              : (size_t) c->tiled_max_staged * sizeof(double4) + 6 * nt2 * sizeof(double);
  LJOne one = {c->lj.h_onetype[0], c->lj.h_onetype[1], c->lj.h_onetype[2], c->lj.h_onetype[3], c->lj.h_onetype[4], c->lj.h_onetype[5]};
  if (ev) TRY(dev_ensure(c, c->partials, (size_t) nseg * 8));
  static bool attr_done[64] = {};  // per device: a process may hold contexts on several GPUs
  if (!attr_done[c->device & 63]) {
    CU(cudaFuncSetAttribute(k_lj_tiled<0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CU(cudaFuncSetAttribute(k_lj_tiled<0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CU(cudaFuncSetAttribute(k_lj_tiled<1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CU(cudaFuncSetAttribute(k_lj_tiled<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_done[c->device & 63] = true;
  }
  {
    KProf kp(c, KC_PAIR);
#define T_ARGS T, c->bin_start.p, c->cell_list.p, c->pos.p, c->numneigh.p, c->neigh16.p, c->maxneigh, c->nlocal, c->tiled_max_staged, one, c->lj.d, c->lj.ntypes, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p
    if (ev) {
      if (onetype) k_lj_tiled<1, 1><<<nseg, 128, smem, c->stream>>>(T_ARGS);
      else k_lj_tiled<1, 0><<<nseg, 128, smem, c->stream>>>(T_ARGS);
    } else {
      if (onetype) k_lj_tiled<0, 1><<<nseg, 128, smem, c->stream>>>(T_ARGS);
      else k_lj_tiled<0, 0><<<nseg, 128, smem, c->stream>>>(T_ARGS);
    }
#undef T_ARGS
    LAUNCH_CHECK();
  }
  if (ev) TRY(reduce_partials(c, nseg, 7, 0));
  return B200MD_OK;
}
