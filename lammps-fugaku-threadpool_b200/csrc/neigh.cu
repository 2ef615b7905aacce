// neigh.cu -- the reneighbor branch of Verlet::run (src/verlet.cpp:283-306) on the device:
//   Domain::pbc            (src/domain.cpp:530-600)      -> k_pbc
//   Atom::sort + NBinStandard::bin_atoms (src/atom.cpp:2246-2345, src/nbin_standard.cpp:221-261)
//                                                        -> counting sort k_bin_count / scan / k_bin_place /
//                                                           k_bin_sortseg / k_permute_owned
//   Comm::exchange / CommBrick::borders                  -> comm.cu (single-stage, 26 directions)
//   NPair::build           (src/npair_half_bin_atomonly_newton.cpp:33-113, full-list semantics of
//                           src/npair_full_bin_atomonly.cpp) -> k_build_seg (k_build_fp64: the all-FP64 fallback)
//   Neighbor::check_distance (src/neighbor.cpp:2322-2374) -> k_check_distance
// This is a from-scratch design: bins are >= cutneighmax wide (27-bin stencil, 9 contiguous x-rows),
// owned atoms are physically re-ordered by bin at every rebuild, ghosts are made in ONE stage from the
// 26 direction masks instead of the reference's 6 staged swaps (same ghost set when every sub-domain is
// at least one ghost cutoff wide), and the list is a full list in a blocked-transposed layout.
#include <type_traits>
#include <utility>

#include "ctx.cuh"

// ------------------------------------------------------------------------------------------------
// binning helpers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int bin_of(const BinGeom &g, double x, double y, double z)
{
  int bx = (int) floor((x - g.lo[0]) * g.inv[0]) + BIN_MARGIN;
  int by = (int) floor((y - g.lo[1]) * g.inv[1]) + BIN_MARGIN;
  int bz = (int) floor((z - g.lo[2]) * g.inv[2]) + BIN_MARGIN;
  bx = min(max(bx, 0), g.mbin[0] - 1);
  by = min(max(by, 0), g.mbin[1] - 1);
  bz = min(max(bz, 0), g.mbin[2] - 1);
  return (bz * g.mbin[1] + by) * g.mbin[0] + bx;
}

// Domain::pbc for an orthogonal box: wrap owned atoms, update image flags (src/domain.cpp:546-600)
__global__ void k_pbc(double4 *__restrict__ pos, int *__restrict__ image, int n, double3 lo, double3 hi, double3 prd,
                      int px, int py, int pz, int *__restrict__ flags)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 p = pos[i];
  if (!(isfinite(p.x) && isfinite(p.y) && isfinite(p.z))) { flags[0] = 1; return; }
  int img = image[i];
  if (px) {
    if (p.x < lo.x) { p.x += prd.x; int idim = img & IMGMASK; int other = img ^ idim; idim--; idim &= IMGMASK; img = other | idim; }
    if (p.x >= hi.x) { p.x -= prd.x; p.x = fmax(p.x, lo.x); int idim = img & IMGMASK; int other = img ^ idim; idim++; idim &= IMGMASK; img = other | idim; }
  }
  if (py) {
    if (p.y < lo.y) { p.y += prd.y; int idim = (img >> IMGBITS) & IMGMASK; int other = img ^ (idim << IMGBITS); idim--; idim &= IMGMASK; img = other | (idim << IMGBITS); }
    if (p.y >= hi.y) { p.y -= prd.y; p.y = fmax(p.y, lo.y); int idim = (img >> IMGBITS) & IMGMASK; int other = img ^ (idim << IMGBITS); idim++; idim &= IMGMASK; img = other | (idim << IMGBITS); }
  }
  if (pz) {
    if (p.z < lo.z) { p.z += prd.z; int idim = (img >> IMG2BITS) & IMGMASK; int other = img ^ (idim << IMG2BITS); idim--; idim &= IMGMASK; img = other | (idim << IMG2BITS); }
    if (p.z >= hi.z) { p.z -= prd.z; p.z = fmax(p.z, lo.z); int idim = (img >> IMG2BITS) & IMGMASK; int other = img ^ (idim << IMG2BITS); idim++; idim &= IMGMASK; img = other | (idim << IMG2BITS); }
  }
  pos[i] = p;
  image[i] = img;
}

// sub-cell of an atom inside its bin: 4x4x4 cells in Morton order.  Owned atoms are ordered by (bin, sub-cell,
// previous index) so that CL consecutive atoms -- one cluster of the cluster neighbor list -- are spatially
// compact (about half of a cluster's candidate neighbors then serve more than one of its atoms).
__device__ __forceinline__ int subcell_of(const BinGeom &g, double x, double y, double z)
{
  double fx = (x - g.lo[0]) * g.inv[0], fy = (y - g.lo[1]) * g.inv[1], fz = (z - g.lo[2]) * g.inv[2];
  int sx = min(3, max(0, (int) ((fx - floor(fx)) * 4.0)));
  int sy = min(3, max(0, (int) ((fy - floor(fy)) * 4.0)));
  int sz = min(3, max(0, (int) ((fz - floor(fz)) * 4.0)));
  return (sx & 1) | ((sy & 1) << 1) | ((sz & 1) << 2) | ((sx & 2) << 2) | ((sy & 2) << 3) | ((sz & 2) << 4);
}

template <int WITH_SUB>
__global__ void k_bin_count(const double4 *__restrict__ pos, int n, int nowned, BinGeom g, int *__restrict__ atom_bin,
                            int *__restrict__ bin_count, unsigned char *__restrict__ atom_sub)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 p = pos[i];
  int b;
  if (i < nowned) {
    // owned atoms always land in a bin of the sub-domain proper, also when (x - lo) * inv rounds up to nbin at the
    // upper face (the case NBin::coord2bin's `x >= hi` branch exists for, src/nbin.cpp:150-157): the warp-per-bin
    // build walks exactly these bins
    int bx = min(max((int) floor((p.x - g.lo[0]) * g.inv[0]), 0), g.nbin[0] - 1) + BIN_MARGIN;
    int by = min(max((int) floor((p.y - g.lo[1]) * g.inv[1]), 0), g.nbin[1] - 1) + BIN_MARGIN;
    int bz = min(max((int) floor((p.z - g.lo[2]) * g.inv[2]), 0), g.nbin[2] - 1) + BIN_MARGIN;
    b = (bz * g.mbin[1] + by) * g.mbin[0] + bx;
  } else
    b = bin_of(g, p.x, p.y, p.z);
  atom_bin[i] = b;
  if (WITH_SUB) atom_sub[i] = (unsigned char) subcell_of(g, p.x, p.y, p.z);
  atomicAdd(&bin_count[b], 1);
}

// ---- exclusive scan over n ints (three small kernels; n = mbins+1, at most a few million) ----------
#define SCAN_BLOCK 256
#define SCAN_ITEMS 4
__global__ void k_scan_block(const int *__restrict__ in, int *__restrict__ out, int *__restrict__ sums, int n)
{
  __shared__ int wsum[SCAN_BLOCK / 32];
  int base = blockIdx.x * SCAN_BLOCK * SCAN_ITEMS + threadIdx.x * SCAN_ITEMS;
  int v[SCAN_ITEMS], t = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { v[k] = (base + k < n) ? in[base + k] : 0; t += v[k]; }
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int incl = t;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
  if (lane == 31) wsum[w] = incl;
  __syncthreads();
  int woff = 0;
  for (int k = 0; k < w; k++) woff += wsum[k];
  int excl = woff + incl - t;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { if (base + k < n) out[base + k] = excl; excl += v[k]; }
  if (threadIdx.x == SCAN_BLOCK - 1) sums[blockIdx.x] = woff + incl;
}

__global__ void k_scan_sums(int *__restrict__ sums, int nblk)  // one block, in-place exclusive scan
{
  __shared__ int wsum[32];
  __shared__ int carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int start = 0; start < nblk; start += blockDim.x) {
    int i = start + threadIdx.x;
    int t = (i < nblk) ? sums[i] : 0;
    int incl = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) wsum[w] = incl;
    __syncthreads();
    int woff = 0;
    for (int k = 0; k < w; k++) woff += wsum[k];
    int carry = carry_s;
    if (i < nblk) sums[i] = carry + woff + incl - t;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry_s = carry + woff + incl;
    __syncthreads();
  }
}

__global__ void k_scan_add(int *__restrict__ out, const int *__restrict__ sums, int n)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] += sums[i / (SCAN_BLOCK * SCAN_ITEMS)];
}

static int exclusive_scan(b200md_ctx *c, const int *in, int *out, int n)
{
  int nblk = cdiv(n, SCAN_BLOCK * SCAN_ITEMS);
  TRY(dev_ensure(c, c->scan_tmp, (size_t) nblk + 1));
  k_scan_block<<<nblk, SCAN_BLOCK, 0, c->stream>>>(in, out, c->scan_tmp.p, n);
  LAUNCH_CHECK();
  if (nblk > 1) {
    k_scan_sums<<<1, 1024, 0, c->stream>>>(c->scan_tmp.p, nblk);
    LAUNCH_CHECK();
    k_scan_add<<<cdiv(n, 256), 256, 0, c->stream>>>(out, c->scan_tmp.p, n);
    LAUNCH_CHECK();
  }
  return B200MD_OK;
}

__global__ void k_bin_place(const int *__restrict__ atom_bin, int n, const int *__restrict__ bin_start,
                            int *__restrict__ cursor, int *__restrict__ list)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b = atom_bin[i];
  int slot = bin_start[b] + atomicAdd(&cursor[b], 1);
  list[slot] = i;
}

// the atomics above leave each bin's segment in arbitrary order: sort it so that the atom order (hence every
// floating-point summation order downstream) is reproducible run to run.  Key = (sub-cell, index) for the
// owned-atom sort, index alone for the owned+ghost cell list ("ascending index within a bin",
// src/nbin_standard.cpp:254-259).
#define SORTSEG_LOCAL 48
template <int WITH_SUB>
__global__ void k_bin_sortseg(const int *__restrict__ bin_start, int mbins, int *__restrict__ list,
                              const unsigned char *__restrict__ atom_sub)
{
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= mbins) return;
  const int s = bin_start[b], e = bin_start[b + 1], n = e - s;
  if (n <= 1) return;
  if (n <= SORTSEG_LOCAL) {
    long long key[SORTSEG_LOCAL];
    for (int t = 0; t < n; t++) {
      int idx = list[s + t];
      key[t] = WITH_SUB ? (((long long) atom_sub[idx] << 32) | (long long) idx) : (long long) idx;
    }
    for (int a = 1; a < n; a++) {
      long long kk = key[a];
      int k = a - 1;
      while (k >= 0 && key[k] > kk) { key[k + 1] = key[k]; k--; }
      key[k + 1] = kk;
    }
    for (int t = 0; t < n; t++) list[s + t] = (int) (key[t] & 0xffffffffLL);
  } else {
    for (int a = s + 1; a < e; a++) {
      int idx = list[a];
      long long kk = WITH_SUB ? (((long long) atom_sub[idx] << 32) | (long long) idx) : (long long) idx;
      int k = a - 1;
      while (k >= s) {
        int jdx = list[k];
        long long kj = WITH_SUB ? (((long long) atom_sub[jdx] << 32) | (long long) jdx) : (long long) jdx;
        if (kj <= kk) break;
        list[k + 1] = jdx;
        k--;
      }
      list[k + 1] = idx;
    }
  }
}

__global__ void k_permute_owned(const int *__restrict__ perm, int n, const double4 *__restrict__ pos,
                                double4 *__restrict__ pos_o, const double *__restrict__ vx, const double *__restrict__ vy,
                                const double *__restrict__ vz, double *__restrict__ vx_o, double *__restrict__ vy_o,
                                double *__restrict__ vz_o, const int *__restrict__ tag, int *__restrict__ tag_o,
                                const int *__restrict__ image, int *__restrict__ image_o, double4 *__restrict__ xhold)
{
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  int i = perm[s];
  const double4 p = pos[i];
  pos_o[s] = p;
  if (xhold) xhold[s] = p;   // Neighbor::build's xhold = x of the owned atoms at this build (src/neighbor.cpp:2412-2430)
  vx_o[s] = vx[i];
  vy_o[s] = vy[i];
  vz_o[s] = vz[i];
  tag_o[s] = tag[i];
  image_o[s] = image[i];
}

// counting sort of atoms [0,n) into bins: list = atom indices ordered by bin; within a bin by (sub-cell, index)
// when with_sub, else by index
static int bin_sort(b200md_ctx *c, int n, int *list, bool with_sub)
{
  const BinGeom &g = c->bg;
  TRY(dev_ensure(c, c->bin_count, (size_t) g.mbins + 1));
  TRY(dev_ensure(c, c->bin_start, (size_t) g.mbins + 1));
  TRY(dev_ensure(c, c->atom_bin, (size_t) n));
  if (with_sub) TRY(dev_ensure(c, c->atom_sub, (size_t) n));
  KProf kp(c, KC_SORT);
  CU(cudaMemsetAsync(c->bin_count.p, 0, ((size_t) g.mbins + 1) * sizeof(int), c->stream));
  if (with_sub) k_bin_count<1><<<cdiv(n, 256), 256, 0, c->stream>>>(c->pos.p, n, c->nlocal, g, c->atom_bin.p, c->bin_count.p, c->atom_sub.p);
  else k_bin_count<0><<<cdiv(n, 256), 256, 0, c->stream>>>(c->pos.p, n, c->nlocal, g, c->atom_bin.p, c->bin_count.p, nullptr);
  LAUNCH_CHECK();
  TRY(exclusive_scan(c, c->bin_count.p, c->bin_start.p, g.mbins + 1));
  CU(cudaMemsetAsync(c->bin_count.p, 0, ((size_t) g.mbins + 1) * sizeof(int), c->stream));
  k_bin_place<<<cdiv(n, 256), 256, 0, c->stream>>>(c->atom_bin.p, n, c->bin_start.p, c->bin_count.p, list);
  LAUNCH_CHECK();
  if (with_sub) k_bin_sortseg<1><<<cdiv(g.mbins, 128), 128, 0, c->stream>>>(c->bin_start.p, g.mbins, list, c->atom_sub.p);
  else k_bin_sortseg<0><<<cdiv(g.mbins, 128), 128, 0, c->stream>>>(c->bin_start.p, g.mbins, list, nullptr);
  LAUNCH_CHECK();
  return B200MD_OK;
}

// ------------------------------------------------------------------------------------------------
// bins: NBinStandard::setup_bins (src/nbin_standard.cpp:82-215) re-designed for a 27-bin stencil
// ------------------------------------------------------------------------------------------------
int neigh_setup_bins(b200md_ctx *c)
{
  if (!c->have_box || !c->have_neigh) return b200md_fail(c, B200MD_ERR_ARG, "box and neighbor settings must be set first");
  BinGeom &g = c->bg;
  double target = c->binsize_user > c->cutneighmax ? c->binsize_user : c->cutneighmax;
  long long mb = 1;
  for (int k = 0; k < 3; k++) {
    double len = c->subhi[k] - c->sublo[k];
    bool decomposed_or_periodic = c->periodic[k] || c->procgrid[k] > 1;
    if (!c->host_ghosts && decomposed_or_periodic && len < c->cutneighmax)
      return b200md_fail(c, B200MD_ERR_DOMAIN, "sub-domain (%g) thinner than the ghost cutoff (%g) in dim %d", len,
                         c->cutneighmax, k);
    int nb = (int) (len / target);
    if (nb < 1) nb = 1;
    while (nb > 1 && len / nb < c->cutneighmax * (1.0 + 1e-10)) nb--;
    double inv = nb / len;
    if (c->host_ghosts && len / nb < c->cutneighmax * (1.0 + 1e-10)) {
      // host-ghost mode accepts boxes thinner than the cutoff (the caller made several layers of periodic
      // images): bins stay >= cutneighmax wide, so the 27-bin stencil still covers the cutoff sphere, and simply
      // overhang the sub-domain
      const double bs = c->cutneighmax * (1.0 + 1e-10);
      nb = (int) ceil(len / bs);
      if (nb < 1) nb = 1;
      inv = 1.0 / bs;
    }
    g.nbin[k] = nb;
    g.mbin[k] = nb + 2 * BIN_MARGIN;
    g.lo[k] = c->sublo[k];
    g.inv[k] = inv;
    mb *= g.mbin[k];
  }
  if (mb > 2000000000LL) return b200md_fail(c, B200MD_ERR_ARG, "too many bins");
  g.mbins = (int) mb;
  return B200MD_OK;
}

// ------------------------------------------------------------------------------------------------
// neighbor build, all-FP64 form (fallback: boxes so large relative to the cutoff that the FP32 prefilter's error
// band would exceed 1 % of cutneighsq, see build_tolerance): one thread per owned atom scans its 27-bin stencil.
// Membership is decided exactly as the reference decides it: rsq = delx*delx + dely*dely + delz*delz (no FMA), keep if
// rsq <= cutneighsq[itype][jtype] (src/npair_half_bin_atomonly_newton.cpp:80-87), so the lists are bit-identical to
// the reference's as tag-pair sets.
// Layout: entry k of atom i at [(i>>5)*32*maxneigh + k*32 + (i&31)] (one coalesced 128-byte row per k per warp).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
k_build_fp64(const double4 *__restrict__ pos, int nlocal, BinGeom g, const int *__restrict__ bin_start,
             const int *__restrict__ cell_list, const double *__restrict__ cutneighsq, int ntypes, int maxneigh,
             int *__restrict__ numneigh, int *__restrict__ neigh, int *__restrict__ flags)
{
  extern __shared__ double s_cut[];
  const int nt1 = ntypes + 1;
  for (int k = threadIdx.x; k < nt1 * nt1; k += blockDim.x) s_cut[k] = cutneighsq[k];
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const double4 p = pos[i];
  const int trow = pos_type(p) * nt1;
  const int bx = min(max((int) floor((p.x - g.lo[0]) * g.inv[0]) + BIN_MARGIN, 1), g.mbin[0] - 2);
  const int by = min(max((int) floor((p.y - g.lo[1]) * g.inv[1]) + BIN_MARGIN, 1), g.mbin[1] - 2);
  const int bz = min(max((int) floor((p.z - g.lo[2]) * g.inv[2]) + BIN_MARGIN, 1), g.mbin[2] - 2);
  const size_t base = (size_t) (i >> 5) * TILE * maxneigh + (i & 31);
  int n = 0;
  for (int dz = -1; dz <= 1; dz++)
    for (int dy = -1; dy <= 1; dy++) {
      const int row = ((bz + dz) * g.mbin[1] + by + dy) * g.mbin[0] + bx - 1;
      const int s = bin_start[row], e = bin_start[row + 3];   // the 3 bins of an x-row are contiguous in cell_list
      for (int k = s; k < e; k++) {
        const int j = cell_list[k];
        const double4 pj = ld_pos256(pos + j);
        const double ddx = __dsub_rn(p.x, pj.x), ddy = __dsub_rn(p.y, pj.y), ddz = __dsub_rn(p.z, pj.z);
        const double rsq = __dadd_rn(__dadd_rn(__dmul_rn(ddx, ddx), __dmul_rn(ddy, ddy)), __dmul_rn(ddz, ddz));
        if (rsq <= s_cut[trow + pos_type(pj)] && j != i) {
          if (n < maxneigh) neigh[base + (size_t) n * TILE] = j;
          n++;
        }
      }
    }
  numneigh[i] = n;
  unsigned am = __activemask();
  int nmax = __reduce_max_sync(am, n);
  if ((threadIdx.x & 31) == (__ffs(am) - 1)) atomicMax(&flags[1], nmax);  // flags[1] = longest list seen
}

// ------------------------------------------------------------------------------------------------
// neighbor build, the default: ONE THREAD PER OWNED ATOM over candidates staged in shared memory.
//   * k_rec_fill writes every binned atom (owned + ghost, in cell_list order, bins padded to the vector width) as a
//     record {float x, y, z relative to the bin-grid origin; j | (type-1) << 28 (; tag)}, in groups of VW records.
//   * a block owns a SEGMENT of tg consecutive bins of one x-row (about 112 owned atoms).  Bins are >= the neighbor
//     cutoff wide, so every candidate of its atoms lies in the 3 x 3 x (tg + 2) bins around it: 9 CONTIGUOUS record
//     ranges, brought into shared memory by 9 bulk copies (cp.async.bulk + mbarrier).
//   * every thread then walks the 27 bins around ITS atom's bin (9 sub-ranges of the staged rows) with LDS.128 --
//     the lanes of a warp sit in 2 or 3 bins, so a step reads 2 or 3 distinct addresses -- VW candidates
//     per iteration through the packed FP32 pipe (fadd2 / fmul2 / ffma2).  The decision is final when
//     |rsq_f - cutneighsq| > tol, where tol rigorously bounds the FP32 error (derivation at build_tolerance()).  Inside
//     the band (about 1 candidate in 10^4) the reference's own FP64 expression decides:
//     rsq = delx*delx + dely*dely + delz*delz without FMA contraction, keep if rsq <= cutneighsq[itype][jtype]
//     (src/npair_half_bin_atomonly_newton.cpp:80-87).  List membership is therefore bit-identical to the all-FP64
//     kernel above and to the reference's list as tag-pair sets.
//   * entry order per atom = (dz, dy) rows, then cell_list order -- what every earlier form of this kernel emitted, so
//     the force summation order (and every bit of the forces) is unchanged.
// Why this form (ncu, profiles/r02_ncu_build.md): the warp-per-bin kernel of round 1 (lanes = the 18.5 atoms of a bin,
// candidates broadcast from global memory) was ISSUE-bound -- 81 % issue-slot utilisation, 757 warp instructions per
// atom: 27 candidate steps per atom (500 candidates over 18.5 useful lanes) at 28 instructions each.  Here a lane
// scans its own 500 candidates (15.6 steps per atom at 32 lanes) at 12 instructions per candidate (48 per group of 4;
// profiles/r02_ncu_build_eam_v2.md: now the L1 data pipe, 67 %, is the busiest unit).
// ------------------------------------------------------------------------------------------------
#define BR_JMASK 0x0FFFFFFF
#define SEG_FAR 1.0e18f   /* coordinate of a padding record: rsq = inf, never <= any cutoff */

// candidate records in cell_list order, every bin padded to the build's vector width VW (2 or 4) with records that can
// never be accepted, laid out in GROUPS of VW records: {x[VW], y[VW], z[VW], jt[VW] (, tag[VW])} -- NF = 4 or 5 32-bit
// fields -- so that the build reads a group's fields at fixed offsets from ONE pointer (LDS.128 / LDS.64 per field) and a
// staged row is one contiguous range.  k_pad_count: padded population of each bin (its exclusive scan is pbin_start)
__global__ void k_pad_count(const int *__restrict__ bin_start, int mbins, int vw, int *__restrict__ cnt)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > mbins) return;
  cnt[b] = b < mbins ? ((bin_start[b + 1] - bin_start[b] + vw - 1) & ~(vw - 1)) : 0;
}

// word of field f of record d (lg = log2 VW)
__host__ __device__ __forceinline__ size_t rec_word(int d, int f, int lg, int nf)
{
  return ((((size_t) (d >> lg)) * nf + f) << lg) + (d & ((1 << lg) - 1));
}

// record of cell_list position k: FP32 coordinates relative to the bin-grid origin, j | (type-1) << 28, tag (nf == 5)
__global__ void k_rec_fill(const double4 *__restrict__ pos, const int *__restrict__ cell_list, const int *__restrict__ atom_bin,
                           const int *__restrict__ bin_start, const int *__restrict__ pbin_start, int nall, double ox, double oy,
                           double oz, float *__restrict__ rec, int lg, int nf, const int *__restrict__ tag)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nall) return;
  const int j = cell_list[k];
  const int b = atom_bin[j];
  const int d = pbin_start[b] + k - bin_start[b];
  const double4 p = pos[j];
  int *reci = reinterpret_cast<int *>(rec);
  rec[rec_word(d, 0, lg, nf)] = (float) (p.x - ox);
  rec[rec_word(d, 1, lg, nf)] = (float) (p.y - oy);
  rec[rec_word(d, 2, lg, nf)] = (float) (p.z - oz);
  reci[rec_word(d, 3, lg, nf)] = j | ((pos_type(p) - 1) << 28);
  if (nf == 5) reci[rec_word(d, 4, lg, nf)] = tag[j];  // bonded / molecular systems: the candidate's tag, for the special-partner test
  if (k == bin_start[b + 1] - 1) {   // the bin's last record also writes its padding: never accepted
    for (int e = d + 1; e < pbin_start[b + 1]; e++) {
      rec[rec_word(e, 0, lg, nf)] = rec[rec_word(e, 1, lg, nf)] = rec[rec_word(e, 2, lg, nf)] = SEG_FAR;
      reci[rec_word(e, 3, lg, nf)] = BR_JMASK;
      if (nf == 5) reci[rec_word(e, 4, lg, nf)] = 0;
    }
  }
}

// ---- 1-D bulk copies global -> shared through the TMA engine (cp.async.bulk + mbarrier; SASS: UBLKCP, SYNCS) ----
#ifndef B200MD_CUDAEMU
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned phase)
{
  unsigned done = 0;
  while (!done)
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(done)
                 : "r"(smem_u32(bar)), "r"(phase)
                 : "memory");
}
#endif

// BONDED == 1 / 3 (in.chain): a candidate within the cutoff whose tag is one of the atom's 1-2 partners is NOT listed
// (special_bonds lj weight 0, NPair::find_special src/npair.h:125-158 -> -1) but recorded as the device index of that
// bond partner; of several periodic images the closest wins (Domain::closest_image).
struct BondBuild {
  const int *tag;              // [nall]
  const int *partner_by_tag;   // [ntags][maxb], 0 = empty
  int *bond_idx;               // [nlocal][maxb] out
  int maxb;
  // BONDED == 2 (host-ghost molecular systems): the general special-bond lists, NPair::find_special (src/npair.h:125-158)
  const int *spec_n3;          // [nlocal][3] cumulative counts of 1-2, 1-3, 1-4 partners
  const int *spec_tags;        // [nlocal][maxspecial]
  int maxspecial;
  int flag[4];                 // 0 drop, 1 ordinary neighbor, 2 keep with the class in bits 28-29 of the entry
};
#define SEG_THREADS 128       /* default block: about 112 owned atoms */
#define SEG_MAX_THREADS 256   /* "build_threads" (b200md_tune) may ask for this many */
// shared-memory plan of a segment: `cap` candidate records in the grouped layout of k_rec_fill (a thread reads the VW
// candidates of a group with one LDS.128 (VW = 4, dense bins) or LDS.64 (VW = 2, sparse bins such as in.chain's 3 beads
// per bin, where padding to 4 would add a third more candidates) per field, all at fixed offsets from one pointer), every
// bin range VW-aligned: no per-candidate range tests.  A staged row (the 3 x (tg + 2) bins of one (dz, dy)) is ONE
// contiguous range of the global records: where a group is a multiple of 16 bytes it goes to shared memory as one 1-D
// bulk copy through the TMA engine (cp.async.bulk, completion on an mbarrier) -- 9 copies issued by one thread.
//   float rec[cap * NF];
//   int bs[9][tg+3]  cell_list position of each staged bin's first record (+ the end);
//   int po[9][tg+3]  staged offset (in records) of each staged bin (+ the end)
// BONDED: 0 atomic; 1 bonds, any box (of several periodic images of a partner the closest wins); 2 the general special
// lists of host-ghost molecular systems; 3 bonds, at most 2 per atom, in a box wider than 2 cutoffs in every periodic
// dimension (a tag then appears at most once inside the cutoff: no image choice to make) -- in.chain.
struct RecArrays {
  const float *rec;         // [rec_cap * NF]
  const int *pbin_start;    // [mbins + 1]
};
template <int ONETYPE, int BONDED, int VW>
__global__ void __launch_bounds__(SEG_MAX_THREADS)
k_build_seg(const double4 *__restrict__ pos, RecArrays rec, int nlocal, BinGeom g,
            const int *__restrict__ bin_start, const double *__restrict__ cutneighsq, int ntypes, float tol, float hw, int maxneigh,
            int *__restrict__ numneigh, int *__restrict__ neigh, int *__restrict__ flags, int *__restrict__ tile_ghost,
            BondBuild bb, int tg, int cap)
{
  constexpr int NF = BONDED ? 5 : 4;   // 32-bit fields per record
  constexpr int GW = NF * VW;          // words per group
  constexpr int LG = VW == 4 ? 2 : 1;
  extern __shared__ int4 s_dyn[];
  float *s_rec = reinterpret_cast<float *>(s_dyn);
  int *s_bs = reinterpret_cast<int *>(s_dyn) + (size_t) cap * NF;
  const int W = tg + 3;
  int *s_po = s_bs + 9 * W;
  __shared__ double s_cut[(B200MD_MAXTYPES + 1) * (B200MD_MAXTYPES + 1)];
  __shared__ float s_cutf[(B200MD_MAXTYPES + 1) * (B200MD_MAXTYPES + 1)];
  __shared__ int s_row[10];                       // staged offset of each row (+ the total)
  __shared__ int s_g0[9], s_len[9];               // global padded start and length (records) of each row
  __shared__ unsigned long long s_bar;            // mbarrier of the bulk copies (8-byte aligned by its type)
  const int nt1 = ntypes + 1;
  for (int k = threadIdx.x; k < nt1 * nt1; k += blockDim.x) { s_cut[k] = cutneighsq[k]; s_cutf[k] = (float) cutneighsq[k]; }
  const int nsegx = (g.nbin[0] + tg - 1) / tg;
  const int seg = blockIdx.x;
  const int sx = seg % nsegx, oby = (seg / nsegx) % g.nbin[1], obz = seg / (nsegx * g.nbin[1]);
  const int gx = min(tg, g.nbin[0] - sx * tg);
  const int bx0 = BIN_MARGIN + sx * tg, by = oby + BIN_MARGIN, bz = obz + BIN_MARGIN;
  const int nq = gx + 3;   // bin starts per row: bins bx0-1 .. bx0+gx, plus the end of the last
  // s_po first holds the padded global start of every staged bin; turned into staged offsets below
  for (int idx = threadIdx.x; idx < 9 * nq; idx += blockDim.x) {
    const int r = idx / nq, q = idx - r * nq;
    const int first = ((bz + r / 3 - 1) * g.mbin[1] + (by + r % 3 - 1)) * g.mbin[0] + bx0 - 1;
    s_bs[r * W + q] = bin_start[first + q];
    s_po[r * W + q] = rec.pbin_start[first + q];
  }
  constexpr bool BULK_OK = (GW * 4) % 16 == 0;   // a row's bytes and its two addresses are multiples of 16
#ifndef B200MD_CUDAEMU
  if (BULK_OK && threadIdx.x == 0) mbar_init(&s_bar, 1);
#endif
  __syncthreads();
  if (threadIdx.x == 0) {
    int off = 0;
    for (int r = 0; r < 9; r++) {
      s_row[r] = off;
      s_g0[r] = s_po[r * W];
      s_len[r] = s_po[r * W + gx + 2] - s_po[r * W];
      off += s_len[r];
    }
    s_row[9] = off;
  }
  __syncthreads();
  const int total = s_row[9];
  if (total > cap) {   // denser than the staging area allows: report the size wanted, the host grows it and redoes the build
    if (threadIdx.x == 0) atomicMax(&flags[6], total);
    return;
  }
  // stage the 9 rows
  bool bulk = false;
#ifndef B200MD_CUDAEMU
  bulk = BULK_OK;
  if (bulk && threadIdx.x == 0) {
    mbar_expect_tx(&s_bar, (unsigned) total * (unsigned) (NF * 4));
    for (int r = 0; r < 9; r++) {
      const int gs = s_g0[r], len = s_len[r], so = s_row[r];
      if (len == 0) continue;
      bulk_g2s(s_rec + (size_t) so * NF, rec.rec + (size_t) gs * NF, (unsigned) len * (unsigned) (NF * 4), &s_bar);
    }
  }
#endif
  if (!bulk) {   // groups of 40 bytes (VW = 2 with tags): 8-byte units
#pragma unroll 1
    for (int r = 0; r < 9; r++) {
      const int gs = s_g0[r], len = s_len[r], so = s_row[r];
      const float2 *src = reinterpret_cast<const float2 *>(rec.rec + (size_t) gs * NF);
      float2 *dst = reinterpret_cast<float2 *>(s_rec + (size_t) so * NF);
      for (int k = threadIdx.x; k < len * NF / 2; k += blockDim.x) dst[k] = __ldg(src + k);
    }
  }
  // global padded starts -> staged offsets (every element on its own: the row's start was saved in s_g0)
  for (int idx = threadIdx.x; idx < 9 * nq; idx += blockDim.x) {
    const int r = idx / nq, q = idx - r * nq;
    s_po[r * W + q] += s_row[r] - s_g0[r];
  }
#ifndef B200MD_CUDAEMU
  if (bulk) mbar_wait(&s_bar, 0);
#endif
  __syncthreads();
  const int *s_reci = reinterpret_cast<const int *>(s_rec);
  auto word = [](int d, int f) -> int { return (((d >> LG) * NF + f) << LG) + (d & (VW - 1)); };
  const float cut1 = s_cutf[nt1 + 1];
  const int own_s = s_bs[4 * W + 1], own_e = s_bs[4 * W + 1 + gx];
  // a group's appends start at row min(n, ncap) and take at most VW rows: they stay inside the tile's rows without a
  // clamp per entry.  Past ncap the rows are wrong, but the count stays true and the host redoes every build whose
  // longest list exceeds maxneigh - 4 with longer rows.
  const int ncap = maxneigh - VW;
  int nmax = 0;
  // the segment's own records (centre row, bins 1..gx); owned atoms come first in every bin's records (cell_list is
  // index-ordered within a bin, ghosts have index >= nlocal)
  for (int slot = own_s + threadIdx.x; slot < own_e; slot += blockDim.x) {
    // my bin inside the segment, from my position in cell_list (exact; a float coordinate is not): the last bl in
    // [0, gx) with s_bs[4W + bl + 1] <= slot (bin starts ascend; empty bins repeat a start)
    int bl = 0;
    for (int hi = gx - 1; bl < hi;) {
      const int mid = (bl + hi + 1) >> 1;
      if (s_bs[4 * W + mid + 1] <= slot) bl = mid;
      else hi = mid - 1;
    }
    const int mine = s_po[4 * W + bl + 1] + slot - s_bs[4 * W + bl + 1];
    const int myjt = s_reci[word(mine, 3)];
    const int i = myjt & BR_JMASK;
    if (i >= nlocal) continue;
    const int trow = ONETYPE ? 0 : (((unsigned) myjt >> 28) + 1) * nt1;
    int *const nbp = neigh + ((size_t) (i >> 5) * TILE * maxneigh + (i & 31));
    const float mx = s_rec[word(mine, 0)], my = s_rec[word(mine, 1)], mz = s_rec[word(mine, 2)];
    const float2 nx = make_float2(-mx, -mx), ny = make_float2(-my, -my), nz = make_float2(-mz, -mz);
    const float2 ncut = make_float2(-cut1, -cut1);
    int n = 0;
    bool ghost_near = false;   // overlap mode: a ghost record among my candidates (a superset of "a ghost in my list")
    int ptag[B200MD_MAXBONDS], pbest[B200MD_MAXBONDS];
    float prsq[B200MD_MAXBONDS];
    int ptag0 = 0, ptag1 = 0, pb0 = -1, pb1 = -1;
    if (BONDED == 1) {
      const int ti = bb.tag[i];
#pragma unroll
      for (int m = 0; m < B200MD_MAXBONDS; m++) {
        ptag[m] = (m < bb.maxb) ? bb.partner_by_tag[(size_t) (ti - 1) * bb.maxb + m] : 0;
        pbest[m] = -1;
        prsq[m] = 3.0e38f;
      }
    }
    if (BONDED == 3) {
      const int ti = s_reci[word(mine, 4)];   // my own tag, from my staged record
      ptag0 = bb.partner_by_tag[(size_t) (ti - 1) * bb.maxb];
      ptag1 = bb.maxb > 1 ? bb.partner_by_tag[(size_t) (ti - 1) * bb.maxb + 1] : 0;
    }
    // the ambiguity band (about 1 candidate in 10^4): the reference's FP64 expression decides
    auto exact = [&](const int jt) -> bool {
      const double4 pi = pos[i], pj = pos[jt & BR_JMASK];
      const double ddx = __dsub_rn(pi.x, pj.x), ddy = __dsub_rn(pi.y, pj.y), ddz = __dsub_rn(pi.z, pj.z);
      const double r2 = __dadd_rn(__dadd_rn(__dmul_rn(ddx, ddx), __dmul_rn(ddy, ddy)), __dmul_rn(ddz, ddz));
      return r2 <= s_cut[(ONETYPE ? nt1 : trow) + (ONETYPE ? 1 : ((unsigned) jt >> 28) + 1)];
    };
    // the VW verdicts of a group appended in order.  SELF: the centre row alone holds my own record.
    auto append_group = [&](auto self_c, const bool (&a)[VW], const int (&js)[VW], const int (&ts)[VW], const float (&rs)[VW]) {
      constexpr bool SELF = decltype(self_c)::value;
      const int n0 = min(n, ncap);
      int row = n0;
#pragma unroll
      for (int u = 0; u < VW; u++) {
        const int j = ONETYPE ? js[u] : (js[u] & BR_JMASK);   // one type: the type bits are zero
        bool accept = a[u];
        if (SELF) accept = accept && j != i;
        int entry = j;
        if (BONDED == 1) {
          // branch-free: with 2 bonds per bead some lane of a warp meets a partner at most steps, and a divergent body
          // run by 3-4 lanes was a fifth of this kernel's instructions (ncu source view, profiles/r02_ncu_build.md).
          // Empty slots hold tag 0, padding records carry tag 0 but are never accepted.
          const int tj = ts[u];
          bool partner = false;
#pragma unroll
          for (int m = 0; m < B200MD_MAXBONDS; m++) {
            const bool hit = accept && tj == ptag[m];
            const bool closer = hit && rs[u] < prsq[m];
            partner |= hit;
            prsq[m] = closer ? rs[u] : prsq[m];
            pbest[m] = closer ? j : pbest[m];
          }
          accept = accept && !partner;
        }
        if (BONDED == 3) {   // a partner inside the cutoff is unique: no distance to compare
          const int tj = ts[u];
          const bool h0 = accept && tj == ptag0, h1 = accept && tj == ptag1;
          pb0 = h0 ? j : pb0;
          pb1 = h1 ? j : pb1;
          accept = accept && !(h0 || h1);
        }
        if (BONDED == 2 && accept) {   // NPair::find_special: first match in the atom's 1-2 | 1-3 | 1-4 list decides
          const int tj = ts[u];
          const int n1 = bb.spec_n3[3 * i], n2 = bb.spec_n3[3 * i + 1], n3 = bb.spec_n3[3 * i + 2];
          const int *sp = bb.spec_tags + (size_t) i * bb.maxspecial;
          for (int m = 0; m < n3; m++)
            if (sp[m] == tj) {
              const int which = m < n1 ? 1 : (m < n2 ? 2 : 3);
              if (bb.flag[which] == 0) accept = false;
              else if (bb.flag[which] == 2) entry = j | (which << CLBITS);
              break;
            }
        }
        if (accept) {
          nbp[(size_t) row * TILE] = entry;
          row++;
        }
      }
      n += row - n0;
    };
    // one group: VW squared distances through the packed FP32 pipe (fadd2 / fmul2 / ffma2; same operations and rounding
    // as the scalar form rsq = fma(dz, dz, fma(dy, dy, dx * dx)) that build_tolerance() bounds), d = rsq - cutoff, the
    // verdict d <= 0 on the common path, and ONE rare branch for the group: taken iff some |d| <= hw (hw covers the
    // rigorous FP32 error bound tol and the rounding of d), where the FP64 expression decides for those candidates
    auto group = [&](auto self_c, const float *gp) {
      float rs[VW], d[VW];
      int js[VW], ts[VW];
      if constexpr (VW == 4) {
        const float4 X = *reinterpret_cast<const float4 *>(gp), Y = *reinterpret_cast<const float4 *>(gp + 4),
                     Z = *reinterpret_cast<const float4 *>(gp + 8);
        const int4 J = *reinterpret_cast<const int4 *>(gp + 12);
        const float2 dxa = __fadd2_rn(make_float2(X.x, X.y), nx), dxb = __fadd2_rn(make_float2(X.z, X.w), nx);
        const float2 dya = __fadd2_rn(make_float2(Y.x, Y.y), ny), dyb = __fadd2_rn(make_float2(Y.z, Y.w), ny);
        const float2 dza = __fadd2_rn(make_float2(Z.x, Z.y), nz), dzb = __fadd2_rn(make_float2(Z.z, Z.w), nz);
        const float2 ra = __ffma2_rn(dza, dza, __ffma2_rn(dya, dya, __fmul2_rn(dxa, dxa)));
        const float2 rb = __ffma2_rn(dzb, dzb, __ffma2_rn(dyb, dyb, __fmul2_rn(dxb, dxb)));
        rs[0] = ra.x; rs[1] = ra.y; rs[2] = rb.x; rs[3] = rb.y;
        js[0] = J.x; js[1] = J.y; js[2] = J.z; js[3] = J.w;
        if (BONDED) {
          const int4 T = *reinterpret_cast<const int4 *>(gp + 16);
          ts[0] = T.x; ts[1] = T.y; ts[2] = T.z; ts[3] = T.w;
        }
        if (ONETYPE) {
          const float2 da = __fadd2_rn(ra, ncut), db = __fadd2_rn(rb, ncut);
          d[0] = da.x; d[1] = da.y; d[2] = db.x; d[3] = db.y;
        }
      } else {
        const float2 X = *reinterpret_cast<const float2 *>(gp), Y = *reinterpret_cast<const float2 *>(gp + 2),
                     Z = *reinterpret_cast<const float2 *>(gp + 4);
        const int2 J = *reinterpret_cast<const int2 *>(gp + 6);
        const float2 dxa = __fadd2_rn(X, nx), dya = __fadd2_rn(Y, ny), dza = __fadd2_rn(Z, nz);
        const float2 ra = __ffma2_rn(dza, dza, __ffma2_rn(dya, dya, __fmul2_rn(dxa, dxa)));
        rs[0] = ra.x; rs[1] = ra.y;
        js[0] = J.x; js[1] = J.y;
        if (BONDED) {
          const int2 T = *reinterpret_cast<const int2 *>(gp + 8);
          ts[0] = T.x; ts[1] = T.y;
        }
        if (ONETYPE) {
          const float2 da = __fadd2_rn(ra, ncut);
          d[0] = da.x; d[1] = da.y;
        }
      }
      if (!ONETYPE) {
#pragma unroll
        for (int u = 0; u < VW; u++) d[u] = rs[u] - s_cutf[trow + ((unsigned) js[u] >> 28) + 1];
      }
      bool acc[VW];
      float dmin = fabsf(d[0]);
#pragma unroll
      for (int u = 0; u < VW; u++) {
        acc[u] = d[u] <= 0.f;
        if (u) dmin = fminf(dmin, fabsf(d[u]));
      }
      if (__builtin_expect(dmin <= hw, 0)) {
        bool a2[VW];
#pragma unroll
        for (int u = 0; u < VW; u++) a2[u] = fabsf(d[u]) <= hw ? exact(js[u]) : acc[u];
        append_group(self_c, a2, js, ts, rs);
      } else {
        append_group(self_c, acc, js, ts, rs);
      }
    };
#pragma unroll 1
    for (int r = 0; r < 9; r++) {
      const int lo = s_po[r * W + bl], hi = s_po[r * W + bl + 3];   // both multiples of VW
      if (tile_ghost) {   // ghosts are the last records of a bin: one look per stencil bin
#pragma unroll
        for (int q = 0; q < 3; q++) {
          const int cnt = s_bs[r * W + bl + q + 1] - s_bs[r * W + bl + q];
          if (cnt > 0) ghost_near |= (s_reci[word(s_po[r * W + bl + q] + cnt - 1, 3)] & BR_JMASK) >= nlocal;
        }
      }
      const float *gp = s_rec + (size_t) lo * NF, *const ge = s_rec + (size_t) hi * NF;
      if (r == 4) {
        for (; gp < ge; gp += GW) group(std::true_type{}, gp);
      } else {
        for (; gp < ge; gp += GW) group(std::false_type{}, gp);
      }
    }
    numneigh[i] = n;
    if (BONDED == 1) {
#pragma unroll
      for (int m = 0; m < B200MD_MAXBONDS; m++)
        if (m < bb.maxb) {
          bb.bond_idx[(size_t) i * bb.maxb + m] = pbest[m];
          if (ptag[m] != 0 && pbest[m] < 0) flags[5] = 1;  // "Bond atoms missing" (src/ntopo_bond_all.cpp:56-61)
        }
    }
    if (BONDED == 3) {
      bb.bond_idx[(size_t) i * bb.maxb] = pb0;
      if (bb.maxb > 1) bb.bond_idx[(size_t) i * bb.maxb + 1] = pb1;
      if ((ptag0 != 0 && pb0 < 0) || (ptag1 != 0 && pb1 < 0)) flags[5] = 1;
    }
    if (tile_ghost && ghost_near) tile_ghost[i >> 5] = 1;   // several lanes may write the same 1
    nmax = max(nmax, n);
  }
  nmax = __reduce_max_sync(0xffffffffu, nmax);
  // flags[1] = longest list seen; the plain read only skips atomics that could not raise it
  if ((threadIdx.x & 31) == 0 && nmax > flags[1]) atomicMax(&flags[1], nmax);
}

// overlap mode: the tiles as two compact lists, ghost-free tiles first (ascending), then the others (ascending), from
// the flags the build left and their exclusive scan; flags[7] = number of tiles in the second list
__global__ void k_tile_lists(const int *__restrict__ flag, const int *__restrict__ pos, int ntiles, int *__restrict__ list,
                             int *__restrict__ flags)
{
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ntiles) return;
  const int nghost = pos[ntiles - 1] + (flag[ntiles - 1] != 0);
  const int gflag = flag[t] != 0, p = pos[t];
  if (gflag) list[ntiles - nghost + p] = t;
  else list[t - p] = t;
  if (t == 0) flags[7] = nghost;
}

// Rigorous bound on |rsq_f - rsq| of k_build_seg for any pair with r <= 2*rc (pairs farther apart differ from the
// cutoff by more than 3*rc^2, far outside any band): coordinates are stored relative to the bin-grid origin, so
// |x_f - x| <= u*M with u = 2^-24 and M the grid extent; dx_f carries both conversion errors plus one rounding:
// |dx_f - dx| <= 2u(M + r); three squared terms and three roundings give
// |rsq_f - rsq| <= 2*sqrt(3)*r*2u(M + r) + 3u*rsq + (float)cutneighsq conversion (u*rc^2).
// tol is that bound at r = 2*rc with another factor 2 of slack.  The FP32 path is used only while tol stays
// below 1 % of the smallest cutneighsq (huge boxes with tiny cutoffs fall back to the all-FP64 kernel).
static double build_tolerance(const b200md_ctx *c, double *cutsq_min_out)
{
  const BinGeom &g = c->bg;
  double M = 0.0;
  for (int k = 0; k < 3; k++) M = fmax(M, g.mbin[k] / g.inv[k]);
  const double u = 1.0 / 16777216.0, rc = c->cutneighmax, r = 2.0 * rc;
  double cmin = 1e300;
  for (int i = 1; i <= c->ntypes; i++)
    for (int j = 1; j <= c->ntypes; j++) cmin = fmin(cmin, c->cutneighsq_h[i * (c->ntypes + 1) + j]);
  if (cutsq_min_out) *cutsq_min_out = cmin;
  return 2.0 * (2.0 * 1.7320508075688772 * r * 2.0 * u * (M + r) + 3.0 * u * r * r + u * rc * rc);
}

__global__ void k_check_distance(const double4 *__restrict__ pos, const double4 *__restrict__ xhold, int n, double triggersq,
                                 int *__restrict__ flags)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (moved_beyond_trigger(pos[i], xhold[i], triggersq)) flags[2] = 1;
}

int neigh_check_distance(b200md_ctx *c, int *moved)
{
  if (c->check_ahead) {   // the previous step's fused integration already compared these positions with xhold
    c->check_ahead = false;
  } else {
    CU(cudaMemsetAsync(c->flags_d + 2, 0, sizeof(int), c->stream));
    k_check_distance<<<cdiv(c->nlocal, 256), 256, 0, c->stream>>>(c->pos.p, c->xhold.p, c->nlocal, c->triggersq, c->flags_d);
    LAUNCH_CHECK();
  }
  TRY(comm_allreduce_max_int(c, c->flags_d + 2, 1));  // the MPI_Allreduce of Neighbor::check_distance (src/neighbor.cpp:2371)
  CU(cudaMemcpyAsync(c->flags_h + 2, c->flags_d + 2, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  *moved = c->flags_h[2];
  return B200MD_OK;
}

// ------------------------------------------------------------------------------------------------
// the whole reneighbor sequence
// ------------------------------------------------------------------------------------------------
// development aid: B200MD_DEBUG_TIMING=1 prints the wall time of every phase of a rebuild (adds a sync per phase)
#include <chrono>
struct PhaseTimer {
  b200md_ctx *c;
  bool on;
  std::chrono::steady_clock::time_point t0;
  PhaseTimer(b200md_ctx *ctx) : c(ctx), on(ctx->debug_timing != 0), t0(std::chrono::steady_clock::now()) {}
  void mark(const char *what)
  {
    if (!on) return;
    cudaStreamSynchronize(c->stream);
    auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "[b200md rebuild] %-14s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};

int neigh_rebuild(b200md_ctx *c)
{
  PhaseTimer ph(c);
  if (c->nlocal <= 0) return b200md_fail(c, B200MD_ERR_ARG, "no atoms uploaded");
  TRY(neigh_setup_bins(c));
  if (c->bonds.maxb) {  // "Bad FENE bond" raised by k_bond_fene since the last build (bond_fene.cpp:88)
    CU(cudaMemcpyAsync(c->flags_h + 4, c->flags_d + 4, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (c->flags_h[4]) return b200md_fail(c, B200MD_ERR_ARG, "Bad FENE bond");
  }
  CU(cudaMemsetAsync(c->flags_d, 0, 8 * sizeof(int), c->stream));
  int n = c->nlocal;

  // Domain::pbc
  k_pbc<<<cdiv(n, 256), 256, 0, c->stream>>>(c->pos.p, c->image.p, n, make_double3(c->boxlo[0], c->boxlo[1], c->boxlo[2]),
                                              make_double3(c->boxhi[0], c->boxhi[1], c->boxhi[2]),
                                              make_double3(c->prd[0], c->prd[1], c->prd[2]), c->periodic[0], c->periodic[1],
                                              c->periodic[2], c->flags_d);
  LAUNCH_CHECK();

  ph.mark("pbc");
  // Comm::exchange: atoms that left this rank's sub-domain migrate (multi-rank only)
  TRY(comm_exchange(c));
  ph.mark("exchange");
  n = c->nlocal;
  // fix langevin replay: the reference's own Atom::sort reorders the atoms its random stream is consumed over
  // (Verlet::setup src/verlet.cpp:121, Verlet::run :291)
  // (counter mode draws by (tag, timestep): no order to replay, and after a migration the order list would be stale)
  if (c->langevin.on && c->langevin.rng_mode == 0 && c->langevin.sortfreq > 0 && c->ntimestep >= c->langevin.nextsort) TRY(langevin_resort(c));

  // spatial sort of owned atoms (Atom::sort) by neighbor bin
  TRY(dev_ensure(c, c->perm, (size_t) n));
  TRY(bin_sort(c, n, c->perm.p, true));
  ph.mark("sort owned");
  // same capacity as the live arrays, no slack: the pairs are swapped at every rebuild and must not leap-frog
  TRY(dev_ensure(c, c->pos_tmp, c->pos.n, false, false));
  TRY(dev_ensure(c, c->tag_tmp, c->tag.n, false, false));
  TRY(dev_ensure(c, c->image_tmp, c->image.n, false, false));
  for (int k = 0; k < 3; k++) TRY(dev_ensure(c, c->v_tmp[k], c->v[k].n, false, false));
  if (c->dist_check) TRY(dev_ensure(c, c->xhold, (size_t) n));
  k_permute_owned<<<cdiv(n, 256), 256, 0, c->stream>>>(c->perm.p, n, c->pos.p, c->pos_tmp.p, c->v[0].p, c->v[1].p, c->v[2].p,
                                                        c->v_tmp[0].p, c->v_tmp[1].p, c->v_tmp[2].p, c->tag.p, c->tag_tmp.p,
                                                        c->image.p, c->image_tmp.p, c->dist_check ? c->xhold.p : nullptr);
  LAUNCH_CHECK();
  c->xhold_by_permute = c->dist_check != 0;
  std::swap(c->pos, c->pos_tmp);
  std::swap(c->tag, c->tag_tmp);
  std::swap(c->image, c->image_tmp);
  for (int k = 0; k < 3; k++) std::swap(c->v[k], c->v_tmp[k]);

  ph.mark("permute");
  // ghosts
  TRY(comm_borders(c));
  ph.mark("borders");
  return neigh_build_lists(c);
}

// bin owned + ghost atoms and build the list from the positions now on the device (also the entry point of the
// host-buffer mode, where the caller supplies the ghosts: hostmode.cu)
int neigh_build_lists(b200md_ctx *c)
{
  PhaseTimer ph(c);
  const int n = c->nlocal;
  const int nall = n + c->nghost;
  if (nall > CLJMASK) return b200md_fail(c, B200MD_ERR_ARG, "more than 2^28 owned+ghost atoms on one GPU");

  // bin owned + ghost
  TRY(dev_ensure(c, c->cell_list, (size_t) nall));
  TRY(bin_sort(c, nall, c->cell_list.p, false));
  ph.mark("bin all");

  // per-atom arrays that follow nlocal / nall
  TRY(dev_ensure(c, c->numneigh, (size_t) n));
  for (int k = 0; k < 3; k++) TRY(dev_ensure(c, c->f[k], (size_t) n));
  if (c->pair_style == 2) {
    TRY(dev_ensure(c, c->rho, (size_t) n));
    TRY(dev_ensure(c, c->fp, (size_t) nall));
  }

  if (c->maxneigh == 0) {
    double vol = 1.0;
    for (int k = 0; k < 3; k++) vol *= c->subhi[k] - c->sublo[k];
    double expect = 4.18879020478639 * c->cutneighmax * c->cutneighmax * c->cutneighmax * n / vol;
    int m = (int) (expect * 1.2) + 24;
    c->maxneigh = (m + 7) / 8 * 8;
  }
  const int ntiles = cdiv(n, TILE);
  CU(cudaMemsetAsync(c->flags_d + 6, 0, 2 * sizeof(int), c->stream));
  for (int attempt = 0; attempt < 8; attempt++) {
    TRY(dev_ensure(c, c->neigh, (size_t) ntiles * TILE * c->maxneigh));
    size_t smem = (size_t) (c->ntypes + 1) * (c->ntypes + 1) * sizeof(double);
    double cutsq_min = 0.0;
    const double tol = build_tolerance(c, &cutsq_min);
    if (tol < 0.01 * cutsq_min && !c->build_fp64) {
      const BinGeom &g = c->bg;
      KProf kp(c, KC_NEIGH);
      const int nown = g.nbin[0] * g.nbin[1] * g.nbin[2];
      int *tile_ghost = nullptr;
      if (c->overlap) {  // tile flags for the interior / boundary split of the force kernels
        TRY(dev_ensure(c, c->tile_ghost, (size_t) ntiles));
        CU(cudaMemsetAsync(c->tile_ghost.p, 0, (size_t) ntiles * sizeof(int), c->stream));
        tile_ghost = c->tile_ghost.p;
      }
      c->tiles_valid = tile_ghost != nullptr;
      BondBuild bb = {nullptr, nullptr, nullptr, 0, nullptr, nullptr, 0, {1, 1, 1, 1}};
      const bool bonded = c->bonds.maxb > 0, special = c->special.on && !bonded;
      if (special) {
        if (!c->host_ghosts) return b200md_fail(c, B200MD_ERR_ARG, "general special-bond lists are a host-ghost mode feature (b200md_host_special)");
        bb.spec_n3 = c->special.n3.p;
        bb.spec_tags = c->special.tags.p;
        bb.maxspecial = c->special.maxspecial;
        for (int k = 0; k < 4; k++) bb.flag[k] = c->special.flag[k];
      }
      if (bonded) {
        TRY(dev_ensure(c, c->bonds.idx, (size_t) n * c->bonds.maxb));
        bb.tag = c->tag.p;
        bb.partner_by_tag = c->bonds.partner_by_tag.p;
        bb.bond_idx = c->bonds.idx.p;
        bb.maxb = c->bonds.maxb;
      }
      // bins per block: about 112 owned atoms on 128 threads (lj melt: 18.5 atoms per bin -> 6 bins; chain melt: 3 per
      // bin -> 37); "build_gb" overrides (tests)
      const double per_bin = (double) n / (nown > 0 ? nown : 1);
      const int bthreads = c->build_threads == SEG_MAX_THREADS ? SEG_MAX_THREADS : SEG_THREADS;
      int tg = c->build_gb_user > 0 ? c->build_gb_user : (int) (0.875 * bthreads / (per_bin > 0.1 ? per_bin : 0.1) + 0.5);
      tg = tg < 1 ? 1 : (tg > g.nbin[0] ? g.nbin[0] : tg);
      // staging area: the 3 x 3 x (tg + 2) bins of a segment with 8 % head room; a denser segment reports its size
      // through flags[6] and the build is redone with room for it (like the maxneigh retry below)
      const size_t rec_bytes = (bonded || special) ? 20 : 16;   // 4 or 5 fields per staged record
      auto seg_smem = [&](int t, int cap) { return (size_t) cap * rec_bytes + (size_t) (2 * 9 * (t + 3)) * sizeof(int); };
      if (c->seg_cap_tg != tg) { c->seg_cap = 0; c->seg_cap_tg = tg; }
      if (c->seg_cap == 0 && c->seg_cap_user > 0) c->seg_cap = (c->seg_cap_user + 3) & ~3;   // test knob: start too small, exercise the retry
      // (every staged bin is padded to a multiple of vw records: + (vw - 1) / 2 on average)
      const int vw = (per_bin >= 8.0 && !(c->special.on && c->bonds.maxb == 0)) ? 4 : 2;
      // the candidate records: FP32, groups of vw, every bin padded to vw (k_pad_count -> scan -> k_rec_fill)
      const int nf = (bonded || special) ? 5 : 4, lg = vw == 4 ? 2 : 1;
      {
        const int mb = g.mbins;
        const size_t cap_rec = (((size_t) nall + (size_t) (vw - 1) * (size_t) (mb < nall ? mb : nall)) + 63) & ~(size_t) 63;
        TRY(dev_ensure(c, c->pbin_start, (size_t) mb + 1));
        TRY(dev_ensure(c, c->rec, cap_rec * nf));
        // bin_count is free after the sort: it holds the padded populations until the scan has run
        k_pad_count<<<cdiv(mb + 1, 256), 256, 0, c->stream>>>(c->bin_start.p, mb, vw, c->bin_count.p);
        LAUNCH_CHECK();
        TRY(exclusive_scan(c, c->bin_count.p, c->pbin_start.p, mb + 1));
        k_rec_fill<<<cdiv(nall, 256), 256, 0, c->stream>>>(c->pos.p, c->cell_list.p, c->atom_bin.p, c->bin_start.p, c->pbin_start.p, nall,
                                                          g.lo[0] - BIN_MARGIN / g.inv[0], g.lo[1] - BIN_MARGIN / g.inv[1],
                                                          g.lo[2] - BIN_MARGIN / g.inv[2], c->rec.p, lg, nf, c->tag.p);
        LAUNCH_CHECK();
      }
      const RecArrays rec = {c->rec.p, c->pbin_start.p};
      // bonds without an image choice: at most 2 per atom and every periodic dimension wider than 2 cutoffs
      bool unique_images = c->bonds.maxb <= 2 && !c->build_bond_general;
      for (int k = 0; k < 3; k++)
        if (c->periodic[k] && !(c->prd[k] > 2.0 * c->cutneighmax)) unique_images = false;
      // half-width of the band in which the FP64 expression decides: the FP32 error bound + the rounding of rsq - cutoff
      double cutsq_max = 0.0;
      for (int i = 1; i <= c->ntypes; i++)
        for (int j = 1; j <= c->ntypes; j++) cutsq_max = fmax(cutsq_max, c->cutneighsq_h[i * (c->ntypes + 1) + j]);
      const float hw = (float) (tol * 1.0001 + cutsq_max * 2.4e-7);
      if (c->seg_cap == 0) c->seg_cap = (int) (1.08 * 9.0 * (tg + 2) * ((double) nall / (nown > 0 ? nown : 1) + 0.5 * (vw - 1))) + 64;
      if (c->seg_cap > nall + 27 * (tg + 2)) c->seg_cap = nall + 27 * (tg + 2);   // no segment can stage more than exists
      c->seg_cap = (c->seg_cap + 3) & ~3;
      while (tg > 1 && seg_smem(tg, c->seg_cap) > 160 * 1024) {   // never needed for liquids; a guard for pathological density
        tg = (tg + 1) / 2;
        c->seg_cap = (c->seg_cap + 1) / 2 + 64;
        c->seg_cap_tg = tg;
      }
      if (seg_smem(tg, c->seg_cap) > 200 * 1024)
        return b200md_fail(c, B200MD_ERR_NEIGH_OVERFLOW, "neighbor build: %d atoms in the 27 bins around one bin do not fit shared memory", c->seg_cap);
      c->build_gb = tg;
      static bool attr_done[64] = {};  // per device: a process may hold contexts on several GPUs
      if (!attr_done[c->device & 63]) {
        CU(cudaFuncSetAttribute(k_build_seg<1, 0, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<0, 0, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<1, 1, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<0, 1, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<1, 3, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<0, 3, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<1, 0, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<0, 0, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<1, 1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<0, 1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<1, 3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<0, 3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(cudaFuncSetAttribute(k_build_seg<0, 2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_done[c->device & 63] = true;
      }
      const size_t bsmem = seg_smem(tg, c->seg_cap);
#define BB_ARGS c->pos.p, rec, n, g, c->bin_start.p, c->cutneighsq_d.p, c->ntypes, (float) tol, hw, c->maxneigh, c->numneigh.p, c->neigh.p, c->flags_d, tile_ghost, bb, tg, c->seg_cap
#define BB_LAUNCH(OT, BD, V) k_build_seg<OT, BD, V><<<bgrid, bthreads, bsmem, c->stream>>>(BB_ARGS)
      const int bgrid = cdiv(g.nbin[0], tg) * g.nbin[1] * g.nbin[2];
      const int ot = c->ntypes == 1;
      if (special) {   // molecular systems under the stock loop: the multi-type kernel, narrow vectors (small systems)
        BB_LAUNCH(0, 2, 2);
      } else if (vw == 4) {
        if (bonded && unique_images) { if (ot) BB_LAUNCH(1, 3, 4); else BB_LAUNCH(0, 3, 4); }
        else if (bonded) { if (ot) BB_LAUNCH(1, 1, 4); else BB_LAUNCH(0, 1, 4); }
        else { if (ot) BB_LAUNCH(1, 0, 4); else BB_LAUNCH(0, 0, 4); }
      } else {
        if (bonded && unique_images) { if (ot) BB_LAUNCH(1, 3, 2); else BB_LAUNCH(0, 3, 2); }
        else if (bonded) { if (ot) BB_LAUNCH(1, 1, 2); else BB_LAUNCH(0, 1, 2); }
        else { if (ot) BB_LAUNCH(1, 0, 2); else BB_LAUNCH(0, 0, 2); }
      }
#undef BB_LAUNCH
      if (tile_ghost) {   // (the flags are 0 / 1, so their scan counts the ghost tiles before each tile)
        TRY(dev_ensure(c, c->tile_list, (size_t) ntiles));
        TRY(dev_ensure(c, c->tile_pos, (size_t) ntiles));
        TRY(exclusive_scan(c, tile_ghost, c->tile_pos.p, ntiles));
        k_tile_lists<<<cdiv(ntiles, 256), 256, 0, c->stream>>>(tile_ghost, c->tile_pos.p, ntiles, c->tile_list.p, c->flags_d);
        c->launches++;
      }
#undef BB_ARGS
      LAUNCH_CHECK();
    } else {
      if (c->bonds.maxb > 0)
        return b200md_fail(c, B200MD_ERR_ARG, "bonded systems are built by the bin-per-warp kernel only (cluster size 1, FP32 prefilter)");
      c->tiles_valid = false;
      KProf kp(c, KC_NEIGH);
      k_build_fp64<<<cdiv(n, 128), 128, smem, c->stream>>>(c->pos.p, n, c->bg, c->bin_start.p, c->cell_list.p, c->cutneighsq_d.p,
                                                          c->ntypes, c->maxneigh, c->numneigh.p, c->neigh.p, c->flags_d);
      LAUNCH_CHECK();
    }
    CU(cudaMemcpyAsync(c->flags_h, c->flags_d, 8 * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (c->flags_h[0]) return b200md_fail(c, B200MD_ERR_NONFINITE, "Non-numeric positions - simulation unstable");
    if (c->flags_h[5]) return b200md_fail(c, B200MD_ERR_LOST_ATOMS, "Bond atoms missing: a bond partner is not within the neighbor cutoff");
    c->bonds.idx_valid = c->bonds.maxb > 0;
    c->maxneigh_seen = c->flags_h[1];
    c->ntiles_ghost = c->tiles_valid ? c->flags_h[7] : 0;
    if (c->flags_h[6] > c->seg_cap) {   // a segment's candidates did not fit the staging area: make room and redo
      c->seg_cap = (c->flags_h[6] + c->flags_h[6] / 8 + 64 + 3) & ~3;
      CU(cudaMemsetAsync(c->flags_d + 6, 0, sizeof(int), c->stream));
      CU(cudaMemsetAsync(c->flags_d + 1, 0, sizeof(int), c->stream));
      continue;
    }
    // (k_build_seg appends a group of up to 4 entries without a per-entry clamp: its rows are right only while every
    // list stays within maxneigh - 4)
    if (c->flags_h[1] > c->maxneigh - 4) {  // "Neighbor list overflow": grow and redo instead of aborting
      c->maxneigh = (c->flags_h[1] + c->flags_h[1] / 8 + 4 + 7) / 8 * 8;
      CU(cudaMemsetAsync(c->flags_d + 1, 0, sizeof(int), c->stream));
      continue;
    }
    break;
  }
  if (c->flags_h[1] > c->maxneigh - 4) return b200md_fail(c, B200MD_ERR_NEIGH_OVERFLOW, "Neighbor list overflow");

  ph.mark("build");
  if (c->dist_check && !c->xhold_by_permute) {   // (the rebuild's permutation already wrote it; this is the host-ghost entry)
    TRY(dev_ensure(c, c->xhold, (size_t) n));
    CU(cudaMemcpyAsync(c->xhold.p, c->pos.p, (size_t) n * sizeof(double4), cudaMemcpyDeviceToDevice, c->stream));
  }
  c->xhold_by_permute = false;
  ph.mark("xhold");
  c->ago = 0;
  c->nbuilds++;
  c->list_valid = true;
  return B200MD_OK;
}
