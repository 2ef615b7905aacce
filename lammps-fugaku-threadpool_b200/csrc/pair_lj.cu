// pair_lj.cu -- pair_style lj/cut on the device: PairLJCut::compute (src/pair_lj_cut.cpp:69-139).
// Full neighbor list, one thread per owned atom, no atomics and no ghost-force reverse_comm; the
// force of atom i is written once (this also replaces Verlet::force_clear, src/verlet.cpp:376-421).
// Energy / virial (EVFLAG) follow Pair::ev_tally_full (src/pair.cpp:1115-1160): half of each pair term
// per visit, reduced per block and then in a fixed order (reproducible run to run).
#include "ctx.cuh"
#include "bond_dev.cuh"

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

// The force kernel: one thread per owned atom walks its list; x_j is one 32-byte LDG.256 gather through L1.
// What bounds it (ncu, profiles/r02_ncu_lj.md): the L1 data pipe at 95 % of its wavefront peak -- one wavefront per
// distinct 128-byte line of a warp-wide gather, 21.8 lines per warp step on a melt (tools/line_model.py reproduces
// the counter), against 8.5 cycles of FP64 issue per step.  Measured and dropped: lists per
// cluster of 2/4 atoms or lanes, shared-memory staging of the bin neighbourhood with 16-bit slot lists (the random
// LDS.64 gathers conflict 2.4-way: 15.9 wavefronts per step, and occupancy falls to 26 %), register tiles with
// indexed shuffles (25-31 % lane efficiency), a Newton-3 half list (no native 64-bit shared-memory atomic add on
// sm_100: ATOMS.CAST.SPIN loops at 104 cycles per warp step).
// SPLIT 1 (overlap mode): warp w works on the 32-atom tile tile_list[w], w < part -- the launch covers either the
// ghost-free tiles (they run while the halo is still in flight) or the others, each as a compact list made at the
// rebuild, so neither launch carries idle blocks.  SPLIT 2 (host-buffer pipeline): the atom range [part, nlocal).
// INTEG (steps that neither tally nor end a run, fix nve alone): the thread that holds f_i also does what follows it in
// the timestep -- FixNVE::final_integrate of this step and initial_integrate of the next (src/fix_nve.cpp:64-141; the
// operations and their order are k_nve_final_initial's, nve.cu) -- and writes the new position into the second
// position buffer (other threads still gather the old ones).  f is not stored; the streaming pass over x, v, f of the
// separate kernel (136 B per atom) is gone.  (NVEFuse, nve_fused_update: ctx.cuh)
template <int EVFLAG, int ONETYPE, int SPLIT, int INTEG>
__global__ void __launch_bounds__(128)
k_lj_force(const double4 *__restrict__ pos, const int *__restrict__ numneigh, const int *__restrict__ neigh, int maxneigh,
           int nlocal, LJOne one, const double *__restrict__ tables, int ntypes, double *__restrict__ fx,
           double *__restrict__ fy, double *__restrict__ fz, double *__restrict__ partials,
           const int *__restrict__ tile_list, int part, NVEFuse nve, ChainPost cp)
{
  extern __shared__ double s_tab[];   // 6 tables of (ntypes+1)^2, then special_lj[4] ([0] = 1), when !ONETYPE
  const int nt1 = ntypes + 1, nt2 = nt1 * nt1;
  if (!ONETYPE) {
    for (int k = threadIdx.x; k < 6 * nt2 + 4; k += blockDim.x) s_tab[k] = tables[k];
    __syncthreads();
  }
  // SPLIT == 2 (host-buffer pipeline): the atom range [part, nlocal), part a multiple of the block size
  int i = (SPLIT == 2 ? part : 0) + blockIdx.x * blockDim.x + threadIdx.x;
  if (SPLIT == 1) {
    const int w = i >> 5;
    i = w < part ? tile_list[w] * TILE + (threadIdx.x & 31) : nlocal;
  }
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  const bool active = i < nlocal;
  if (active) {
    const double4 pi = pos[i];
    const int trow = ONETYPE ? 0 : pos_type(pi) * nt1;
    const int n = numneigh[i];
    const int *nb = neigh + (size_t) (i >> 5) * TILE * maxneigh + (i & 31);
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
        const bool in = rsq < cutsq && (k0 + u < n);
        // factor_lj of src/pair_lj_cut.cpp:103-105: the special-bond class sits in bits 28-29 of the entry (0 = ordinary
        // neighbor, weight 1); single-type systems are atomic, their entries carry no class
        const double factor_lj = ONETYPE ? 1.0 : s_tab[6 * nt2 + ((ent[u] >> CLBITS) & 3u)];
        const double r2inv = rcp_pos(rsq);
        const double r6inv = r2inv * r2inv * r2inv;
        const double forcelj = r6inv * (lj1 * r6inv - lj2);
        const double fpair = in ? factor_lj * forcelj * r2inv : 0.0;
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
          acc[0] += in ? 0.5 * factor_lj * (r6inv * (lj3 * r6inv - lj4) - off) : 0.0;
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
    if (INTEG == 2) {   // in.chain: bonds and thermostat kick of atom i, then fix nve, all on the force in registers
      chain_post_atom(cp, nve, pos, i, pi, fxi, fyi, fzi);
    } else if (INTEG) {
      nve_fused_update(nve, i, pi, fxi, fyi, fzi);
    } else {
      fx[i] = fxi;
      fy[i] = fyi;
      fz[i] = fzi;
    }
  }
  if (EVFLAG) block_partials<7>(acc, partials);
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

int pair_lj_compute(b200md_ctx *c, int eflag, int vflag)
{
  const int n = c->nlocal;
  const int threads = 128, blocks = cdiv(n, threads);
  const int ev = (eflag || vflag) ? 1 : 0;
  if (ev) TRY(dev_ensure(c, c->partials, (size_t) blocks * 8));
  LJOne one = {c->lj.h_onetype[0], c->lj.h_onetype[1], c->lj.h_onetype[2], c->lj.h_onetype[3], c->lj.h_onetype[4], c->lj.h_onetype[5]};
  const int onetype = (c->lj.ntypes == 1) && !c->special.on;
  size_t smem = onetype ? 0 : ((size_t) 6 * (c->lj.ntypes + 1) * (c->lj.ntypes + 1) + 4) * sizeof(double);
  NVEFuse nve = {};
  ChainPost cp = {};
  const bool integ = !ev && nve_fuse_args(c, nve);
  if (integ && c->fuse.chain) TRY(chain_post_args(c, c->ntimestep, cp));
  {
    KProf kp(c, KC_PAIR);
#define LJ_ARGS c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, one, c->lj.d, c->lj.ntypes, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p, nullptr, -1, nve, cp
    if (ev) {
      if (onetype) k_lj_force<1, 1, 0, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
      else k_lj_force<1, 0, 0, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    } else if (integ && c->fuse.chain) {
      if (onetype) k_lj_force<0, 1, 0, 2><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
      else k_lj_force<0, 0, 0, 2><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
      c->fuse.done = true;
    } else if (integ) {
      if (onetype) k_lj_force<0, 1, 0, 1><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
      else k_lj_force<0, 0, 0, 1><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
      c->fuse.done = true;
    } else {
      if (onetype) k_lj_force<0, 1, 0, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
      else k_lj_force<0, 0, 0, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    }
#undef LJ_ARGS
    LAUNCH_CHECK();
  }
  if (ev) TRY(reduce_partials(c, blocks, 7, 0));
  return B200MD_OK;
}

// overlap mode (capi.cu): forces of the ghost-free tiles (part 0) or of the others (part 1), no energy/virial tally
int pair_lj_compute_part(b200md_ctx *c, int part)
{
  const int n = c->nlocal;
  const int ntiles = cdiv(n, TILE), cnt = part ? c->ntiles_ghost : ntiles - c->ntiles_ghost;
  if (cnt == 0) return B200MD_OK;
  const int *list = c->tile_list.p + (part ? ntiles - c->ntiles_ghost : 0);
  const int threads = 128, blocks = cdiv(cnt * TILE, threads);
  LJOne one = {c->lj.h_onetype[0], c->lj.h_onetype[1], c->lj.h_onetype[2], c->lj.h_onetype[3], c->lj.h_onetype[4], c->lj.h_onetype[5]};
  const int onetype = (c->lj.ntypes == 1) && !c->special.on;
  size_t smem = onetype ? 0 : ((size_t) 6 * (c->lj.ntypes + 1) * (c->lj.ntypes + 1) + 4) * sizeof(double);
  NVEFuse nve = {};
  ChainPost cp = {};
  const bool integ = nve_fuse_args(c, nve);
  if (integ && c->fuse.chain) TRY(chain_post_args(c, c->ntimestep, cp));
  KProf kp(c, KC_PAIR);
#define LJ_ARGS c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, one, c->lj.d, c->lj.ntypes, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p, list, cnt, nve, cp
  if (integ && c->fuse.chain) {
    if (onetype) k_lj_force<0, 1, 1, 2><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    else k_lj_force<0, 0, 1, 2><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    c->fuse.done = true;
  } else if (integ) {
    if (onetype) k_lj_force<0, 1, 1, 1><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    else k_lj_force<0, 0, 1, 1><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    c->fuse.done = true;
  } else {
    if (onetype) k_lj_force<0, 1, 1, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    else k_lj_force<0, 0, 1, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
  }
#undef LJ_ARGS
  LAUNCH_CHECK();
  return B200MD_OK;
}

// host-buffer pipeline (capi.cu): forces of the owned atoms [first, first + count) only, no energy/virial tally, so
// the download of one range's results runs under the force kernel of the next
bool pair_lj_range_usable(const b200md_ctx *c) { return c->pair_style == 1; }

int pair_lj_compute_range(b200md_ctx *c, int first, int count)
{
  const int threads = 128, blocks = cdiv(count, threads);
  if (first % threads) return b200md_fail(c, B200MD_ERR_ARG, "pair range must start at a multiple of %d", threads);
  LJOne one = {c->lj.h_onetype[0], c->lj.h_onetype[1], c->lj.h_onetype[2], c->lj.h_onetype[3], c->lj.h_onetype[4], c->lj.h_onetype[5]};
  const int onetype = (c->lj.ntypes == 1) && !c->special.on;
  size_t smem = onetype ? 0 : ((size_t) 6 * (c->lj.ntypes + 1) * (c->lj.ntypes + 1) + 4) * sizeof(double);
  KProf kp(c, KC_PAIR);
#define LJ_ARGS c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, first + count, one, c->lj.d, c->lj.ntypes, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p, nullptr, first, NVEFuse{}, ChainPost{}
  if (onetype) k_lj_force<0, 1, 2, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
  else k_lj_force<0, 0, 2, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
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
    const unsigned ent = (unsigned) __ldg(nb + (size_t) k * TILE);
    const int j = (int) (ent & CLJMASK);
    const double factor_lj = __ldg(tables + 6 * nt2 + ((ent >> CLBITS) & 3u));
    const double4 pj = pos[j];
    const double dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
    const double rsq = dx * dx + dy * dy + dz * dz;
    const int t = trow + pos_type(pj);
    if (rsq >= __ldg(tables + t)) continue;
    const double r2inv = 1.0 / rsq, r6inv = r2inv * r2inv * r2inv;
    const double fpair = factor_lj * r6inv * (__ldg(tables + nt2 + t) * r6inv - __ldg(tables + 2 * nt2 + t)) * r2inv;
    acc[0] += 0.5 * factor_lj * (r6inv * (__ldg(tables + 3 * nt2 + t) * r6inv - __ldg(tables + 4 * nt2 + t)) - __ldg(tables + 5 * nt2 + t));
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

