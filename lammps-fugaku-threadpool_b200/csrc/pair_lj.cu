// pair_lj.cu -- pair_style lj/cut on the device: PairLJCut::compute (src/pair_lj_cut.cpp:69-139).
// Full neighbor list, one thread per owned atom, no atomics and no ghost-force reverse_comm; the
// force of atom i is written once (this also replaces Verlet::force_clear, src/verlet.cpp:376-421).
// Energy / virial (EVFLAG) follow Pair::ev_tally_full (src/pair.cpp:1115-1160): half of each pair term
// per visit, reduced per block and then in a fixed order (reproducible run to run).
#include "ctx.cuh"

struct LJOne { double cutsq, lj1, lj2, lj3, lj4, offset; };

__device__ __forceinline__ double4 ld_pos(const double4 *__restrict__ pos, int j)
{
  const double2 *q = reinterpret_cast<const double2 *>(pos + j);
  double2 a = __ldg(q), b = __ldg(q + 1);
  return make_double4(a.x, a.y, b.x, b.y);
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

template <int EVFLAG, int ONETYPE>
__global__ void __launch_bounds__(128)
k_lj_full(const double4 *__restrict__ pos, const int *__restrict__ numneigh, const int *__restrict__ neigh, int maxneigh,
          int nlocal, LJOne one, const double *__restrict__ tables, int ntypes, double *__restrict__ fx,
          double *__restrict__ fy, double *__restrict__ fz, double *__restrict__ partials)
{
  extern __shared__ double s_tab[];  // 6 tables of (ntypes+1)^2 when !ONETYPE
  const int nt1 = ntypes + 1, nt2 = nt1 * nt1;
  if (!ONETYPE) {
    for (int k = threadIdx.x; k < 6 * nt2; k += blockDim.x) s_tab[k] = tables[k];
    __syncthreads();
  }
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double fxi = 0.0, fyi = 0.0, fzi = 0.0;
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  if (i < nlocal) {
    const double4 pi = pos[i];
    const int n = numneigh[i];
    const int *nb = neigh + (size_t) (i >> 5) * TILE * maxneigh + (i & 31);
    const int trow = ONETYPE ? 0 : pos_type(pi) * nt1;
#pragma unroll 4
    for (int k = 0; k < n; k++) {
      const int j = __ldcs(nb + (size_t) k * TILE) & NEIGHMASK;
      const double4 pj = ld_pos(pos, j);
      const double dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
      const double rsq = dx * dx + dy * dy + dz * dz;
      double cutsq, lj1, lj2;
      if (ONETYPE) {
        cutsq = one.cutsq; lj1 = one.lj1; lj2 = one.lj2;
      } else {
        const int t = trow + pos_type(pj);
        cutsq = s_tab[t]; lj1 = s_tab[nt2 + t]; lj2 = s_tab[2 * nt2 + t];
      }
      if (rsq < cutsq) {
        const double r2inv = 1.0 / rsq;
        const double r6inv = r2inv * r2inv * r2inv;
        const double forcelj = r6inv * (lj1 * r6inv - lj2);
        const double fpair = forcelj * r2inv;
        fxi += dx * fpair;
        fyi += dy * fpair;
        fzi += dz * fpair;
        if (EVFLAG) {
          double lj3, lj4, off;
          if (ONETYPE) {
            lj3 = one.lj3; lj4 = one.lj4; off = one.offset;
          } else {
            const int t = trow + pos_type(pj);
            lj3 = s_tab[3 * nt2 + t]; lj4 = s_tab[4 * nt2 + t]; off = s_tab[5 * nt2 + t];
          }
          acc[0] += 0.5 * (r6inv * (lj3 * r6inv - lj4) - off);
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
  const int onetype = (c->lj.ntypes == 1);
  size_t smem = onetype ? 0 : (size_t) 6 * (c->lj.ntypes + 1) * (c->lj.ntypes + 1) * sizeof(double);
  {
  KProf kp(c, KC_PAIR);
#define LJ_ARGS c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, one, c->lj.d, c->lj.ntypes, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p
  if (ev) {
    if (onetype) k_lj_full<1, 1><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    else k_lj_full<1, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
  } else {
    if (onetype) k_lj_full<0, 1><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
    else k_lj_full<0, 0><<<blocks, threads, smem, c->stream>>>(LJ_ARGS);
  }
#undef LJ_ARGS
  LAUNCH_CHECK();
  }
  if (ev) TRY(reduce_partials(c, blocks, 7, 0));
  return B200MD_OK;
}
