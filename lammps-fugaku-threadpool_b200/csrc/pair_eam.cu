// pair_eam.cu -- pair_style eam on the device: PairEAM::compute (src/MANYBODY/pair_eam.cpp:142-328).
// The reference runs rho (half list + reverse_comm) -> fp -> forward_comm(fp) -> force (half list +
// reverse_comm of f).  Here a FULL list makes rho_i and f_i complete on the owning thread, so the two
// reverse_comms disappear; the sequence is
//   k_eam_rho_fp : rho_i = sum_j rhor(r_ij), then fp_i = F'(rho_i) and the embedding energy, fused
//                  (pair_eam.cpp:187-247)
//   halo of fp   : ghost fp from their owners (pair_eam.cpp:251, pack/unpack :855-876)
//   k_eam_force  : pair_eam.cpp:257-325
// Spline rows are the reference's 7 coefficients per knot ([0..2] derivative, [3..6] value;
// PairEAM::interpolate :769-797), 1-based knots m = 1..n, re-packed at init into 32-byte aligned pieces:
//   rhoA[t][m]  = {c3,c4,c5,c6} of rhor_spline            (pass A reads one double4)
//   tabC[t][m]  = {r0,r1,r2,-} {z0,z1,z2,z3} {z4,z5,z6,-}  only for the single-type fast path
#include "ctx.cuh"

__device__ __forceinline__ double4 ld4(const double4 *__restrict__ p)
{
  const double2 *q = reinterpret_cast<const double2 *>(p);
  double2 a = __ldg(q), b = __ldg(q + 1);
  return make_double4(a.x, a.y, b.x, b.y);
}

__device__ __forceinline__ double4 ld_pos_e(const double4 *__restrict__ pos, int j) { return ld_pos256(pos + j); }

__device__ __forceinline__ double warp_sum_e(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}

template <int NC> __device__ __forceinline__ void block_partials_e(double (&acc)[NC], double *__restrict__ partials)
{
  __shared__ double s_red[32][NC];
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int c = 0; c < NC; c++) {
    double v = warp_sum_e(acc[c]);
    if (lane == 0) s_red[w][c] = v;
  }
  __syncthreads();
  if (threadIdx.x < NC) {
    double t = 0.0;
    for (int k = 0; k < nw; k++) t += s_red[k][threadIdx.x];
    partials[(size_t) blockIdx.x * 8 + threadIdx.x] = t;
  }
}

struct EAMParams {
  int ntypes, nrho, nr;
  double rdrho, rdr, rhomax, cutforcesq;
  const double *frho;      // [nfrho][nrho+1][7]
  const double *rhor;      // [nrhor][nr+1][7]
  const double *z2r;       // [nz2r][nr+1][7]
  const double4 *rhoA;     // [nrhor][nr+1]
  const int *type2frho, *type2rhor, *type2z2r;
};

template <int EFLAG, int ONETYPE>
__global__ void __launch_bounds__(128)
k_eam_rho_fp(const double4 *__restrict__ pos, const int *__restrict__ numneigh, const int *__restrict__ neigh, int maxneigh,
             int nlocal, EAMParams P, double *__restrict__ rho_out, double *__restrict__ fp_out, double *__restrict__ partials)
{
  constexpr int CL = 1;   // one list per atom (lists per cluster of 2 / 4 atoms measured slower)
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = q * CL;
  double acc[1] = {0.0};
  if (i0 < nlocal) {
    const int nt1 = P.ntypes + 1;
    double xi[CL], yi[CL], zi[CL], rho[CL];
    int it[CL];
#pragma unroll
    for (int a = 0; a < CL; a++) {
      rho[a] = 0.0;
      if (i0 + a < nlocal) {
        const double4 p = pos[i0 + a];
        xi[a] = p.x; yi[a] = p.y; zi[a] = p.z;
        it[a] = pos_type(p);
      } else {
        xi[a] = yi[a] = zi[a] = 1e150;
        it[a] = 1;
      }
    }
    const int n = numneigh[q];
    const int *nb = neigh + (size_t) (q >> 5) * TILE * maxneigh + (q & 31);
#pragma unroll 2
    for (int k = 0; k < n; k++) {
      const unsigned ent = (unsigned) __ldcs(nb + (size_t) k * TILE);
      const int j = (int) (ent & CLJMASK);
      const double4 pj = ld_pos_e(pos, j);
      const int jt = pos_type(pj);
#pragma unroll
      for (int a = 0; a < CL; a++) {
        const double dx = xi[a] - pj.x, dy = yi[a] - pj.y, dz = zi[a] - pj.z;
        const double rsq = dx * dx + dy * dy + dz * dz;
        if (rsq < P.cutforcesq) {
          double p = sqrt(rsq) * P.rdr + 1.0;
          int m = (int) p;
          m = min(m, P.nr - 1);
          p -= m;
          p = fmin(p, 1.0);
          const int tab = ONETYPE ? 0 : P.type2rhor[jt * nt1 + it[a]];
          const double4 cf = ld4(P.rhoA + (size_t) tab * (P.nr + 1) + m);
          rho[a] += ((cf.x * p + cf.y) * p + cf.z) * p + cf.w;
        }
      }
    }
    // pass B (pair_eam.cpp:231-247)
#pragma unroll
    for (int a = 0; a < CL; a++) {
      if (i0 + a >= nlocal) continue;
      double p = rho[a] * P.rdrho + 1.0;
      int m = (int) p;
      m = max(1, min(m, P.nrho - 1));
      p -= m;
      p = fmin(p, 1.0);
      const double *cf = P.frho + ((size_t) (ONETYPE ? 0 : P.type2frho[it[a]]) * (P.nrho + 1) + m) * 7;
      const double fpi = (__ldg(cf) * p + __ldg(cf + 1)) * p + __ldg(cf + 2);
      rho_out[i0 + a] = rho[a];
      fp_out[i0 + a] = fpi;
      if (EFLAG) {
        double phi = ((__ldg(cf + 3) * p + __ldg(cf + 4)) * p + __ldg(cf + 5)) * p + __ldg(cf + 6);
        if (rho[a] > P.rhomax) phi += fpi * (rho[a] - P.rhomax);
        acc[0] += phi;
      }
    }
  }
  if (EFLAG) block_partials_e<1>(acc, partials);
}

template <int EVFLAG, int ONETYPE>
__global__ void __launch_bounds__(128)
k_eam_force(const double4 *__restrict__ pos, const int *__restrict__ numneigh, const int *__restrict__ neigh, int maxneigh,
            int nlocal, EAMParams P, const double *__restrict__ fp, double *__restrict__ fx, double *__restrict__ fy,
            double *__restrict__ fz, double *__restrict__ partials)
{
  constexpr int CL = 1;   // one list per atom (lists per cluster of 2 / 4 atoms measured slower)
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = q * CL;
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  if (i0 < nlocal) {
    const int nt1 = P.ntypes + 1;
    double xi[CL], yi[CL], zi[CL], fpi[CL], fxi[CL], fyi[CL], fzi[CL];
    int it[CL];
#pragma unroll
    for (int a = 0; a < CL; a++) {
      fxi[a] = fyi[a] = fzi[a] = 0.0;
      if (i0 + a < nlocal) {
        const double4 p = pos[i0 + a];
        xi[a] = p.x; yi[a] = p.y; zi[a] = p.z;
        it[a] = pos_type(p);
        fpi[a] = fp[i0 + a];
      } else {
        xi[a] = yi[a] = zi[a] = 1e150;
        it[a] = 1;
        fpi[a] = 0.0;
      }
    }
    const int n = numneigh[q];
    const int *nb = neigh + (size_t) (q >> 5) * TILE * maxneigh + (q & 31);
    for (int k = 0; k < n; k++) {
      const unsigned ent = (unsigned) __ldcs(nb + (size_t) k * TILE);
      const int j = (int) (ent & CLJMASK);
      const double4 pj = ld_pos_e(pos, j);
      const double fpj = __ldg(fp + j);
      const int jt = pos_type(pj);
#pragma unroll
      for (int a = 0; a < CL; a++) {
        const double dx = xi[a] - pj.x, dy = yi[a] - pj.y, dz = zi[a] - pj.z;
        const double rsq = dx * dx + dy * dy + dz * dz;
        if (rsq < P.cutforcesq) {
          const double r = sqrt(rsq);
          double p = r * P.rdr + 1.0;
          int m = (int) p;
          m = min(m, P.nr - 1);
          p -= m;
          p = fmin(p, 1.0);
          int t_ij = 0, t_ji = 0, t_z = 0;
          if (!ONETYPE) {
            t_ij = P.type2rhor[it[a] * nt1 + jt];
            t_ji = P.type2rhor[jt * nt1 + it[a]];
            t_z = P.type2z2r[it[a] * nt1 + jt];
          }
          const double *c1 = P.rhor + ((size_t) t_ij * (P.nr + 1) + m) * 7;
          const double rhoip = (__ldg(c1) * p + __ldg(c1 + 1)) * p + __ldg(c1 + 2);
          double rhojp = rhoip;
          if (!ONETYPE) {
            const double *c2 = P.rhor + ((size_t) t_ji * (P.nr + 1) + m) * 7;
            rhojp = (__ldg(c2) * p + __ldg(c2 + 1)) * p + __ldg(c2 + 2);
          }
          const double *c3 = P.z2r + ((size_t) t_z * (P.nr + 1) + m) * 7;
          const double z2p = (__ldg(c3) * p + __ldg(c3 + 1)) * p + __ldg(c3 + 2);
          const double z2 = ((__ldg(c3 + 3) * p + __ldg(c3 + 4)) * p + __ldg(c3 + 5)) * p + __ldg(c3 + 6);
          const double recip = 1.0 / r;
          const double phi = z2 * recip;
          const double phip = z2p * recip - phi * recip;
          const double psip = fpi[a] * rhojp + fpj * rhoip + phip;
          const double fpair = -psip * recip;
          fxi[a] += dx * fpair;
          fyi[a] += dy * fpair;
          fzi[a] += dz * fpair;
          if (EVFLAG) {
            acc[0] += 0.5 * phi;
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
    }
#pragma unroll
    for (int a = 0; a < CL; a++)
      if (i0 + a < nlocal) {
        fx[i0 + a] = fxi[a];
        fy[i0 + a] = fyi[a];
        fz[i0 + a] = fzi[a];
      }
  }
  if (EVFLAG) block_partials_e<7>(acc, partials);
}

// ------------------------------------------------------------------------------------------------
// single-type fast path (the Cu benchmark): per-atom list, spline coefficients staged in shared memory as SoA
// rows (random knots then spread over the banks and cost no L1 tag look-ups), r and 1/r from one rsqrt, and the
// force pass gathers ONE 32-byte record {x,y,z,fp} per neighbor instead of position + fp[j] + 10 table words.
// ncu on the generic kernels: 7 L1 sectors per pair in the force pass, LSU data pipe at 84 %.
// ------------------------------------------------------------------------------------------------
// one table copy per 512-thread block: fewer, larger blocks leave more of the 228 KB for L1 (ncu, 256-thread blocks:
// L1 hit rate 17 % in the rho pass with 8 table copies per SM)
#define EAM1T_THREADS 512
template <int EFLAG, int SPLIT>
__global__ void __launch_bounds__(EAM1T_THREADS, 2)
k_eam_rho_fp_1t(const double4 *__restrict__ pos, const int *__restrict__ numneigh, const int *__restrict__ neigh, int maxneigh,
                int nlocal, EAMParams P, const double *__restrict__ tab1, double *__restrict__ rho_out,
                double *__restrict__ fp_out, double4 *__restrict__ posfp, double *__restrict__ partials,
                const int *__restrict__ tile_list, int part, int *__restrict__ numneigh_in, int *__restrict__ neigh_in)
{
  extern __shared__ double s_t[];  // [4][n1]: rhor c3..c6
  const int n1 = P.nr + 1;
  for (int k = threadIdx.x; k < 4 * n1; k += blockDim.x) s_t[k] = tab1[k];
  __syncthreads();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (SPLIT == 1) {   // overlap mode, see k_lj_force: warp w takes tile tile_list[w], w < part
    const int w = i >> 5;
    i = w < part ? tile_list[w] * TILE + (threadIdx.x & 31) : nlocal;
  }
  double acc[1] = {0.0};
  const bool active = i < nlocal;
  if (active) {
    const double4 pi = pos[i];
    const int n = numneigh[i];
    const int *nb = neigh + (size_t) (i >> 5) * TILE * maxneigh + (i & 31);
    double rho = 0.0;
    // the entries inside the force cutoff (58 % of the Cu melt's list: cutforce 4.95 of cutneigh 5.95), in list order, for
    // the force pass of this step: it then gathers and evaluates those alone
    int *const w0 = neigh_in + (size_t) (i >> 5) * TILE * maxneigh + (i & 31);
    int *wp = w0;
    // software pipeline (see k_lj_force): next batch of list entries in flight while PF gathers are issued together
    constexpr int PF = 4;
    const int nlast = max(n - 1, 0);
    unsigned enx[PF], cur[PF];
#pragma unroll
    for (int u = 0; u < PF; u++) enx[u] = (unsigned) __ldcs(nb + (size_t) min(u, nlast) * TILE);
    for (int k0 = 0; k0 < n; k0 += PF) {
      double4 pj[PF];
#pragma unroll
      for (int u = 0; u < PF; u++) pj[u] = ld_pos256(pos + (enx[u] & CLJMASK));
#pragma unroll
      for (int u = 0; u < PF; u++) {
        cur[u] = enx[u];
        enx[u] = (unsigned) __ldcs(nb + (size_t) min(k0 + PF + u, nlast) * TILE);
      }
#pragma unroll
      for (int u = 0; u < PF; u++) {
        const double dx = pi.x - pj[u].x, dy = pi.y - pj[u].y, dz = pi.z - pj[u].z;
        const double rsq = dx * dx + dy * dy + dz * dz;
        if (rsq < P.cutforcesq && k0 + u < n) {
          *wp = (int) cur[u];
          wp += TILE;
          double p = rsq * rsqrt(rsq) * P.rdr + 1.0;
          int m = (int) p;
          m = min(m, P.nr - 1);
          p -= m;
          p = fmin(p, 1.0);
          rho += ((s_t[m] * p + s_t[n1 + m]) * p + s_t[2 * n1 + m]) * p + s_t[3 * n1 + m];
        }
      }
    }
    numneigh_in[i] = (int) ((wp - w0) / TILE);
    double p = rho * P.rdrho + 1.0;
    int m = (int) p;
    m = max(1, min(m, P.nrho - 1));
    p -= m;
    p = fmin(p, 1.0);
    const double *cf = P.frho + (size_t) m * 7;
    const double fpi = (__ldg(cf) * p + __ldg(cf + 1)) * p + __ldg(cf + 2);
    rho_out[i] = rho;
    fp_out[i] = fpi;
    posfp[i] = make_double4(pi.x, pi.y, pi.z, fpi);
    if (EFLAG) {
      double phi = ((__ldg(cf + 3) * p + __ldg(cf + 4)) * p + __ldg(cf + 5)) * p + __ldg(cf + 6);
      if (rho > P.rhomax) phi += fpi * (rho - P.rhomax);
      acc[0] = phi;
    }
  }
  if (EFLAG) block_partials_e<1>(acc, partials);
}

// numneigh / neigh here are the IN-RANGE lists the rho pass of the same step left (numneigh_in, neigh_in): a subsequence of
// the full list in the same order, so the sums run over the same terms in the same order as over the full list
// INTEG: fix nve fused behind the force of atom i, see k_lj_force (pos = the position records {x, y, z, type})
template <int EVFLAG, int SPLIT, int INTEG>
__global__ void __launch_bounds__(EAM1T_THREADS, 2)
k_eam_force_1t(const double4 *__restrict__ posfp, const int *__restrict__ numneigh, const int *__restrict__ neigh, int maxneigh,
               int nlocal, EAMParams P, const double *__restrict__ tab1, double *__restrict__ fx, double *__restrict__ fy,
               double *__restrict__ fz, double *__restrict__ partials, const int *__restrict__ tile_list, int part,
               const double4 *__restrict__ pos, NVEFuse nve)
{
  extern __shared__ double s_t[];  // [7][n1]: rhor c0..c2 (derivative), z2r c3..c6 (value; derivative derived)
  const int n1 = P.nr + 1;
  for (int k = threadIdx.x; k < 7 * n1; k += blockDim.x) s_t[k] = tab1[4 * n1 + k];
  __syncthreads();
  int i = (SPLIT == 2 ? part : 0) + blockIdx.x * blockDim.x + threadIdx.x;  // 2: range mode, see k_lj_force
  if (SPLIT == 1) {
    const int w = i >> 5;
    i = w < part ? tile_list[w] * TILE + (threadIdx.x & 31) : nlocal;
  }
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  const bool active = i < nlocal;
  if (active) {
    const double4 pi = posfp[i];
    const int n = numneigh[i];
    const int *nb = neigh + (size_t) (i >> 5) * TILE * maxneigh + (i & 31);
    double fxi = 0.0, fyi = 0.0, fzi = 0.0;
    constexpr int PF = 4;
    const int nlast = max(n - 1, 0);
    unsigned enx[PF];
#pragma unroll
    for (int u = 0; u < PF; u++) enx[u] = (unsigned) __ldcs(nb + (size_t) min(u, nlast) * TILE);
    for (int k0 = 0; k0 < n; k0 += PF) {
      double4 pjv[PF];
#pragma unroll
      for (int u = 0; u < PF; u++) pjv[u] = ld_pos256(posfp + (enx[u] & CLJMASK));
#pragma unroll
      for (int u = 0; u < PF; u++) enx[u] = (unsigned) __ldcs(nb + (size_t) min(k0 + PF + u, nlast) * TILE);
#pragma unroll
      for (int u = 0; u < PF; u++) {
        const double4 pj = pjv[u];
        const double dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
        const double rsq = dx * dx + dy * dy + dz * dz;
        if (rsq < P.cutforcesq && k0 + u < n) {
          const double recip = rsqrt(rsq);
          const double r = rsq * recip;
          double p = r * P.rdr + 1.0;
          int m = (int) p;
          m = min(m, P.nr - 1);
          p -= m;
          p = fmin(p, 1.0);
          const double rhoip = (s_t[m] * p + s_t[n1 + m]) * p + s_t[2 * n1 + m];
          const double z3 = s_t[3 * n1 + m], z4 = s_t[4 * n1 + m], z5 = s_t[5 * n1 + m], z6 = s_t[6 * n1 + m];
          const double z2 = ((z3 * p + z4) * p + z5) * p + z6;
          const double z2p = ((3.0 * z3 * p + 2.0 * z4) * p + z5) * P.rdr;  // = spline[0..2] of array2spline (:752-754)
          const double phi = z2 * recip;
          const double phip = z2p * recip - phi * recip;
          const double psip = (pi.w + pj.w) * rhoip + phip;  // one type: rhojp == rhoip
          const double fpair = -psip * recip;
          fxi += dx * fpair;
          fyi += dy * fpair;
          fzi += dz * fpair;
          if (EVFLAG) {
            acc[0] += 0.5 * phi;
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
    }
    if (INTEG) {
      nve_fused_update(nve, i, pos[i], fxi, fyi, fzi);
    } else {
      fx[i] = fxi;
      fy[i] = fyi;
      fz[i] = fzi;
    }
  }
  if (EVFLAG) block_partials_e<7>(acc, partials);
}

static void launch_eam_rho(b200md_ctx *c, int blocks, int threads, int ev, int onetype, const EAMParams &P, int n)
{
#define A_ARGS c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, P, c->rho.p, c->fp.p, c->partials.p
  if (ev) {
    if (onetype) k_eam_rho_fp<1, 1><<<blocks, threads, 0, c->stream>>>(A_ARGS);
    else k_eam_rho_fp<1, 0><<<blocks, threads, 0, c->stream>>>(A_ARGS);
  } else {
    if (onetype) k_eam_rho_fp<0, 1><<<blocks, threads, 0, c->stream>>>(A_ARGS);
    else k_eam_rho_fp<0, 0><<<blocks, threads, 0, c->stream>>>(A_ARGS);
  }
#undef A_ARGS
}

static void launch_eam_force(b200md_ctx *c, int blocks, int threads, int ev, int onetype, const EAMParams &P, int n)
{
#define C_ARGS c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, P, c->fp.p, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p
  if (ev) {
    if (onetype) k_eam_force<1, 1><<<blocks, threads, 0, c->stream>>>(C_ARGS);
    else k_eam_force<1, 0><<<blocks, threads, 0, c->stream>>>(C_ARGS);
  } else {
    if (onetype) k_eam_force<0, 1><<<blocks, threads, 0, c->stream>>>(C_ARGS);
    else k_eam_force<0, 0><<<blocks, threads, 0, c->stream>>>(C_ARGS);
  }
#undef C_ARGS
}

static bool eam_use_1t(const b200md_ctx *c)
{
  const EAMTables &T = c->eam;
  const int onetype = (T.ntypes == 1 && T.nrhor == 1 && T.nz2r == 1);
  return onetype && (size_t) 7 * (T.nr + 1) * 8 <= 96 * 1024;
}

static EAMParams eam_params(const b200md_ctx *c)
{
  const EAMTables &T = c->eam;
  EAMParams P = {T.ntypes, T.nrho, T.nr, T.rdrho, T.rdr, T.rhomax, T.cutforcesq, T.frho, T.rhor, T.z2r, T.rhoA,
                 T.type2frho, T.type2rhor, T.type2z2r};
  return P;
}

static int eam_attrs_1t(b200md_ctx *c)
{
  static bool attr_done[64] = {};  // per device: a process may hold contexts on several GPUs
  if (!attr_done[c->device & 63]) {
    CU(cudaFuncSetAttribute(k_eam_force_1t<0, 0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CU(cudaFuncSetAttribute(k_eam_force_1t<1, 0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CU(cudaFuncSetAttribute(k_eam_force_1t<0, 0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CU(cudaFuncSetAttribute(k_eam_force_1t<0, 1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CU(cudaFuncSetAttribute(k_eam_rho_fp_1t<0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CU(cudaFuncSetAttribute(k_eam_rho_fp_1t<1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CU(cudaFuncSetAttribute(k_eam_rho_fp_1t<0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CU(cudaFuncSetAttribute(k_eam_force_1t<0, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CU(cudaFuncSetAttribute(k_eam_force_1t<0, 2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    attr_done[c->device & 63] = true;
  }
  return B200MD_OK;
}

// the in-range lists of the single-type path: same shape as the full list (sized with it, so a grown maxneigh regrows them)
static int eam_inrange_ensure(b200md_ctx *c)
{
  const size_t ntiles = (size_t) cdiv(c->nlocal, TILE);
  TRY(dev_ensure(c, c->numneigh_in, (size_t) c->nlocal));
  return dev_ensure(c, c->neigh_in, ntiles * TILE * c->maxneigh);
}

// {x,y,z,fp} records of atoms [first, first+n) from pos + fp (host-ghost mode: ghost fp arrives through the host)
__global__ void k_posfp_fill(const double4 *__restrict__ pos, const double *__restrict__ fp, double4 *__restrict__ posfp,
                             int first, int n)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 p = pos[first + i];
  p.w = fp[first + i];
  posfp[first + i] = p;
}

// passes A+B (pair_eam.cpp:187-247): rho, fp and the embedding energy (-> red_d[8]) of the owned atoms
int pair_eam_rho(b200md_ctx *c, int ev)
{
  const int n = c->nlocal;
  const EAMTables &T = c->eam;
  const EAMParams P = eam_params(c);
  if (eam_use_1t(c)) {
    const int threads = EAM1T_THREADS, blocks = cdiv(n, threads);
    const size_t n1 = (size_t) T.nr + 1, smemA = 4 * n1 * 8;
    TRY(dev_ensure(c, c->posfp, (size_t) n + c->nghost));
    TRY(eam_inrange_ensure(c));
    if (ev) TRY(dev_ensure(c, c->partials, (size_t) blocks * 8));
    TRY(eam_attrs_1t(c));
    {
      KProf kp(c, KC_EAM_RHO);
#define A1 c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, P, T.tab1, c->rho.p, c->fp.p, c->posfp.p, c->partials.p, nullptr, -1, c->numneigh_in.p, c->neigh_in.p
      if (ev) k_eam_rho_fp_1t<1, 0><<<blocks, threads, smemA, c->stream>>>(A1);
      else k_eam_rho_fp_1t<0, 0><<<blocks, threads, smemA, c->stream>>>(A1);
#undef A1
      LAUNCH_CHECK();
    }
    if (ev) TRY(reduce_partials(c, blocks, 1, 8));
    return B200MD_OK;
  }
  const int threads = 128, blocks = cdiv(n, threads);
  if (ev) TRY(dev_ensure(c, c->partials, (size_t) blocks * 8));
  const int onetype = (T.ntypes == 1 && T.nrhor == 1 && T.nz2r == 1);
  {
    KProf kp(c, KC_EAM_RHO);
    launch_eam_rho(c, blocks, threads, ev, onetype, P, n);
    LAUNCH_CHECK();
  }
  if (ev) TRY(reduce_partials(c, blocks, 1, 8));  // embedding energy -> red_d[8]
  return B200MD_OK;
}

// pass C (pair_eam.cpp:257-325); fp of owned AND ghost atoms must be in place (ghost_from_fp: build the ghosts'
// {x,y,z,fp} records of the single-type path from pos + fp instead of expecting them from the halo)
int pair_eam_force(b200md_ctx *c, int ev, bool ghost_from_fp)
{
  const int n = c->nlocal;
  const EAMTables &T = c->eam;
  const EAMParams P = eam_params(c);
  if (eam_use_1t(c)) {
    const int threads = EAM1T_THREADS, blocks = cdiv(n, threads);
    const size_t n1 = (size_t) T.nr + 1, smemC = 7 * n1 * 8;
    if (ev) TRY(dev_ensure(c, c->partials, (size_t) blocks * 8));
    if (ghost_from_fp && c->nghost > 0) {
      k_posfp_fill<<<cdiv(c->nghost, 256), 256, 0, c->stream>>>(c->pos.p, c->fp.p, c->posfp.p, n, c->nghost);
      LAUNCH_CHECK();
    }
    NVEFuse nve = {};
    const bool integ = !ev && nve_fuse_args(c, nve);
    {
      KProf kp(c, KC_PAIR);
#define C1 c->posfp.p, c->numneigh_in.p, c->neigh_in.p, c->maxneigh, n, P, T.tab1, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p, nullptr, -1, c->pos.p, nve
      if (ev) k_eam_force_1t<1, 0, 0><<<blocks, threads, smemC, c->stream>>>(C1);
      else if (integ) {
        k_eam_force_1t<0, 0, 1><<<blocks, threads, smemC, c->stream>>>(C1);
        c->fuse.done = true;
      } else k_eam_force_1t<0, 0, 0><<<blocks, threads, smemC, c->stream>>>(C1);
#undef C1
      LAUNCH_CHECK();
    }
    if (ev) TRY(reduce_partials(c, blocks, 7, 0));
    return B200MD_OK;
  }
  const int threads = 128, blocks = cdiv(n, threads);
  if (ev) TRY(dev_ensure(c, c->partials, (size_t) blocks * 8));
  const int onetype = (T.ntypes == 1 && T.nrhor == 1 && T.nz2r == 1);
  {
    KProf kp(c, KC_PAIR);
    launch_eam_force(c, blocks, threads, ev, onetype, P, n);
    LAUNCH_CHECK();
  }
  if (ev) TRY(reduce_partials(c, blocks, 7, 0));  // pair energy + virial -> red_d[0..6]; eng = red[0] + red[8]
  return B200MD_OK;
}

int pair_eam_compute(b200md_ctx *c, int eflag, int vflag)
{
  const int ev = (eflag || vflag) ? 1 : 0;
  TRY(pair_eam_rho(c, ev));
  // PairEAM's forward_comm of fp (pair_eam.cpp:251): single-type path ships {x + shift, fp} in one message
  if (eam_use_1t(c)) TRY(halo_forward_vec4(c, c->posfp.p));
  else TRY(halo_forward_scalar(c, c->fp.p));
  return pair_eam_force(c, ev, false);
}

// ---- overlap mode (capi.cu): the single-type passes restricted to the tiles whose ghost flag equals `part` ----
bool pair_eam_can_split(const b200md_ctx *c) { return eam_use_1t(c); }

int pair_eam_rho_part(b200md_ctx *c, int part)
{
  const int n = c->nlocal;
  const EAMTables &T = c->eam;
  const EAMParams P = eam_params(c);
  const int ntiles = cdiv(n, TILE), cnt = part ? c->ntiles_ghost : ntiles - c->ntiles_ghost;
  const int *list = c->tile_list.p + (part ? ntiles - c->ntiles_ghost : 0);
  // the small second launch (6-9 % of the tiles) in 128-thread blocks: finer tail than 3 waves of 512-thread blocks
  const int threads = part ? 128 : EAM1T_THREADS, blocks = cdiv(cnt * TILE, threads);
  const size_t n1 = (size_t) T.nr + 1, smemA = 4 * n1 * 8;
  TRY(dev_ensure(c, c->posfp, (size_t) n + c->nghost));
  TRY(eam_inrange_ensure(c));
  if (cnt == 0) return B200MD_OK;
  TRY(eam_attrs_1t(c));
  KProf kp(c, KC_EAM_RHO);
  k_eam_rho_fp_1t<0, 1><<<blocks, threads, smemA, c->stream>>>(c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, P, T.tab1, c->rho.p,
                                                                c->fp.p, c->posfp.p, c->partials.p, list, cnt, c->numneigh_in.p,
                                                                c->neigh_in.p);
  LAUNCH_CHECK();
  return B200MD_OK;
}

int pair_eam_force_part(b200md_ctx *c, int part)
{
  const int n = c->nlocal;
  const EAMTables &T = c->eam;
  const EAMParams P = eam_params(c);
  const int ntiles = cdiv(n, TILE), cnt = part ? c->ntiles_ghost : ntiles - c->ntiles_ghost;
  if (cnt == 0) return B200MD_OK;
  const int *list = c->tile_list.p + (part ? ntiles - c->ntiles_ghost : 0);
  // the small second launch (6-9 % of the tiles) in 128-thread blocks: finer tail than 3 waves of 512-thread blocks
  const int threads = part ? 128 : EAM1T_THREADS, blocks = cdiv(cnt * TILE, threads);
  const size_t n1 = (size_t) T.nr + 1, smemC = 7 * n1 * 8;
  TRY(eam_attrs_1t(c));
  NVEFuse nve = {};
  const bool integ = nve_fuse_args(c, nve);
  KProf kp(c, KC_PAIR);
#define C1 c->posfp.p, c->numneigh_in.p, c->neigh_in.p, c->maxneigh, n, P, T.tab1, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p, list, cnt, c->pos.p, nve
  if (integ) {
    k_eam_force_1t<0, 1, 1><<<blocks, threads, smemC, c->stream>>>(C1);
    c->fuse.done = true;
  } else k_eam_force_1t<0, 1, 0><<<blocks, threads, smemC, c->stream>>>(C1);
#undef C1
  LAUNCH_CHECK();
  return B200MD_OK;
}

// ---- host-buffer pipeline (capi.cu): passes A+B and the fp halo for all atoms, then pass C range by range ----
bool pair_eam_range_usable(const b200md_ctx *c) { return eam_use_1t(c); }

int pair_eam_before_ranges(b200md_ctx *c)
{
  TRY(pair_eam_rho(c, 0));
  return halo_forward_vec4(c, c->posfp.p);
}

int pair_eam_force_range(b200md_ctx *c, int first, int count)
{
  const EAMTables &T = c->eam;
  const EAMParams P = eam_params(c);
  const int threads = EAM1T_THREADS, blocks = cdiv(count, threads);
  if (first % threads) return b200md_fail(c, B200MD_ERR_ARG, "pair range must start at a multiple of %d", threads);
  const size_t n1 = (size_t) T.nr + 1, smemC = 7 * n1 * 8;
  TRY(eam_attrs_1t(c));
  KProf kp(c, KC_PAIR);
  k_eam_force_1t<0, 2, 0><<<blocks, threads, smemC, c->stream>>>(c->posfp.p, c->numneigh_in.p, c->neigh_in.p, c->maxneigh, first + count, P, T.tab1,
                                                                  c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p, nullptr, first, c->pos.p, NVEFuse{});
  LAUNCH_CHECK();
  return B200MD_OK;
}

// ---- per-atom tallies (see pair_lj.cu): eatom = F(rho_i) + half of each pair's phi, vatom = half of each pair's
// del x del * fpair (pair_eam.cpp:240-246, 300-323 with eflag_atom / vflag_atom).  Runs after a force evaluation at
// the current positions: rho of the owned atoms and fp of owned + ghost atoms are in place.
__global__ void __launch_bounds__(128)
k_eam_peratom(const double4 *__restrict__ pos, const int *__restrict__ numneigh, const int *__restrict__ neigh, int maxneigh,
              int nlocal, EAMParams P, const double *__restrict__ rho, const double *__restrict__ fp, const double4 *__restrict__ posfp,
              double *__restrict__ out, size_t stride)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const int nt1 = P.ntypes + 1;
  const double4 pi = pos[i];
  const int it = pos_type(pi);
  const double fpi = posfp ? posfp[i].w : fp[i];
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  {  // embedding energy of atom i (pass B)
    double p = rho[i] * P.rdrho + 1.0;
    int m = (int) p;
    m = max(1, min(m, P.nrho - 1));
    p -= m;
    p = fmin(p, 1.0);
    const double *cf = P.frho + ((size_t) P.type2frho[it] * (P.nrho + 1) + m) * 7;
    double phi = ((__ldg(cf + 3) * p + __ldg(cf + 4)) * p + __ldg(cf + 5)) * p + __ldg(cf + 6);
    if (rho[i] > P.rhomax) phi += fpi * (rho[i] - P.rhomax);
    acc[0] = phi;
  }
  const int n = numneigh[i];
  const int *nb = neigh + (size_t) (i >> 5) * TILE * maxneigh + (i & 31);
  for (int k = 0; k < n; k++) {
    const int j = __ldg(nb + (size_t) k * TILE) & CLJMASK;
    const double4 pj = pos[j];
    const double dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
    const double rsq = dx * dx + dy * dy + dz * dz;
    if (rsq >= P.cutforcesq) continue;
    const int jt = pos_type(pj);
    const double fpj = posfp ? posfp[j].w : fp[j];
    const double r = sqrt(rsq);
    double p = r * P.rdr + 1.0;
    int m = (int) p;
    m = min(m, P.nr - 1);
    p -= m;
    p = fmin(p, 1.0);
    const double *c1 = P.rhor + ((size_t) P.type2rhor[it * nt1 + jt] * (P.nr + 1) + m) * 7;
    const double *c2 = P.rhor + ((size_t) P.type2rhor[jt * nt1 + it] * (P.nr + 1) + m) * 7;
    const double *c3 = P.z2r + ((size_t) P.type2z2r[it * nt1 + jt] * (P.nr + 1) + m) * 7;
    const double rhoip = (__ldg(c1) * p + __ldg(c1 + 1)) * p + __ldg(c1 + 2);
    const double rhojp = (__ldg(c2) * p + __ldg(c2 + 1)) * p + __ldg(c2 + 2);
    const double z2p = (__ldg(c3) * p + __ldg(c3 + 1)) * p + __ldg(c3 + 2);
    const double z2 = ((__ldg(c3 + 3) * p + __ldg(c3 + 4)) * p + __ldg(c3 + 5)) * p + __ldg(c3 + 6);
    const double recip = 1.0 / r;
    const double phi = z2 * recip;
    const double phip = z2p * recip - phi * recip;
    const double psip = fpi * rhojp + fpj * rhoip + phip;
    const double hf = -0.5 * psip * recip;
    acc[0] += 0.5 * phi;
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

int pair_eam_peratom(b200md_ctx *c, double *out, size_t stride)
{
  const int n = c->nlocal;
  const EAMParams P = eam_params(c);
  // single-type path: fp of owned and ghost atoms lives in the {x,y,z,fp} records; generic path: in fp[nall]
  const double4 *posfp = eam_use_1t(c) ? c->posfp.p : nullptr;
  k_eam_peratom<<<cdiv(n, 128), 128, 0, c->stream>>>(c->pos.p, c->numneigh.p, c->neigh.p, c->maxneigh, n, P, c->rho.p, c->fp.p, posfp, out, stride);
  LAUNCH_CHECK();
  return B200MD_OK;
}
