// bond_dev.cuh -- per-atom device functions of what bench/in.chain adds behind the pair force (bond.cu): the FENE bonds
// of one atom and the counter-mode Langevin kick.  Shared by the kernels of bond.cu and by the force kernel that runs
// them in its epilogue (k_lj_force, INTEG 2, pair_lj.cu).
#pragma once
#include "ctx.cuh"

// all bonds of owned atom i (tag ti, position pi), both directions being stored: the force on i into (fxi, fyi, fzi);
// with EVFLAG half of each bond's energy and virial into acc
template <int EVFLAG>
__device__ __forceinline__ void fene_atom(const double4 *__restrict__ pos, int i, const double4 pi, int ti, int maxb,
                                          const int *__restrict__ bond_idx, const int *__restrict__ bond_type_by_tag,
                                          const FeneParams &P, int *__restrict__ flags, double &fxi, double &fyi, double &fzi,
                                          double (&acc)[7])
{
  const double CUBEROOT2 = 1.25992104989487316476;  // MY_CUBEROOT2 (src/math_const.h)
  for (int m = 0; m < maxb; m++) {
    const int j = bond_idx[(size_t) i * maxb + m];
    if (j < 0) continue;  // no bond in this slot
    const int type = bond_type_by_tag[(size_t) (ti - 1) * maxb + m];
    const double4 pj = ld_pos256(pos + j);
    const double delx = pi.x - pj.x, dely = pi.y - pj.y, delz = pi.z - pj.z;
    const double rsq = delx * delx + dely * dely + delz * delz;
    const double r0sq = P.r0[type] * P.r0[type];
    double rlogarg = 1.0 - rsq / r0sq;
    if (rlogarg < 0.1) {
      if (rlogarg <= -3.0) flags[4] = 1;  // "Bad FENE bond" (bond_fene.cpp:88)
      rlogarg = 0.1;
    }
    double fbond = -P.k[type] / rlogarg, sr6 = 0.0;
    const double s2 = P.sigma[type] * P.sigma[type];
    const bool lj = rsq < CUBEROOT2 * s2;
    if (lj) {
      const double sr2 = s2 / rsq;
      sr6 = sr2 * sr2 * sr2;
      fbond += 48.0 * P.epsilon[type] * sr6 * (sr6 - 0.5) / rsq;
    }
    fxi += delx * fbond;
    fyi += dely * fbond;
    fzi += delz * fbond;
    if (EVFLAG) {
      double ebond = -0.5 * P.k[type] * r0sq * log(rlogarg);
      if (lj) ebond += 4.0 * P.epsilon[type] * sr6 * (sr6 - 1.0) + P.epsilon[type];
      const double hf = 0.5 * fbond;  // each bond is visited from both of its atoms
      acc[0] += 0.5 * ebond;
      acc[1] += delx * delx * hf;
      acc[2] += dely * dely * hf;
      acc[3] += delz * delz * hf;
      acc[4] += delx * dely * hf;
      acc[5] += delx * delz * hf;
      acc[6] += dely * delz * hf;
    }
  }
}


// ---- counter-based mode: Philox4x32-10 (Salmon et al., SC'11) keyed by the fix's seed, counter = (tag, timestep) ----
// The three numbers of an atom depend on nothing but (seed, tag, step): independent of the rank count, of the atom
// order and of how a run is split into calls; no host work and no copy per step.  Statistically equivalent to the
// reference's thermostat, not stream-compatible with it (that is what the replay mode above is for).
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                                                       uint32_t out[4])
{
#pragma unroll
  for (int r = 0; r < 10; r++) {
    const uint64_t p0 = (uint64_t) 0xD2511F53u * c0, p1 = (uint64_t) 0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t) (p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t) p1, n2 = (uint32_t) (p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t) p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
// uniform in (0,1) with 32 random bits, symmetric about 1/2 (so u - 0.5 has exactly zero mean over the 2^32 values)
__host__ __device__ __forceinline__ double u01_32(uint32_t x) { return ((double) x + 0.5) * 2.3283064365386963e-10; }
__host__ __device__ __forceinline__ void langevin_counter_uniforms(int seed, int tag, int64_t step, double u[3])
{
  uint32_t r[4];
  philox4x32_10((uint32_t) tag, (uint32_t) (uint64_t) step, (uint32_t) ((uint64_t) step >> 32), 0x4c616e67u /* "Lang" */, (uint32_t) seed,
                0x62323030u /* "b200" */, r);
  u[0] = u01_32(r[0]);
  u[1] = u01_32(r[1]);
  u[2] = u01_32(r[2]);
}

struct LangevinGamma { double g1[B200MD_MAXTYPES + 1], g2[B200MD_MAXTYPES + 1]; };
// f += gamma1 * v + gamma2 * (u - 0.5) for one atom, u from the counter generator
__device__ __forceinline__ void langevin_counter_kick(const LangevinGamma &G, int type, int seed, int tagi, int64_t step, double vx,
                                                      double vy, double vz, double &fx, double &fy, double &fz)
{
  const double g1 = G.g1[type], g2 = G.g2[type];
  double u[3];
  langevin_counter_uniforms(seed, tagi, step, u);
  fx = __dadd_rn(fx, __dadd_rn(__dmul_rn(g1, vx), __dmul_rn(g2, u[0] - 0.5)));
  fy = __dadd_rn(fy, __dadd_rn(__dmul_rn(g1, vy), __dmul_rn(g2, u[1] - 0.5)));
  fz = __dadd_rn(fz, __dadd_rn(__dmul_rn(g1, vz), __dmul_rn(g2, u[2] - 0.5)));
}

// what the epilogue of a force kernel needs to run the chain's post-force work for its atom (bond.cu fills it)
struct ChainPost {
  const int *tag;
  int maxb;
  const int *bond_idx, *bond_type_by_tag;
  FeneParams P;
  int *flags;
  int langevin, seed;
  int64_t step;
  LangevinGamma G;
};
// pair force (fx, fy, fz) of atom i in registers -> + FENE bonds -> + thermostat kick -> fix nve (final + next initial);
// the operations and order of k_bond_fene, k_langevin_counter, k_nve_final_initial
__device__ __forceinline__ void chain_post_atom(const ChainPost &cp, const NVEFuse &nve, const double4 *__restrict__ pos, int i,
                                                const double4 pi, double fx, double fy, double fz)
{
  const int ti = cp.tag[i];
  double acc[7];
  double bx = 0.0, by = 0.0, bz = 0.0;
  if (cp.maxb > 0) fene_atom<0>(pos, i, pi, ti, cp.maxb, cp.bond_idx, cp.bond_type_by_tag, cp.P, cp.flags, bx, by, bz, acc);
  double a = fx + bx, b = fy + by, cz = fz + bz;
  if (cp.langevin) langevin_counter_kick(cp.G, pos_type(pi), cp.seed, ti, cp.step, nve.vx[i], nve.vy[i], nve.vz[i], a, b, cz);
  nve_fused_update(nve, i, pi, a, b, cz);
}
int chain_post_args(b200md_ctx *c, int64_t ntimestep, ChainPost &cp);   // bond.cu
