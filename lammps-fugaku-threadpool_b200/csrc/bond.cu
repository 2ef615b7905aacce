// bond.cu -- what bench/in.chain adds to the pair path:
//   bond_style fene        BondFENE::compute (src/MOLECULE/bond_fene.cpp:52-127) over the bond list that
//                          NTopoBondAll::build + Domain::closest_image produce (src/ntopo_bond_all.cpp:37-76,
//                          src/domain.cpp:1054-1082)
//   fix langevin           FixLangevin::post_force, plain instance (src/fix_langevin.cpp:583-716)
// Device design, in the spirit of the full neighbor list: every owned atom evaluates ALL of its bonds itself (both
// directions are stored), so there are no atomics and no reverse_comm of ghost forces; energy and virial are tallied
// half per visit like Pair::ev_tally_full.  The bond topology is a table indexed by atom TAG (partner tags and bond
// types, b200md_set_bond_topology), so it needs no permutation when atoms are re-sorted or migrate.  The device index
// of each partner's closest image is resolved inside the neighbor build (k_build_seg, BONDED): a 1-2 partner always
// lies within the neighbor cutoff (FENE: r < R0), is dropped from the pair list there (special_bonds lj weight 0) and
// recorded in bond_idx instead.
#include "ctx.cuh"
#include "bond_dev.cuh"

__device__ __forceinline__ double warp_sum_b(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}

template <int EVFLAG>
__global__ void __launch_bounds__(256)
k_bond_fene(const double4 *__restrict__ pos, const int *__restrict__ tag, int nlocal, int maxb,
            const int *__restrict__ bond_idx, const int *__restrict__ bond_type_by_tag, FeneParams P, double *__restrict__ fx,
            double *__restrict__ fy, double *__restrict__ fz, double *__restrict__ partials, int *__restrict__ flags)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  if (i < nlocal) {
    double fxi = 0.0, fyi = 0.0, fzi = 0.0;
    fene_atom<EVFLAG>(pos, i, pos[i], tag[i], maxb, bond_idx, bond_type_by_tag, P, flags, fxi, fyi, fzi, acc);
    fx[i] += fxi;
    fy[i] += fyi;
    fz[i] += fzi;
  }
  if (EVFLAG) {
    __shared__ double s_red[8][7];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int c = 0; c < 7; c++) {
      const double v = warp_sum_b(acc[c]);
      if (lane == 0) s_red[w][c] = v;
    }
    __syncthreads();
    if (threadIdx.x < 7) {
      double t = 0.0;
      for (int k = 0; k < (int) (blockDim.x >> 5); k++) t += s_red[k][threadIdx.x];
      partials[(size_t) blockIdx.x * 8 + threadIdx.x] = t;
    }
  }
}

// f += gamma1 * v + gamma2 * (u - 0.5), u = the three uniform numbers the host drew for this atom (by tag) in the
// reference's order; same operation order as fix_langevin.cpp:655-713 (fdrag + fran first, then added to f)
__global__ void k_langevin(const double4 *__restrict__ pos, const int *__restrict__ tag, int nlocal, const double *__restrict__ vx,
                           const double *__restrict__ vy, const double *__restrict__ vz, const double *__restrict__ rand_by_tag,
                           const double *__restrict__ gamma1, const double *__restrict__ gamma2, double *__restrict__ fx,
                           double *__restrict__ fy, double *__restrict__ fz)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const int type = __double2loint(pos[i].w);
  const double g1 = gamma1[type], g2 = gamma2[type];
  const double *u = rand_by_tag + (size_t) (tag[i] - 1) * 3;
  const double fran0 = __dmul_rn(g2, u[0] - 0.5), fran1 = __dmul_rn(g2, u[1] - 0.5), fran2 = __dmul_rn(g2, u[2] - 0.5);
  const double fdrag0 = __dmul_rn(g1, vx[i]), fdrag1 = __dmul_rn(g1, vy[i]), fdrag2 = __dmul_rn(g1, vz[i]);
  fx[i] = __dadd_rn(fx[i], __dadd_rn(fdrag0, fran0));
  fy[i] = __dadd_rn(fy[i], __dadd_rn(fdrag1, fran1));
  fz[i] = __dadd_rn(fz[i], __dadd_rn(fdrag2, fran2));
}

// ---- counter-based mode: Philox4x32-10 keyed by the fix's seed, counter = (tag, timestep): bond_dev.cuh ----
__global__ void k_langevin_counter(const double4 *__restrict__ pos, const int *__restrict__ tag, int nlocal, const double *__restrict__ vx,
                                   const double *__restrict__ vy, const double *__restrict__ vz, int seed, int64_t step, LangevinGamma G,
                                   double *__restrict__ fx, double *__restrict__ fy, double *__restrict__ fz)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  double a = fx[i], b = fy[i], cz = fz[i];
  langevin_counter_kick(G, pos_type(pos[i]), seed, tag[i], step, vx[i], vy[i], vz[i], a, b, cz);
  fx[i] = a;
  fy[i] = b;
  fz[i] = cz;
}

// What follows the pair force on a step of in.chain that neither tallies nor ends the run, in ONE pass per atom instead
// of three streaming kernels: the FENE bonds of atom i (k_bond_fene), the thermostat kick (k_langevin_counter; counter
// mode only -- the replay mode draws on the host), FixNVE::final_integrate and the next step's initial_integrate
// (k_nve_final_initial) -- the same operations in the same order, so the trajectory is bit-identical.  Bond partners are
// gathered from the OLD positions; the new ones go to the second position buffer (NVEFuse, ctx.cuh).
__global__ void __launch_bounds__(256)
k_chain_post(const double4 *__restrict__ pos, int nlocal, const double *__restrict__ fx, const double *__restrict__ fy,
             const double *__restrict__ fz, ChainPost cp, NVEFuse nve)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  chain_post_atom(cp, nve, pos, i, pos[i], fx[i], fy[i], fz[i]);
}

int bond_compute(b200md_ctx *c, int ev)
{
  if (!c->bonds.maxb) return B200MD_OK;
  if (!c->bonds.idx_valid) return b200md_fail(c, B200MD_ERR_ARG, "bond partners were not resolved by the last neighbor build");
  const int n = c->nlocal, threads = 256, blocks = cdiv(n, threads);
  if (ev) TRY(dev_ensure(c, c->partials, (size_t) blocks * 8));
  {
    KProf kp(c, KC_PAIR);
#define B_ARGS c->pos.p, c->tag.p, n, c->bonds.maxb, c->bonds.idx.p, c->bonds.type_by_tag.p, c->bonds.fene, c->f[0].p, c->f[1].p, c->f[2].p, c->partials.p, c->flags_d
    if (ev) k_bond_fene<1><<<blocks, threads, 0, c->stream>>>(B_ARGS);
    else k_bond_fene<0><<<blocks, threads, 0, c->stream>>>(B_ARGS);
#undef B_ARGS
    LAUNCH_CHECK();
  }
  if (ev) {
    TRY(reduce_partials(c, blocks, 7, 24));  // bond energy + virial -> red_d[24..30]
    CU(cudaMemcpyAsync(c->red_h + 24, c->red_d + 24, 7 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->bonds.eng_bond = c->red_h[24];
    for (int k = 0; k < 6; k++) c->bonds.virial_bond[k] = c->red_h[25 + k];
  }
  return B200MD_OK;
}

// ---- RanMars (src/random_mars.cpp:30-100), host side: the reference's stream, replayed in the reference's order ----
void RanMarsHost::init(int seed)
{
  for (double &q : u) q = 0.0;
  int ij = (seed - 1) / 30082, kl = (seed - 1) - 30082 * ij;
  int i = (ij / 177) % 177 + 2, j = ij % 177 + 2, k = (kl / 169) % 178 + 1, l = kl % 169;
  for (int ii = 1; ii <= 97; ii++) {
    double s = 0.0, t = 0.5;
    for (int jj = 1; jj <= 24; jj++) {
      const int m = ((i * j) % 179) * k % 179;
      i = j; j = k; k = m;
      l = (53 * l + 1) % 169;
      if ((l * m) % 64 >= 32) s = s + t;
      t = 0.5 * t;
    }
    u[ii] = s;
  }
  c = 362436.0 / 16777216.0;
  cd = 7654321.0 / 16777216.0;
  cm = 16777213.0 / 16777216.0;
  i97 = 97;
  j97 = 33;
  uniform();
}

double RanMarsHost::uniform()
{
  double uni = u[i97] - u[j97];
  if (uni < 0.0) uni += 1.0;
  u[i97] = uni;
  i97--;
  if (i97 == 0) i97 = 97;
  j97--;
  if (j97 == 0) j97 = 97;
  c -= cd;
  if (c < 0.0) c += cm;
  uni -= c;
  if (uni < 0.0) uni += 1.0;
  return uni;
}

// FixLangevin::post_force for timestep `ntimestep` of a run [beginstep, endstep] (compute_target, :773-787)
// gfactor1 / gfactor2 of the step (FixLangevin::compute_target :773-787, init :284-298; ratio = 1), g2 times sqrt(T)
static void langevin_gammas(const b200md_ctx *c, int64_t ntimestep, double *g1, double *g2)
{
  const LangevinState &L = c->langevin;
  double delta = (double) (ntimestep - L.beginstep);
  if (delta != 0.0) delta /= (double) (L.endstep - L.beginstep);
  const double t_target = L.t_start + delta * (L.t_stop - L.t_start);
  const double tsqrt = sqrt(t_target);
  for (int t = 1; t <= c->ntypes; t++) {
    double a = -c->mass_h[t] / L.t_period / c->ftm2v;
    double b = sqrt(c->mass_h[t]) / c->ftm2v;
    b *= sqrt(24.0 * c->boltz / L.t_period / L.dt / c->mvv2e);
    a *= 1.0 / 1.0;
    b *= 1.0 / sqrt(1.0);
    g1[t] = a;
    g2[t] = b * tsqrt;
  }
  g1[0] = g2[0] = 0.0;
}

// bonds + thermostat + fix nve of a step that neither tallies nor ends the run, fused (k_chain_post); false if this
// system's extras cannot be fused (replay-mode thermostat)
bool chain_post_usable(const b200md_ctx *c) { return !c->langevin.on || c->langevin.rng_mode == 1; }

// the arguments of chain_post_atom for this step (bond tables, thermostat factors)
int chain_post_args(b200md_ctx *c, int64_t ntimestep, ChainPost &cp)
{
  if (c->bonds.maxb && !c->bonds.idx_valid) return b200md_fail(c, B200MD_ERR_ARG, "bond partners were not resolved by the last neighbor build");
  cp = {};
  cp.tag = c->tag.p;
  cp.maxb = c->bonds.maxb;
  cp.bond_idx = c->bonds.idx.p;
  cp.bond_type_by_tag = c->bonds.type_by_tag.p;
  cp.P = c->bonds.fene;
  cp.flags = c->flags_d;
  cp.langevin = c->langevin.on ? 1 : 0;
  cp.seed = c->langevin.seed;
  cp.step = ntimestep;
  if (c->langevin.on) {
    double g1[B200MD_MAXTYPES + 1], g2[B200MD_MAXTYPES + 1];
    langevin_gammas(c, ntimestep, g1, g2);
    for (int t = 0; t <= c->ntypes; t++) { cp.G.g1[t] = g1[t]; cp.G.g2[t] = g2[t]; }
  }
  return B200MD_OK;
}

int chain_post_fused(b200md_ctx *c, int64_t ntimestep)
{
  NVEFuse nve = {};
  if (!nve_fuse_args(c, nve)) return b200md_fail(c, B200MD_ERR_ARG, "chain_post_fused without a fused-integration request");
  ChainPost cp;
  TRY(chain_post_args(c, ntimestep, cp));
  const int n = c->nlocal;
  KProf kp(c, KC_NVE);
  k_chain_post<<<cdiv(n, 256), 256, 0, c->stream>>>(c->pos.p, n, c->f[0].p, c->f[1].p, c->f[2].p, cp, nve);
  LAUNCH_CHECK();
  c->fuse.done = true;
  return B200MD_OK;
}

int langevin_apply(b200md_ctx *c, int64_t ntimestep)
{
  LangevinState &L = c->langevin;
  if (!L.on) return B200MD_OK;
  const int n = c->nlocal;
  if (L.rng_mode == 0) {
    if (c->nranks > 1)
      return b200md_fail(c, B200MD_ERR_ARG, "fix langevin replay mode supports one rank (RanMars(seed + me) per rank); use the counter mode (b200md_langevin_rng)");
    if ((int) L.order.size() != n) return b200md_fail(c, B200MD_ERR_ARG, "fix langevin: b200md_langevin_order must list every owned atom");
  }
  double g1[B200MD_MAXTYPES + 1], g2[B200MD_MAXTYPES + 1];
  langevin_gammas(c, ntimestep, g1, g2);
  if (L.rng_mode == 1) {
    LangevinGamma G;
    for (int t = 0; t <= B200MD_MAXTYPES; t++) { G.g1[t] = t <= c->ntypes ? g1[t] : 0.0; G.g2[t] = t <= c->ntypes ? g2[t] : 0.0; }
    k_langevin_counter<<<cdiv(n, 256), 256, 0, c->stream>>>(c->pos.p, c->tag.p, n, c->v[0].p, c->v[1].p, c->v[2].p, L.seed, ntimestep, G,
                                                            c->f[0].p, c->f[1].p, c->f[2].p);
    LAUNCH_CHECK();
    return B200MD_OK;
  }
  // three numbers per atom, drawn in the reference's local-index order, stored by tag
  for (int k = 0; k < n; k++) {
    double *u = L.rand_h + (size_t) (L.order[k] - 1) * 3;
    u[0] = L.rng.uniform();
    u[1] = L.rng.uniform();
    u[2] = L.rng.uniform();
  }
  TRY(dev_ensure(c, L.rand_d, (size_t) 3 * L.ntags));
  TRY(dev_ensure(c, L.gamma_d, (size_t) 2 * (B200MD_MAXTYPES + 1)));
  CU(cudaMemcpyAsync(L.rand_d.p, L.rand_h, (size_t) 3 * L.ntags * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(L.gamma_d.p, g1, (c->ntypes + 1) * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(L.gamma_d.p + (B200MD_MAXTYPES + 1), g2, (c->ntypes + 1) * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_langevin<<<cdiv(n, 256), 256, 0, c->stream>>>(c->pos.p, c->tag.p, n, c->v[0].p, c->v[1].p, c->v[2].p, L.rand_d.p, L.gamma_d.p,
                                                  L.gamma_d.p + (B200MD_MAXTYPES + 1), c->f[0].p, c->f[1].p, c->f[2].p);
  LAUNCH_CHECK();
  // g1/g2 live on this stack frame and rand_h is refilled next step: finish the copies before returning
  CU(cudaStreamSynchronize(c->stream));
  return B200MD_OK;
}

// Atom::sort (src/atom.cpp:2203-2290, setup_sort_bins :2296-2400) applied to the REFERENCE's local order of the owned
// atoms: bins of userbinsize (or half the neighbor cutoff) over the sub-domain, atoms keep their relative order
// inside a bin.  Only the order is needed here (the device keeps its own layout); positions are the ones the
// reference sorts on, i.e. after Domain::pbc.
int langevin_resort(b200md_ctx *c)
{
  LangevinState &L = c->langevin;
  L.nextsort = (c->ntimestep / L.sortfreq) * L.sortfreq + L.sortfreq;
  const int n = c->nlocal;
  if ((int) L.order.size() != n) return b200md_fail(c, B200MD_ERR_ARG, "fix langevin: b200md_langevin_order must list every owned atom");
  const double binsize = L.userbinsize > 0.0 ? L.userbinsize : 0.5 * c->cutneighmax;
  if (binsize <= 0.0) return B200MD_OK;
  int nb[3];
  double inv[3];
  for (int d = 0; d < 3; d++) {
    const double len = c->boxhi[d] - c->boxlo[d];
    nb[d] = (int) (len * (1.0 / binsize));
    if (nb[d] == 0) nb[d] = 1;
    inv[d] = nb[d] / len;
  }
  const int64_t nbins64 = (int64_t) nb[0] * nb[1] * nb[2];
  if (nbins64 == 1) return B200MD_OK;
  if (nbins64 > (int64_t) 1 << 30) return b200md_fail(c, B200MD_ERR_ARG, "Too many atom sorting bins");
  const int nbins = (int) nbins64;
  std::vector<double4> pos((size_t) n);
  std::vector<int> tag((size_t) n);
  CU(cudaMemcpyAsync(pos.data(), c->pos.p, (size_t) n * sizeof(double4), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(tag.data(), c->tag.p, (size_t) n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  std::vector<int> dev_of_tag((size_t) L.ntags + 1, -1);
  for (int i = 0; i < n; i++) {
    if (tag[i] < 1 || tag[i] > L.ntags) return b200md_fail(c, B200MD_ERR_ARG, "fix langevin: atom tag outside 1..ntags");
    dev_of_tag[tag[i]] = i;
  }
  std::vector<int> binhead((size_t) nbins, -1), next((size_t) n), sorted((size_t) n);
  for (int k = n - 1; k >= 0; k--) {
    const int i = dev_of_tag[L.order[k]];
    if (i < 0) return b200md_fail(c, B200MD_ERR_ARG, "fix langevin: an atom of the order list is not on the device");
    int ix = (int) ((pos[i].x - c->boxlo[0]) * inv[0]);
    int iy = (int) ((pos[i].y - c->boxlo[1]) * inv[1]);
    int iz = (int) ((pos[i].z - c->boxlo[2]) * inv[2]);
    ix = std::min(std::max(ix, 0), nb[0] - 1);
    iy = std::min(std::max(iy, 0), nb[1] - 1);
    iz = std::min(std::max(iz, 0), nb[2] - 1);
    const int ibin = iz * nb[1] * nb[0] + iy * nb[0] + ix;
    next[k] = binhead[ibin];
    binhead[ibin] = k;
  }
  int m = 0;
  for (int b = 0; b < nbins; b++)
    for (int k = binhead[b]; k >= 0; k = next[k]) sorted[m++] = L.order[k];
  L.order.swap(sorted);
  return B200MD_OK;
}

void bond_release(b200md_ctx *c)
{
  dev_free(c, c->bonds.partner_by_tag);
  dev_free(c, c->bonds.type_by_tag);
  dev_free(c, c->bonds.idx);
  dev_free(c, c->bonds.best);
  dev_free(c, c->langevin.rand_d);
  dev_free(c, c->langevin.gamma_d);
  if (c->langevin.rand_h) cudaFreeHost(c->langevin.rand_h);
  c->langevin.rand_h = nullptr;
}

// ---- C ABI (include/b200md.h, "in.chain additions") ------------------------------------------------
#define NEED_CTX()                                                                 \
  if (!c) return b200md_fail(nullptr, B200MD_ERR_ARG, "NULL context");             \
  if (cudaSetDevice(c->device) != cudaSuccess) return b200md_fail(c, B200MD_ERR_CUDA, "cudaSetDevice failed")

extern "C" int b200md_set_bonds(b200md_ctx *c, int n, const int *tag, int bond_per_atom, const int *num_bond, const int *bond_type,
                                const int *bond_atom, int newton_bond)
{
  NEED_CTX();
  BondState &B = c->bonds;
  if (n == 0 || bond_per_atom == 0) {  // no bonds: back to the plain atomic path
    B.maxb = 0;
    B.idx_valid = false;
    c->list_valid = false;
    return B200MD_OK;
  }
  if (n < 0 || bond_per_atom < 0 || !tag || !num_bond || !bond_type || !bond_atom)
    return b200md_fail(c, B200MD_ERR_ARG, "set_bonds: tag, num_bond, bond_type, bond_atom are required");
  int ntags = 0;
  for (int i = 0; i < n; i++) ntags = std::max(ntags, tag[i]);
  std::vector<int> deg((size_t) ntags + 1, 0);
  for (int i = 0; i < n; i++) {
    if (tag[i] < 1) return b200md_fail(c, B200MD_ERR_ARG, "set_bonds: tags must be >= 1");
    if (num_bond[i] < 0 || num_bond[i] > bond_per_atom) return b200md_fail(c, B200MD_ERR_ARG, "set_bonds: num_bond outside 0..bond_per_atom");
    for (int m = 0; m < num_bond[i]; m++) {
      const int p = bond_atom[(size_t) i * bond_per_atom + m];
      if (p < 1 || p > ntags) return b200md_fail(c, B200MD_ERR_ARG, "Bond atoms %d %d missing", tag[i], p);
      if (bond_type[(size_t) i * bond_per_atom + m] <= 0) continue;  // turned-off bond (ntopo_bond_all.cpp:65)
      deg[tag[i]]++;
      if (newton_bond) deg[p]++;
    }
  }
  int maxb = 0;
  for (int t = 1; t <= ntags; t++) maxb = std::max(maxb, deg[t]);
  if (maxb > B200MD_MAXBONDS)
    return b200md_fail(c, B200MD_ERR_ARG, "set_bonds: an atom has %d bonds, this library holds up to %d per atom", maxb, B200MD_MAXBONDS);
  if (maxb == 0) { B.maxb = 0; return B200MD_OK; }
  std::vector<int> partner((size_t) ntags * maxb, 0), type((size_t) ntags * maxb, 0);
  std::fill(deg.begin(), deg.end(), 0);
  auto add = [&](int a, int b, int t) {
    partner[(size_t) (a - 1) * maxb + deg[a]] = b;
    type[(size_t) (a - 1) * maxb + deg[a]] = t;
    deg[a]++;
  };
  for (int i = 0; i < n; i++)
    for (int m = 0; m < num_bond[i]; m++) {
      const int p = bond_atom[(size_t) i * bond_per_atom + m], t = bond_type[(size_t) i * bond_per_atom + m];
      if (t <= 0) continue;
      if (t > B200MD_MAXBONDTYPES) return b200md_fail(c, B200MD_ERR_ARG, "set_bonds: bond type %d > %d", t, B200MD_MAXBONDTYPES);
      add(tag[i], p, t);
      if (newton_bond) add(p, tag[i], t);
    }
  TRY(dev_ensure(c, B.partner_by_tag, partner.size()));
  TRY(dev_ensure(c, B.type_by_tag, type.size()));
  CU(cudaMemcpyAsync(B.partner_by_tag.p, partner.data(), partner.size() * sizeof(int), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(B.type_by_tag.p, type.data(), type.size() * sizeof(int), cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  B.maxb = maxb;
  B.ntags = ntags;
  B.idx_valid = false;
  c->list_valid = false;  // the pair list must drop the 1-2 partners: rebuild
  c->setup_done = false;
  return B200MD_OK;
}

extern "C" int b200md_bond_fene(b200md_ctx *c, int nbondtypes, const double *k, const double *r0, const double *epsilon,
                                const double *sigma)
{
  NEED_CTX();
  if (nbondtypes < 1 || nbondtypes > B200MD_MAXBONDTYPES || !k || !r0 || !epsilon || !sigma)
    return b200md_fail(c, B200MD_ERR_ARG, "bond_fene: 1..%d bond types, arrays [nbondtypes+1]", B200MD_MAXBONDTYPES);
  FeneParams &P = c->bonds.fene;
  for (int t = 0; t <= B200MD_MAXBONDTYPES; t++) { P.k[t] = 0.0; P.r0[t] = 1.0; P.epsilon[t] = 0.0; P.sigma[t] = 0.0; }
  for (int t = 1; t <= nbondtypes; t++) { P.k[t] = k[t]; P.r0[t] = r0[t]; P.epsilon[t] = epsilon[t]; P.sigma[t] = sigma[t]; }
  c->bonds.fene_set = true;
  return B200MD_OK;
}

extern "C" int b200md_thermo_bond(b200md_ctx *c, double *eng_bond, double *virial)
{
  NEED_CTX();
  if (!c->bonds.maxb) return b200md_fail(c, B200MD_ERR_ARG, "thermo_bond: no bonds defined");
  if (!c->have_ev) return b200md_fail(c, B200MD_ERR_ARG, "no force evaluation has tallied energy/virial yet");
  double loc[7] = {c->bonds.eng_bond, c->bonds.virial_bond[0], c->bonds.virial_bond[1], c->bonds.virial_bond[2],
                   c->bonds.virial_bond[3], c->bonds.virial_bond[4], c->bonds.virial_bond[5]};
  if (c->nranks > 1) {  // collective, like b200md_thermo: Bond::energy / virial summed over the ranks (compute pe, pressure)
    CU(cudaMemcpyAsync(c->red_d + 16, loc, sizeof(loc), cudaMemcpyHostToDevice, c->stream));
    TRY(comm_allreduce_sum(c, c->red_d + 16, 7));
    CU(cudaMemcpyAsync(loc, c->red_d + 16, sizeof(loc), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  if (eng_bond) *eng_bond = loc[0];
  if (virial) for (int k = 0; k < 6; k++) virial[k] = loc[1 + k];
  return B200MD_OK;
}

extern "C" int b200md_fix_langevin(b200md_ctx *c, double t_start, double t_stop, double t_period, int seed, double dt)
{
  NEED_CTX();
  LangevinState &L = c->langevin;
  if (t_period <= 0.0 && seed == 0) {  // unfix
    L.on = false;
    L.seed = 0;
    return B200MD_OK;
  }
  if (t_period <= 0.0) return b200md_fail(c, B200MD_ERR_ARG, "Fix langevin period must be > 0.0");
  if (seed <= 0) return b200md_fail(c, B200MD_ERR_ARG, "Illegal fix langevin command");  // fix_langevin.cpp:84-85
  if (seed > 900000000) return b200md_fail(c, B200MD_ERR_ARG, "Invalid seed for Marsaglia random # generator");
  L.t_start = t_start; L.t_stop = t_stop; L.t_period = t_period; L.dt = dt;
  // the fix owns one generator for its whole life (created in the constructor, fix_langevin.cpp:160): a repeated
  // call for the same fix (same seed, e.g. at the start of every `run`) keeps the stream where it is
  if (!L.on || L.seed != seed) L.rng.init(seed + c->rank);
  L.seed = seed;
  L.on = true;
  c->setup_done = false;
  return B200MD_OK;
}

extern "C" int b200md_langevin_order(b200md_ctx *c, int n, const int *tag, int sortfreq, double userbinsize)
{
  NEED_CTX();
  LangevinState &L = c->langevin;
  if (n <= 0 || !tag || sortfreq < 0) return b200md_fail(c, B200MD_ERR_ARG, "langevin_order: n tags in the reference's local order");
  int ntags = 0;
  for (int i = 0; i < n; i++) {
    if (tag[i] < 1) return b200md_fail(c, B200MD_ERR_ARG, "langevin_order: tags must be >= 1");
    ntags = std::max(ntags, tag[i]);
  }
  L.order.assign(tag, tag + n);
  if (ntags != L.ntags) {
    if (L.rand_h) cudaFreeHost(L.rand_h);
    L.rand_h = nullptr;
    CU(cudaMallocHost((void **) &L.rand_h, (size_t) 3 * ntags * sizeof(double)));
    memset(L.rand_h, 0, (size_t) 3 * ntags * sizeof(double));
    L.ntags = ntags;
  }
  L.sortfreq = sortfreq;
  L.userbinsize = userbinsize;
  L.nextsort = 0;
  return B200MD_OK;
}

extern "C" int b200md_run_bounds(b200md_ctx *c, int64_t beginstep, int64_t endstep)
{
  NEED_CTX();
  if (endstep < beginstep) return b200md_fail(c, B200MD_ERR_ARG, "run_bounds: endstep < beginstep");
  c->langevin.beginstep = beginstep;
  c->langevin.endstep = endstep;
  c->langevin.bounds_set = true;
  return B200MD_OK;
}

// host only, no device needed: the library's RanMars stream (what b200md_fix_langevin consumes)
extern "C" int b200md_ranmars(int seed, int n, double *out)
{
  if (seed <= 0 || seed > 900000000) return b200md_fail(nullptr, B200MD_ERR_ARG, "Invalid seed for Marsaglia random # generator");
  if (n < 0 || (n > 0 && !out)) return b200md_fail(nullptr, B200MD_ERR_ARG, "ranmars: n >= 0 and an output array");
  RanMarsHost r;
  r.init(seed);
  for (int i = 0; i < n; i++) out[i] = r.uniform();
  return B200MD_OK;
}

extern "C" int b200md_langevin_rng(b200md_ctx *c, int mode)
{
  NEED_CTX();
  if (mode != 0 && mode != 1) return b200md_fail(c, B200MD_ERR_ARG, "langevin_rng: 0 = RanMars replay, 1 = counter-based");
  c->langevin.rng_mode = mode;
  return B200MD_OK;
}

// host only: the three uniforms the counter mode gives atom `tag` on timestep `step` (same function as the kernel's)
extern "C" int b200md_langevin_counter_uniforms(int seed, int tag, int64_t step, double *u)
{
  if (!u) return b200md_fail(nullptr, B200MD_ERR_ARG, "langevin_counter_uniforms: u[3] is required");
  langevin_counter_uniforms(seed, tag, step, u);
  return B200MD_OK;
}
// host only: raw Philox4x32-10 block, for known-answer tests
extern "C" int b200md_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
  if (!ctr || !key || !out) return b200md_fail(nullptr, B200MD_ERR_ARG, "philox4x32_10: ctr[4], key[2], out[4]");
  philox4x32_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1], out);
  return B200MD_OK;
}
