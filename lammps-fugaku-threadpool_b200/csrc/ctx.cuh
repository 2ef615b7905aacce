// ctx.cuh -- internal state of libb200md.so (sm_100a).  Not part of the C ABI (include/b200md.h is).
//
// Device data layout (all FP64, atoms resident across timesteps):
//   pos[nall]      double4 {x, y, z, w}; low 32 bits of w hold the atom type (1-based, as in the
//                  reference's atom->type, src/atom.h:70-73).  One 32-byte sector per gathered neighbor.
//                  owned atoms [0,nlocal) are kept in bin order (counting sort at every rebuild);
//                  ghost atoms [nlocal,nall) are grouped by the direction (26 of them) they came from.
//   v[3][nlocal], f[3][nlocal]   SoA, streamed by the integrate kernels.
//   tag[nall], image[nlocal]     int32 (LAMMPS_SMALLBIG, src/lmptype.h:90-130).
//   neigh          FULL neighbor list, "blocked transposed": atoms are grouped in tiles of 32; entry k of
//                  atom i lives at neigh[(i>>5)*32*maxneigh + k*32 + (i&31)], so a warp reads one
//                  coalesced 128-byte row per k and each tile is one contiguous slab in HBM.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "../../include/b200md.h"

#define NEIGHMASK 0x1FFFFFFF  // src/lmptype.h:61-65 (what b200md_neigh_export hands out)
#define SBBITS 30
// device list entry: neighbor index in the low 28 bits; the top 4 bits are free for per-entry flags (special-bond
// class of a molecular system, as the reference keeps in the top 2 bits of its own entries, src/lmptype.h:61-65)
#define CLBITS 28
#define CLJMASK 0x0FFFFFFF
#define IMGMASK 1023          // src/lmptype.h (LAMMPS_SMALLBIG)
#define IMGBITS 10
#define IMG2BITS 20
#define IMGMAX 512

#define B200MD_MAXTYPES 16
#define BIN_MARGIN 2
#define NDIR 26               // halo directions (3^3 - 1)
#define TILE 32

template <class T> struct DevBuf {
  T *p = nullptr;
  size_t n = 0;  // capacity in elements
};

struct HaloPlan {             // passed by value to the halo kernels
  int goff[NDIR + 1];         // ghost group offsets (relative to nlocal), by the direction the atoms travelled in
  int soff[NDIR + 1];         // send-list offsets by direction
  int self[NDIR];             // 1 if the receiver of direction d is this rank itself (periodic image)
  double shift[NDIR][3];      // periodic shift the SENDER applies (src/atom_vec.cpp:363-379)
};

struct BinGeom {
  double lo[3];               // sublo of this rank
  double inv[3];              // 1 / binsize
  int nbin[3];                // bins covering the sub-domain
  int mbin[3];                // nbin + 2*BIN_MARGIN
  int mbins;
};

struct LJTables {             // what PairLJCut::init_one fills (src/pair_lj_cut.cpp:501-557)
  int ntypes;
  double *d;                  // device: 6 tables of (ntypes+1)^2: cutsq, lj1, lj2, lj3, lj4, offset
  double special_lj[4];
  double h_onetype[6];        // ntypes == 1 fast path: the [1][1] entries
};

struct EAMTables {            // what PairEAM::array2spline produces (src/MANYBODY/pair_eam.cpp:744-797)
  int ntypes, nrho, nr, nfrho, nrhor, nz2r;
  double rdrho, rdr, rhomax, cutforcesq;
  double *frho;               // device [nfrho][nrho+1][7]
  double4 *rhoA;              // device [nrhor][nr+1]  value coeffs {c3,c4,c5,c6} of rhor_spline  (pass A)
  double *tab1;               // device, single-type fast path: SoA [11][nr+1] = rhor c3..c6, rhor c0..c2, z2r c3..c6
  double *rhor, *z2r;         // device, original 7-coefficient layout
  int *type2frho, *type2rhor, *type2z2r;  // device
  int h_type2frho[B200MD_MAXTYPES + 1];
};

// ---- in.chain additions (bond.cu) ----
#define B200MD_MAXBONDTYPES 8
#define B200MD_MAXBONDS 4        // bond slots per atom (both directions), held in registers by the neighbor build
struct FeneParams {           // bond_coeff K R0 epsilon sigma (src/MOLECULE/bond_fene.cpp:142-165), 1-based
  double k[B200MD_MAXBONDTYPES + 1], r0[B200MD_MAXBONDTYPES + 1], epsilon[B200MD_MAXBONDTYPES + 1], sigma[B200MD_MAXBONDTYPES + 1];
};
struct BondState {
  int maxb = 0;               // bond slots per atom (both directions stored); 0 = no bonds
  int ntags = 0;              // the tables below have ntags rows (tags 1..ntags)
  DevBuf<int> partner_by_tag; // [ntags][maxb] partner tag, 0 = empty slot
  DevBuf<int> type_by_tag;    // [ntags][maxb] bond type
  DevBuf<int> idx;            // [nlocal][maxb] device index of each partner's closest image (neighbor build), -1 = empty
  DevBuf<unsigned long long> best;  // [nlocal][maxb] scratch of the build: (rsq bits << 32 | index) minimum
  bool idx_valid = false;
  bool fene_set = false;
  FeneParams fene;
  double eng_bond = 0.0, virial_bond[6] = {0, 0, 0, 0, 0, 0};
};
// general special-bond lists of a molecular system whose atoms LAMMPS keeps on the host (host-ghost mode): what
// Atom::nspecial / Atom::special hold (src/atom.h:93-95) and Neighbor::special_flag (src/neighbor.h:82)
struct SpecialState {
  bool on = false;
  int maxspecial = 0;
  int flag[4] = {1, 1, 1, 1};       // 0 exclude, 1 ordinary neighbor, 2 keep with the class in the entry's bits 28-29
  DevBuf<int> n3;                   // [nlocal][3] cumulative counts of 1-2, 1-3, 1-4 partners
  DevBuf<int> tags;                 // [nlocal][maxspecial] partner tags
};
struct RanMarsHost {          // src/random_mars.cpp
  double u[98], c, cd, cm;
  int i97, j97;
  void init(int seed);
  double uniform();
};
struct LangevinState {
  bool on = false;
  double t_start = 0, t_stop = 0, t_period = 1, dt = 0;
  int64_t beginstep = 0, endstep = 0;
  int seed = 0;
  RanMarsHost rng;
  std::vector<int> order;     // tags of the owned atoms in the REFERENCE's local order (the order RanMars is consumed in)
  int sortfreq = 0;           // atom_modify sort: the reference's local order changes at every Atom::sort
  double userbinsize = 0.0;
  int64_t nextsort = 0;
  bool bounds_set = false;    // b200md_run_bounds given (a run split into several b200md_run calls)
  int rng_mode = 0;           // 0 = the reference's RanMars stream replayed on the host, 1 = counter-based on the device
  double *rand_h = nullptr;   // pinned [ntags][3]
  int ntags = 0;
  DevBuf<double> rand_d, gamma_d;
};

// kernel classes for the event timers (b200md_kernel_times)
enum { KC_PAIR = 0, KC_EAM_RHO = 1, KC_NEIGH = 2, KC_NVE = 3, KC_HALO = 4, KC_SORT = 5, KC_N = 6 };
struct KEvent { cudaEvent_t a, b; int cls; };

struct b200md_ctx {
  int device = 0, nranks = 1, rank = 0;
  cudaStream_t stream = nullptr;
  char err[512] = {0};

  // units / box (src/update.cpp:145-207, src/domain.h)
  double boltz = 1, mvv2e = 1, ftm2v = 1, nktv2p = 1;
  double boxlo[3] = {0, 0, 0}, boxhi[3] = {0, 0, 0}, prd[3] = {0, 0, 0};
  int periodic[3] = {1, 1, 1}, procgrid[3] = {1, 1, 1}, myloc[3] = {0, 0, 0};
  double sublo[3], subhi[3];
  bool have_box = false;
  bool host_ghosts = false;   // host-buffer mode: ghosts are supplied by the caller (hostmode.cu)
  int ntypes = 0;
  double mass_h[B200MD_MAXTYPES + 1];
  DevBuf<double> mass_d;

  // atoms
  int nlocal = 0, nghost = 0;
  DevBuf<double4> pos, pos_tmp, xhold;
  DevBuf<double> v[3], v_tmp[3], f[3];
  DevBuf<int> tag, tag_tmp, image, image_tmp;

  // pair styles
  int pair_style = 0;  // 0 none, 1 lj/cut, 2 eam
  LJTables lj = {};
  EAMTables eam = {};
  DevBuf<double> rho, fp;
  DevBuf<double4> posfp;      // EAM single-type path: {x, y, z, fp} per atom (owned + ghost), one sector per gather
  double cutforce = 0;

  // neighbor settings (src/neighbor.cpp:306-695)
  DevBuf<double> cutneighsq_d;
  double cutneighsq_h[(B200MD_MAXTYPES + 1) * (B200MD_MAXTYPES + 1)];
  double cutneighmax = 0, skin = 0, triggersq = 0, binsize_user = 0;
  int every = 1, delay = 0, dist_check = 1;
  bool have_neigh = false;
  int ago = 0;
  int64_t nbuilds = 0, ndanger = 0;
  bool list_valid = false;

  // bins / sort
  BinGeom bg;
  DevBuf<int> bin_count, bin_start, atom_bin, cell_list, perm, scan_tmp;
  DevBuf<unsigned char> atom_sub;
  // candidate records of the list build (neigh.cu): cell_list order, every bin padded to the build's vector width, so the
  // 3 x (tg + 2) bins of a staged row are ONE contiguous, 16-byte aligned range
  DevBuf<float> rec;          // [rec_cap * nf] groups of vw records {x[vw], y[vw], z[vw], jt[vw] (, tag[vw])}: FP32 coordinates
                              // relative to the bin-grid origin, index | (type - 1) << 28, tags for bonded / molecular systems
  DevBuf<int> pbin_start;     // [mbins + 1] padded start of every bin

  // neighbor list
  int maxneigh = 0;
  DevBuf<int> numneigh, neigh;
  DevBuf<int> numneigh_in, neigh_in;   // EAM single-type path: the entries inside the force cutoff, left by the rho pass for the force pass of the same step
  DevBuf<int> tile_ghost;     // [ntiles]: != 0 if an atom of the 32-atom tile has a ghost among its candidates (overlap mode)
  DevBuf<int> tile_list, tile_pos;   // tile numbers, ghost-free tiles first (ascending), then the others; scan scratch
  int ntiles_ghost = 0;       // tiles in the second part
  bool tiles_valid = false;
  int maxneigh_seen = 0;

  // halo / exchange (comm.cu)
  // overlap (default on several ranks, b200md_tune "overlap"): the per-step halo runs on s_comm while the main stream
  // computes the tiles whose lists hold no ghost; hstream != nullptr redirects the halo kernels and NCCL calls there
  int overlap = 0;
  // fix nve fused into the force kernel (b200md_run, steps that neither tally nor end the run; pair_lj.cu): the request
  // for the force evaluation under way, and whether the kernels honoured it (new positions then sit in pos_tmp)
  // check: the fused epilogue also runs Neighbor::check_distance for the next step (which will decide with it);
  // check_ahead: flags_d[2] already holds that verdict for the step now starting
  // chain: the epilogue also runs the FENE bonds and the counter-mode thermostat of its atom (k_lj_force INTEG 2)
  struct { bool on = false, done = false, check = false, chain = false; double dtv = 0.0, dtf = 0.0; } fuse;
  bool check_ahead = false;
  int fuse_nve = 1;           // "fuse_nve" (b200md_tune)
  cudaStream_t s_comm = nullptr, hstream = nullptr;
  cudaEvent_t ev_main = nullptr, ev_halo = nullptr;
  HaloPlan hp;
  void *nccl = nullptr;       // ncclComm_t when nranks > 1
  int peer_send[NDIR], peer_recv[NDIR];
  int sendcount[NDIR], recvcount[NDIR];
  int nsend = 0;
  int64_t nmigrated = 0;
  DevBuf<int> send_idx, border_blockcnt, border_total, sendbuf_i, allcounts;
  DevBuf<signed char> send_dir;
  DevBuf<double4> sendbuf4, recvbuf4;
  DevBuf<double> sendbuf1;
  // per-step halo as a fused pack-and-put over peer memory (comm.cu, "halo_put"): the pack kernel writes every remote
  // record straight into the receiver's landing buffer over NVLink and raises the receiver's per-direction flag; the
  // receiver waits on its own flags and copies the records into the ghost slots.  One cudaMalloc per rank, exported by
  // cudaIpcGetMemHandle: [2][32] int flags (256 B), then two landing buffers (by halo parity) of `cap` records.
  struct PutState {
    int enabled = 0;              // "halo_put" (b200md_tune)
    bool imported = false;        // the peers' allocations are mapped
    bool epoch_ok = false;        // agreed by all ranks at the last borders(): every rank's ghosts fit its landing buffer
    size_t cap = 0;               // records per landing buffer
    char *base = nullptr;         // my allocation
    void *peer_base[64] = {};     // the peers' (mapped), indexed by rank
    size_t peer_cap[64] = {};
    unsigned seq = 0;             // halos put so far (the same number on every rank)
    int *done_d = nullptr;        // [2]: blocks of the pack kernel that finished; a wait that timed out
    int peer_goff[NDIR] = {};     // ghost offset of direction d on the rank that receives my direction d
  } put;

  // reductions / flags
  DevBuf<double> partials;  // [nblocks][8]
  int build_gb = 1;          // bins of an x-row per block (segment) of the last list build (k_build_seg)
  int seg_cap = 0, seg_cap_tg = 0;   // its staging capacity in records (grown on demand) and the segment length it was sized for
  DevBuf<double> peratom;   // [7][peratom_stride]: eatom, vatom xx yy zz xy xz yz of the owned atoms (b200md_pair_peratom)
  size_t peratom_stride = 0;
  int peratom_n = 0;        // owned atoms the tallies are valid for, 0 = none
  double *red_d = nullptr;  // [16] device results
  double *red_h = nullptr;  // pinned mirror
  int *flags_d = nullptr;   // [8]: 0 nonfinite, 1 overflow(max count), 2 moved, 3 lost
  int *flags_h = nullptr;   // pinned mirror
  double eng_vdwl = 0, virial[6] = {0, 0, 0, 0, 0, 0};
  bool have_ev = false;

  // in.chain additions
  BondState bonds;
  SpecialState special;
  LangevinState langevin;

  // run state
  int64_t ntimestep = 0;
  bool setup_done = false;

  // instrumentation: level 1 = per-kernel-class CUDA events recorded on `stream` without any host sync
  // (pairs are parked in a ring and resolved when read), level 2 = also the blocking per-section timers
  double sect_ms[5] = {0, 0, 0, 0, 0};
  int64_t launches = 0;
  int profile = 0;
  std::vector<KEvent> kev;
  size_t kev_used = 0;
  double kc_ms[KC_N] = {0, 0, 0, 0, 0, 0};
  int64_t kc_count[KC_N] = {0, 0, 0, 0, 0, 0};
  double bytes_held = 0;

  // knobs (b200md_tune; defaults chosen from hardware measurements)
  int host_pipeline = 1;      // b200md_step_host: three-stream chunk pipeline (0 = serial copy-compute-copy)
  int hs_min_chunk = 65536;   // smallest system the pipeline splits into ranges
  int build_gb_user = 0;      // test knob: bins of an x-row per block in the list build (0 = by bin population)
  int seg_cap_user = 0;       // test knob: initial staging capacity of the list build (records)
  int build_threads = 128;    // threads per block of the list build (128 or 256; the bins per block follow)
  bool xhold_by_permute = false;   // the rebuild's permutation kernel wrote xhold (neigh_build_lists then does not copy it)
  int build_bond_general = 0; // test knob: bonded build with the closest-image choice (BONDED 1) even where a tag is unique inside the cutoff
  int build_fp64 = 0;         // test knob: force the all-FP64 build kernel (otherwise only huge boxes with tiny cutoffs use it)
  int debug_timing = 0;       // per-phase wall times of every rebuild on stderr (B200MD_DEBUG_TIMING, read at create)

  // host-buffer mode staging and its copy streams (b200md_step_host)
  DevBuf<double> stage;
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t hs_ev[26] = {};
};

// ---------------------------------------------------------------------------------------------
// error handling: no exceptions across the C ABI, int status + text (include/b200md.h)
// ---------------------------------------------------------------------------------------------
int b200md_fail(b200md_ctx *c, int code, const char *fmt, ...);

#define CU(call)                                                                                       \
  do {                                                                                                 \
    cudaError_t e__ = (call);                                                                          \
    if (e__ != cudaSuccess)                                                                            \
      return b200md_fail(c, B200MD_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call,                \
                         cudaGetErrorString(e__));                                                     \
  } while (0)

#define TRY(call)                                                                                      \
  do {                                                                                                 \
    int s__ = (call);                                                                                  \
    if (s__ != B200MD_OK) return s__;                                                                  \
  } while (0)

#define LAUNCH_CHECK()                                                                                 \
  do {                                                                                                 \
    c->launches++;                                                                                     \
    cudaError_t e__ = cudaGetLastError();                                                              \
    if (e__ != cudaSuccess)                                                                            \
      return b200md_fail(c, B200MD_ERR_CUDA, "%s:%d kernel launch: %s", __FILE__, __LINE__,            \
                         cudaGetErrorString(e__));                                                     \
  } while (0)

template <class T> int dev_ensure(b200md_ctx *c, DevBuf<T> &b, size_t n, bool keep = false, bool slack = true)
{
  if (n <= b.n) return B200MD_OK;
  size_t want = slack ? n + n / 8 + 64 : n;
  T *q = nullptr;
  CU(cudaMalloc((void **) &q, want * sizeof(T)));
  if (keep && b.p && b.n) CU(cudaMemcpyAsync(q, b.p, b.n * sizeof(T), cudaMemcpyDeviceToDevice, c->stream));
  if (b.p) {
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaFree(b.p));
    c->bytes_held -= (double) b.n * sizeof(T);
  }
  b.p = q;
  b.n = want;
  c->bytes_held += (double) want * sizeof(T);
  return B200MD_OK;
}

// fix nve fused into a force kernel (k_lj_force / k_eam_force_1t, INTEG): what the kernel needs beside its own arguments
struct NVEFuse {
  double *vx, *vy, *vz;
  const double *mass;
  double4 *pos_next;
  double dtv, dtf;
  // Neighbor::check_distance of the NEXT step, done on the positions just made (nullptr: that step does not check)
  const double4 *xhold;
  double triggersq;
  int *moved_flag;
};
// has the atom moved more than half the skin since the last build?  (Neighbor::check_distance, src/neighbor.cpp:2322-2374)
__device__ __forceinline__ bool moved_beyond_trigger(const double4 a, const double4 b, double triggersq)
{
  const double dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z;
  const double rsq = dx * dx + dy * dy + dz * dz;
  return rsq > triggersq;
}
// the fused integration wanted by b200md_run for the force evaluation under way, if any (capi.cu sets c->fuse; the
// launchers that honour it set c->fuse.done)
static inline bool nve_fuse_args(b200md_ctx *c, NVEFuse &a)
{
  if (!c->fuse.on) return false;
  a = {c->v[0].p, c->v[1].p, c->v[2].p, c->mass_d.p, c->pos_tmp.p, c->fuse.dtv, c->fuse.dtf,
       c->fuse.check ? c->xhold.p : nullptr, c->triggersq, c->flags_d + 2};
  return true;
}
// the epilogue itself: FixNVE::final_integrate of this step + initial_integrate of the next for atom i holding force
// (fx, fy, fz); operations and order of k_nve_final_initial (nve.cu).  p0 = the atom's position record {x, y, z, type}
__device__ __forceinline__ void nve_fused_update(const NVEFuse &nve, int i, const double4 p0, double fx, double fy, double fz)
{
  const double dtfm = nve.dtf / nve.mass[__double2loint(p0.w)];
  const double gx = __dmul_rn(dtfm, fx), gy = __dmul_rn(dtfm, fy), gz = __dmul_rn(dtfm, fz);
  const double a = __dadd_rn(__dadd_rn(nve.vx[i], gx), gx);
  const double b = __dadd_rn(__dadd_rn(nve.vy[i], gy), gy);
  const double cz = __dadd_rn(__dadd_rn(nve.vz[i], gz), gz);
  nve.vx[i] = a;
  nve.vy[i] = b;
  nve.vz[i] = cz;
  const double4 pn = make_double4(__dadd_rn(p0.x, __dmul_rn(nve.dtv, a)), __dadd_rn(p0.y, __dmul_rn(nve.dtv, b)),
                                  __dadd_rn(p0.z, __dmul_rn(nve.dtv, cz)), p0.w);
  nve.pos_next[i] = pn;
  if (nve.xhold && moved_beyond_trigger(pn, nve.xhold[i], nve.triggersq)) *nve.moved_flag = 1;
}

template <class T> void dev_free(b200md_ctx *c, DevBuf<T> &b)
{
  if (b.p) cudaFree(b.p);
  c->bytes_held -= (double) b.n * sizeof(T);
  b.p = nullptr;
  b.n = 0;
}

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// per-kernel-class timing scope: two event records on the launching stream, no host synchronisation
void kprof_drain(b200md_ctx *c);
struct KProf {
  b200md_ctx *c;
  KEvent *e = nullptr;
  KProf(b200md_ctx *ctx, int cls) : c(ctx)
  {
    if (!c->profile) return;
    if (c->kev_used == c->kev.size()) {
      if (c->kev.size() >= 16384) kprof_drain(c);
      else {
        KEvent n;
        n.cls = cls;
        if (cudaEventCreate(&n.a) != cudaSuccess || cudaEventCreate(&n.b) != cudaSuccess) return;
        c->kev.push_back(n);
      }
    }
    e = &c->kev[c->kev_used++];
    e->cls = cls;
    cudaEventRecord(e->a, c->stream);
  }
  ~KProf()
  {
    if (e) cudaEventRecord(e->b, c->stream);
  }
};

// section timers (names as src/timer.h:23-42)
enum { SECT_PAIR = 0, SECT_NEIGH = 1, SECT_COMM = 2, SECT_MODIFY = 3, SECT_OTHER = 4 };

// ---- internal entry points (one per .cu) ------------------------------------------------------
int neigh_setup_bins(b200md_ctx *c);
int neigh_rebuild(b200md_ctx *c);          // pbc + sort + borders + bin + build
int neigh_build_lists(b200md_ctx *c);      // bin owned+ghost + build (second half of neigh_rebuild)
int neigh_check_distance(b200md_ctx *c, int *moved);
int comm_plan(b200md_ctx *c);
int comm_exchange(b200md_ctx *c);
int comm_borders(b200md_ctx *c);
int comm_allreduce_sum(b200md_ctx *c, double *dev, int n);
int comm_allreduce_max_int(b200md_ctx *c, int *dev, int n);
int halo_forward_pos(b200md_ctx *c);
int halo_forward_vec4(b200md_ctx *c, double4 *arr);
int halo_forward_scalar(b200md_ctx *c, double *a);
int pair_lj_compute(b200md_ctx *c, int eflag, int vflag);
int pair_lj_compute_part(b200md_ctx *c, int part);            // overlap mode: part 0 = ghost-free tiles, 1 = the rest
int pair_eam_rho_part(b200md_ctx *c, int part);
int pair_eam_force_part(b200md_ctx *c, int part);
bool pair_lj_range_usable(const b200md_ctx *c);                 // host-buffer pipeline: forces of an atom range
int pair_lj_compute_range(b200md_ctx *c, int first, int count);
bool pair_eam_range_usable(const b200md_ctx *c);
int pair_eam_before_ranges(b200md_ctx *c);                      // passes A+B and the fp halo
int pair_eam_force_range(b200md_ctx *c, int first, int count);
bool pair_eam_can_split(const b200md_ctx *c);
int pair_eam_compute(b200md_ctx *c, int eflag, int vflag);
int pair_eam_rho(b200md_ctx *c, int ev);
int pair_eam_force(b200md_ctx *c, int ev, bool ghost_from_fp);
int nve_initial(b200md_ctx *c, double dtv, double dtf);
int nve_final(b200md_ctx *c, double dtf);
int thermo_ke(b200md_ctx *c, double *sum_mv2);
int reduce_partials(b200md_ctx *c, int nblocks, int ncomp, int dst);  // partials -> red_d[dst..]
int pair_lj_peratom(b200md_ctx *c, double *out, size_t stride);       // eatom / vatom pass, out = [7][stride]
int pair_eam_peratom(b200md_ctx *c, double *out, size_t stride);

// device helpers
// one 32-byte gather of a position record: sm_100 has 256-bit global loads (LDG.E.ENL2.256), which halves the
// LSU requests and L1 tag look-ups of a double4 gather compared with two 128-bit loads
__device__ __forceinline__ double4 ld_pos256(const double4 *__restrict__ p)
{
  double4 r;
  asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
  return r;
}
// 1/x for finite positive x: MUFU.RCP64H seed (about 20 good bits) + two Newton-Raphson steps = 4 DFMA, within
// 1 ulp of the correctly rounded quotient -- the compiler's 1.0/x spends about twice that and carries a
// divergent slow-path call for denormals/infinities that a squared distance never needs
__device__ __forceinline__ double rcp_pos(double x)
{
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double e = fma(-x, y, 1.0);
  y = fma(y, e, y);
  e = fma(-x, y, 1.0);
  y = fma(y, e, y);
  return y;
}
__device__ __forceinline__ int pos_type(const double4 &p) { return __double2loint(p.w); }
__device__ __forceinline__ double type_as_w(int t) { return __hiloint2double(0, t); }
int nve_final_initial(b200md_ctx *c, double dtv, double dtf);
int bond_compute(b200md_ctx *c, int ev);                   // bond.cu
int langevin_apply(b200md_ctx *c, int64_t ntimestep);     // bond.cu
int langevin_resort(b200md_ctx *c);                        // bond.cu: Atom::sort replayed on the reference's local order
void bond_release(b200md_ctx *c);                          // bond.cu
