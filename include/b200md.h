/* ======================================================================================
 * b200md.h -- C ABI of libb200md.so, the B200 (sm_100a) device library behind the LAMMPS
 * `Pair` / `NPair` / `NBin` / `Fix` / `Comm` style classes with suffix `b200`.
 *
 * Conventions (modelled on the reference's own device-library precedent, the `ljl_gpu_*`,
 * `eam_gpu_*` and `lmp_*_device` entry points that src/GPU/pair_lj_cut_gpu.cpp:34-56,
 * src/GPU/pair_eam_gpu.cpp:36-60 and src/GPU/fix_gpu.cpp:43-51 declare):
 *   - plain C, opaque context, raw pointers + sizes; no C++/torch types cross the boundary;
 *   - host arrays are CALLER-owned and may move between calls (LAMMPS re-allocates atom->x on
 *     any avec->grow), device arrays are LIBRARY-owned and persist across timesteps;
 *   - every entry point returns an int status (B200MD_OK == 0); b200md_last_error() gives the
 *     text.  The status values are the three error conditions the reference path raises
 *     (src/nbin.cpp:143 non-finite coordinate, src/npair_half_bin_atomonly_newton.cpp:109
 *     neighbor overflow, src/thermo.cpp lost atoms) plus CUDA/NCCL failures.  The C++ glue turns
 *     a non-zero status into error->one(FLERR,...) like GPU_EXTRA::check_flag does.
 *   - one context per MPI rank / per GPU, called from the rank's main thread only.
 *   - per-type tables are the reference's 1-based layout flattened row-major with leading
 *     dimension (ntypes+1): element [i][j] at index i*(ntypes+1)+j, row/column 0 unused.
 *   - per-atom host arrays are the reference's: double x[n][3] contiguous (src/memory.h
 *     create(double**,n,3)), int type[n], tagint(int32, LAMMPS_SMALLBIG) tag[n],
 *     imageint(int32) image[n].
 *
 * There is NO CPU fallback: every call fails with B200MD_ERR_CUDA when no sm_100 device is
 * usable.
 * ====================================================================================== */
#ifndef B200MD_H
#define B200MD_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200md_ctx b200md_ctx;

enum {
  B200MD_OK = 0,
  B200MD_ERR_NONFINITE = 1,  /* "Non-numeric positions - simulation unstable" (src/nbin.cpp:143)           */
  B200MD_ERR_NEIGH_OVERFLOW = 2, /* "Neighbor list overflow" (src/npair_half_bin_atomonly_newton.cpp:109)  */
  B200MD_ERR_LOST_ATOMS = 3, /* "Lost atoms" (src/thermo.cpp lost_check)                                   */
  B200MD_ERR_CUDA = 4,       /* CUDA or NCCL failure, or no usable device                                   */
  B200MD_ERR_ARG = 5,        /* invalid argument / call order (the reference's error->all on bad input)    */
  B200MD_ERR_DOMAIN = 6      /* sub-domain thinner than the ghost cutoff in a decomposed dimension         */
};

#define B200MD_NCCL_ID_BYTES 128

/* ---- lifetime --------------------------------------------------------------------------
 * replaces lmp_init_device()/lmp_clear_device() (src/GPU/fix_gpu.cpp:43-49,231,242).
 * nranks>1: one context per rank, `nccl_id` is the B200MD_NCCL_ID_BYTES blob produced by
 * b200md_nccl_unique_id() on rank 0 and broadcast by the caller (MPI_Bcast / torch.distributed);
 * this is what replaces the MPI communicator of Comm (src/comm.h:21-170). */
int b200md_nccl_unique_id(void *id_out);
int b200md_create(b200md_ctx **out, int device, int nranks, int rank, const void *nccl_id);
int b200md_destroy(b200md_ctx *ctx);
const char *b200md_last_error(const b200md_ctx *ctx);
/* device bytes currently held (Pair::memory_usage(), ljl_gpu_bytes()) */
double b200md_memory_usage(const b200md_ctx *ctx);

/* ---- domain / units / masses (Domain: src/domain.h; Update::set_units src/update.cpp:145-207) */
int b200md_set_units(b200md_ctx *ctx, double boltz, double mvv2e, double ftm2v, double nktv2p);
/* orthogonal box; procgrid as ProcMap::onelevel_grid would return it (src/procmap.cpp:46) with
 * uniform cuts like CommBrick::setup (src/comm_brick.cpp:191-260).  rank -> grid location is
 * x-fastest: rank = (iz*py + iy)*px + ix. */
int b200md_set_box(b200md_ctx *ctx, const double boxlo[3], const double boxhi[3], const int periodic[3],
                   const int procgrid[3]);
int b200md_set_mass(b200md_ctx *ctx, int ntypes, const double *mass /* [ntypes+1] */);

/* ---- atoms: Atom / AtomVecAtomic (src/atom.h:70-73, src/atom_vec_atomic.cpp) -------------
 * upload replaces the per-step host->device casts of the GPU package (lib/gpu/lal_atom.cu:20);
 * here it happens once, atoms then stay resident.  v and image may be NULL (zeros / 0 image).
 * Atoms must belong to this rank's sub-domain after periodic wrapping. */
int b200md_upload_atoms(b200md_ctx *ctx, int n, const double *x, const double *v, const int *type,
                        const int *tag, const int *image);
int b200md_nlocal(const b200md_ctx *ctx);
int b200md_nghost(const b200md_ctx *ctx);
/* copy owned atoms back (device order; use `tag` to identify).  Any pointer may be NULL.
 * Mirrors atomKK->sync(Host, X_MASK|V_MASK|F_MASK|...) of the KOKKOS package. */
int b200md_download_atoms(b200md_ctx *ctx, double *x, double *v, double *f, int *type, int *tag, int *image);
/* the same copy-back in the CALLER's order: host_tag[n] are the tags of the caller's n = b200md_nlocal() atoms in its own
 * order (Atom::tag), maxtag their maximum; row i of x / v / f / image receives the atom whose tag is host_tag[i] -- the
 * permutation runs on the device, so the copies land straight in Atom::x / v / f / image (what VerletB200 does at every
 * output step on one rank; the role of atomKK->sync(Host, ...) with the host's own atom order, src/KOKKOS/atom_kokkos.h).
 * Needs reasonably dense tags (maxtag <= 4 n + 1024: a direct table), else B200MD_ERR_ARG and b200md_download_atoms
 * serves; a tag the caller does not hold: B200MD_ERR_ARG.  Any output pointer may be NULL. */
int b200md_download_atoms_by_tag(b200md_ctx *ctx, int n, const int *host_tag, int maxtag, double *x, double *v, double *f,
                                 int *image);
/* owned + ghost positions and tags (ghosts after owned), for neighbor-list parity checks */
int b200md_download_all(b200md_ctx *ctx, double *x /*[nall][3]*/, int *tag /*[nall]*/);
/* refresh positions/velocities of the resident owned atoms from host arrays given in TAG order
 * is not offered: use b200md_upload_atoms again (forces a rebuild), as a restart does
 * (src/neighbor.cpp:2297). */

/* ---- pair_style lj/cut : replaces ljl_gpu_init (src/GPU/pair_lj_cut_gpu.cpp:34-38); tables are
 * what PairLJCut::init_one fills (src/pair_lj_cut.cpp:501-557) */
int b200md_lj_init(b200md_ctx *ctx, int ntypes, const double *cutsq, const double *lj1, const double *lj2,
                   const double *lj3, const double *lj4, const double *offset, const double special_lj[4]);

/* ---- pair_style eam : replaces eam_gpu_init (src/GPU/pair_eam_gpu.cpp:36-43); arrays are what
 * PairEAM::file2array/array2spline produce (src/MANYBODY/pair_eam.cpp:533-765):
 * frho_spline[nfrho][nrho+1][7], rhor_spline[nrhor][nr+1][7], z2r_spline[nz2r][nr+1][7] flattened,
 * type2frho[ntypes+1], type2rhor/type2z2r[(ntypes+1)^2]. */
int b200md_eam_init(b200md_ctx *ctx, int ntypes, int nrho, int nr, double rdrho, double rdr, double rhomax,
                    double cutforcesq, int nfrho, const double *frho_spline, int nrhor,
                    const double *rhor_spline, int nz2r, const double *z2r_spline, const int *type2frho,
                    const int *type2rhor, const int *type2z2r);
/* stand-alone hosts: PairEAM::coeff + read_file + file2array + array2spline + init_one for funcfl files
 * (src/MANYBODY/pair_eam.cpp:372-412, 460-526, 533-765, 435-456).  map[1..ntypes] = index into files[] or -1
 * (type not covered by EAM).  mass_out[1..ntypes] receives the file masses (pair_coeff sets the type mass,
 * :404), *cutmax_out the force cutoff.  Ends by calling b200md_eam_init. */
int b200md_eam_init_funcfl(b200md_ctx *ctx, int ntypes, int nfiles, const char *const *files, const int *map,
                           double *mass_out, double *cutmax_out);
/* pair_style eam/alloy (fs = 0, DYNAMO setfl file) and eam/fs (fs = 1, Finnis-Sinclair file: the density
 * function depends on both elements): PairEAMAlloy / PairEAMFS ::coeff + read_file + file2array
 * (src/MANYBODY/pair_eam_alloy.cpp:55-337, pair_eam_fs.cpp:55-345) + PairEAM::array2spline + init_one, i.e.
 * `pair_coeff * * file El1 El2 ...`: elements[1..ntypes] = element name of each atom type or "NULL".  The file's
 * energies are converted between metal and real units as the reference's readers do.  mass_out / cutmax_out as
 * above.  Ends by calling b200md_eam_init; the force kernels are the same multi-type ones. */
int b200md_eam_init_setfl(b200md_ctx *ctx, int ntypes, const char *file, int fs, const char *const *elements,
                          double *mass_out, double *cutmax_out);
/* per-atom EAM intermediates of owned atoms, device order (PairEAM::extract / rho, fp members) */
int b200md_eam_download(b200md_ctx *ctx, double *rho, double *fp);

/* ---- neighbor : Neighbor settings (src/neighbor.cpp:306-695, 2556-2700) ------------------
 * cutneighsq[(ntypes+1)^2] = (cut_ij + skin)^2, cutneighmax = max; every/delay/check as
 * `neigh_modify`; binsize <= 0 selects the device default (one bin per cutneighmax). */
int b200md_neigh_config(b200md_ctx *ctx, int ntypes, const double *cutneighsq, double cutneighmax, double skin,
                        int every, int delay, int check, double binsize);
/* Neighbor::decide + check_distance (src/neighbor.cpp:2293-2374): *flag = 1 if a rebuild is due.
 * Collective over ranks when nranks > 1 (replaces the MPI_Allreduce at :2371). */
int b200md_neigh_decide(b200md_ctx *ctx, int *flag);
/* the whole reneighbor branch of Verlet::run (src/verlet.cpp:283-306): Domain::pbc, Comm::exchange,
 * spatial sort of owned atoms (Atom::sort), Comm::borders, NBin::bin_atoms (counting sort) and
 * NPair::build.  Collective. */
int b200md_neigh_build(b200md_ctx *ctx);
/* the list as the reference exposes it (lammps_neighlist_element_neighbors, src/library.h:224):
 * numneigh[nlocal]; neighbors = concatenation of each owned atom's list (device-order indices into
 * owned+ghost, special bits in the top 2 bits).  The device list is a FULL list (both i->j and j->i).
 * Call with neighbors == NULL to get numneigh and *total only. */
int b200md_neigh_export(b200md_ctx *ctx, int *numneigh, int *neighbors, int64_t *total);
int b200md_neigh_stats(b200md_ctx *ctx, int64_t *nbuilds, int64_t *ndanger, int *maxneigh);

/* ---- Comm : CommBrick::forward_comm (src/comm_brick.cpp:538-591) -------------------------
 * ghost positions from their owners (self copy, peer copy over NVLink/NCCL).  The device list is
 * full, so the reference's reverse_comm of ghost forces (:598-639) has no device counterpart.
 * Comm::exchange and Comm::borders (:652-979) run inside b200md_neigh_build.  Collective. */
int b200md_halo_forward(b200md_ctx *ctx);
/* The per-step halo as ONE fused pack-and-put kernel over peer memory (NVLink P2P) instead of pack + ncclSend/ncclRecv
 * (what src/KOKKOS/comm_kokkos.cpp:184-262 does with GPU-aware MPI, one step further: no send buffer, no library
 * kernel).  Every rank of one box calls b200md_halo_put_export (allocates its landing buffer of 2 x cap_records ghost
 * records and returns the B200MD_IPC_HANDLE_BYTES cudaIpc handle), the host all-gathers handles and capacities
 * (MPI_Allgather / torch.distributed.all_gather), every rank calls b200md_halo_put_import with all of them, then
 * b200md_tune(ctx, "halo_put", 1).  From the next list build on the position / {x,y,z,fp} halos of b200md_run use it
 * whenever every rank's ghosts fit its landing buffer (agreed collectively at each rebuild; NCCL otherwise).  Ranks must
 * be processes of one node with peer access between their GPUs.  Results are bit-identical to the NCCL halo. */
#define B200MD_IPC_HANDLE_BYTES 64
int b200md_halo_put_export(b200md_ctx *ctx, int64_t cap_records, void *handle_out);
int b200md_halo_put_import(b200md_ctx *ctx, int nranks, const void *handles, const int64_t *caps);
/* halos served by the put path so far (0: every halo went through NCCL) */
int64_t b200md_halo_put_count(const b200md_ctx *ctx);
/* CommBrick::setup (src/comm_brick.cpp:191-483) for the single-stage 26-direction scheme, host only (no
 * device needed): peer_send[d] = rank that receives what `rank` sends in direction d, peer_recv[d] = rank
 * whose direction-d messages `rank` receives; -1 = none (non-periodic face).  d = 0..25 enumerates
 * (dx,dy,dz) in {-1,0,1}^3 without (0,0,0), x fastest.  Arrays of 26. */
int b200md_plan_peers(const int procgrid[3], const int periodic[3], int rank, int *peer_send, int *peer_recv);

/* ---- force : Verlet::force_clear (src/verlet.cpp:376) + Pair::compute ---------------------
 * PairLJCut::compute (src/pair_lj_cut.cpp:69-139) or PairEAM::compute incl. its internal
 * forward_comm of fp (src/MANYBODY/pair_eam.cpp:142-328).  eflag/vflag as Integrate::ev_set gives
 * them; when set, eng_vdwl and virial[6] (xx,yy,zz,xy,xz,yz) of THIS rank are returned like the
 * Pair members (src/pair.h:41-44).  eng/virial may be NULL when the flags are 0. */
int b200md_pair_compute(b200md_ctx *ctx, int eflag, int vflag, double *eng_vdwl, double *virial);

/* ---- fix nve : FixNVE::initial_integrate/final_integrate (src/fix_nve.cpp:64-141) ---------
 * dtv = dt, dtf = 0.5*dt*ftm2v (src/fix_nve.cpp:51-58) */
/* eflag_atom / vflag_atom (Pair::ev_tally per-atom tallies, src/pair.cpp:1037-1097; what compute pe/atom and
 * compute stress/atom read from Pair::eatom / vatom): call after a force evaluation at the current positions.
 * The tallies are a pass of their own over the same list (the timestep kernels carry no per-atom stores).
 * b200md_peratom_download: eatom[nlocal], vatom[nlocal][6] (xx yy zz xy xz yz), either may be NULL, in the
 * order of b200md_download_atoms (host-ghost mode: the caller's own order).  Pair contributions only. */
int b200md_pair_peratom(b200md_ctx *ctx);
int b200md_peratom_download(b200md_ctx *ctx, double *eatom, double *vatom);
int b200md_nve_initial(b200md_ctx *ctx, double dtv, double dtf);
int b200md_nve_final(b200md_ctx *ctx, double dtf);

/* ---- thermo: ComputeTemp/ComputePressure/ComputePE (src/compute_temp.cpp:72-100,
 * src/compute_pressure.cpp:228-341, src/compute_pe.cpp:91-100), summed over ranks.
 * out[0]=pe out[1]=ke out[2]=temp out[3]=press out[4..9]=pressure tensor xx,yy,zz,xy,xz,yz;
 * uses eng/virial of the most recent b200md_pair_compute with flags set. */
int b200md_thermo(b200md_ctx *ctx, double natoms_total, double *out);

/* this rank's eng_vdwl and virial[6] from the most recent force evaluation that tallied them: what a pair style
 * publishes in Pair::eng_vdwl / Pair::virial (src/pair.h:41-44) for Compute pe / pressure after b200md_run */
int b200md_last_ev(b200md_ctx *ctx, double *eng_vdwl, double *virial);

/* ---- Verlet::setup / Verlet::run (src/verlet.cpp:93-162, 229-360) kept on the device side so
 * that no O(N) host work remains per step (what VerletB200 calls).  thermo_every as `thermo N`;
 * energy/virial are tallied on those steps and the last one (Integrate::ev_set,
 * src/integrate.cpp:138-183). */
int b200md_setup(b200md_ctx *ctx);
int b200md_run(b200md_ctx *ctx, int nsteps, double dt, int thermo_every);
int64_t b200md_timestep(const b200md_ctx *ctx);
/* update->ntimestep of the caller (reset_timestep, or a context attached in the middle of a script): the thermo
 * schedule of b200md_run and the Atom::sort schedule of the langevin replay count in these steps */
int b200md_set_timestep(b200md_ctx *ctx, int64_t ntimestep);

/* ---- host-buffer mode: one whole timestep with the atoms handed over in HOST arrays every step -- the GPU
 * package's traffic model (x up, results down every step, src/GPU/pair_lj_cut_gpu.cpp:85-118) applied to the
 * whole step.  On entry x, v hold the b200md_nlocal() owned atoms in the device order of the previous call (the
 * order b200md_download_atoms reports); the call runs initial_integrate / [reneighbor | forward_comm] / force /
 * final_integrate and returns x, v, f (and tag, may be NULL) of the *nlocal_out atoms this rank owns afterwards
 * (a reneighboring step re-orders them and, on several ranks, migrates some).  Arrays are [capacity][3];
 * capacity >= the owned count before and after, else B200MD_ERR_ARG.  Pinned host memory lets the copies run
 * asynchronously: uploads, kernels and downloads are pipelined over atom chunks on three streams. */
int b200md_step_host(b200md_ctx *ctx, double dt, int eflag, int vflag, int capacity, double *x, double *v, double *f,
                     int *tag, int *nlocal_out, double *eng_vdwl, double *virial);

/* ---- host-ghost mode: Pair::compute under the STOCK Verlet loop (run_style verlet).  The GPU package's
 * contract (ljl_gpu_compute_n / eam_gpu_compute_n, src/GPU/pair_lj_cut_gpu.cpp:85-118,
 * src/GPU/pair_eam_gpu.cpp:96-139): LAMMPS owns atoms, ghosts (Comm::borders) and the rebuild decision
 * (Neighbor::decide); every call ships x[nall][3] + type[nall] (owned first, then ghosts) up and the owned
 * forces down.  rebuild != 0 when the host reneighbored this step (neighbor->ago == 0): the device list is then
 * rebuilt from these positions (bins over [sublo, subhi] + ghost margin).  After b200md_host_atoms use
 * b200md_pair_compute (lj/cut) or the two EAM passes below, then b200md_host_forces. */
int b200md_host_atoms(b200md_ctx *ctx, int nlocal, int nghost, const double *x, const int *type,
                      const double sublo[3], const double subhi[3], int rebuild);
/* molecular systems in host-ghost mode: the special-bond lists LAMMPS keeps on the host -- Atom::nspecial[nlocal][3]
 * (cumulative counts of 1-2, 1-3, 1-4 partners), Atom::special[nlocal][maxspecial] (their tags), Atom::tag[nall]
 * (src/atom.h:93-95) and Neighbor::special_flag[4] (src/neighbor.h:82; 0 drop, 1 ordinary, 2 keep with the class
 * in the entry).  Call before the b200md_host_atoms that rebuilds the list.  The list build then follows
 * NPair::find_special (src/npair.h:125-158) and pair lj/cut applies factor_lj = special_lj[class]
 * (src/pair_lj_cut.cpp:103-105; the weights given to b200md_lj_init).  maxspecial == 0 turns it off. */
int b200md_host_special(b200md_ctx *ctx, int nlocal, int nall, const int *tag, int maxspecial, const int *nspecial,
                        const int *special, const int special_flag[4]);
/* f[nlocal][3] = (accumulate ? f : 0) + device forces of the owned atoms, host order */
int b200md_host_forces(b200md_ctx *ctx, double *f, int accumulate);
/* PairEAM::compute split at its forward_comm (src/MANYBODY/pair_eam.cpp:251), which stays on the host like in
 * PairEAMGPU::compute: passes A+B -> rho[nlocal], fp[nlocal] (either may be NULL) and the embedding energy;
 * the caller runs comm->forward_comm(this) and hands fp[nall] back for pass C. */
int b200md_host_eam_rho(b200md_ctx *ctx, int eflag, double *rho, double *fp, double *eng_embed);
int b200md_host_eam_force(b200md_ctx *ctx, const double *fp_all, int eflag, int vflag, double *eng_pair,
                          double *virial);

/* ---- bench/in.chain additions: bond_style fene, special_bonds fene, fix langevin -------------------
 * They take effect inside b200md_setup / b200md_run (Verlet::setup / run order:
 * pair force, bond force, post_force fixes, final_integrate; src/verlet.cpp:106-160,232-360).
 *
 * b200md_set_bonds: the molecular atom style's per-atom bond arrays in the caller's local order -- num_bond[n],
 * bond_type[n][bond_per_atom], bond_atom[n][bond_per_atom] = partner tag (src/atom.h:108-110), tag[n]; newton_bond
 * says whether each bond is stored once (on, the data-file default) or on both atoms (src/atom.cpp data_bonds).
 * Replaces NTopoBondAll::build (src/ntopo_bond_all.cpp:37-76): the partner's closest image
 * (Domain::closest_image) is resolved inside the neighbor build, which also leaves 1-2 partners out of the pair
 * list (special_bonds fene = lj 0.0 1.0 1.0, src/npair.h:125-158).  A partner farther than the neighbor cutoff
 * is "Bond atoms missing" (B200MD_ERR_LOST_ATOMS at the next build).  Up to 4 bonds per atom.  n = 0 removes them.
 * The library keeps the topology as a table by atom TAG, so on several ranks every rank passes the COMPLETE
 * topology (all atoms of the system, any order), not only the bonds of the atoms it owns; the table then needs no
 * update when atoms migrate. */
int b200md_set_bonds(b200md_ctx *ctx, int n, const int *tag, int bond_per_atom, const int *num_bond,
                     const int *bond_type, const int *bond_atom, int newton_bond);
/* bond_style fene + bond_coeff K R0 epsilon sigma, arrays [nbondtypes+1] indexed by bond type
 * (BondFENE::coeff / compute, src/MOLECULE/bond_fene.cpp:52-165).  "Bad FENE bond" is reported at the
 * next neighbor build as B200MD_ERR_ARG. */
int b200md_bond_fene(b200md_ctx *ctx, int nbondtypes, const double *k, const double *r0, const double *epsilon,
                     const double *sigma);
/* bond energy (thermo E_bond) and bond virial[6] of the last force evaluation that tallied them, summed over
 * the ranks (collective, like b200md_thermo) (Bond::energy / Bond::virial, src/bond.h:33-35 -> compute pe /
 * pressure); b200md_thermo's pressure already includes the bond virial */
int b200md_thermo_bond(b200md_ctx *ctx, double *eng_bond, double *virial);
/* fix langevin t_start t_stop damp seed, plain instance (no zero / tally / gjf / omega / per-type scale;
 * FixLangevin::post_force_templated<0,0,0,0,0,0>, src/fix_langevin.cpp:583-716), dt = update->dt.
 * The random stream is the reference's RanMars(seed + rank) (src/random_mars.cpp) and is consumed in the
 * reference's local atom order, which the caller states with b200md_langevin_order, so a run reproduces the
 * reference's trajectory.  Calling it again with the same seed (every `run`) updates the parameters and keeps
 * the stream where it is, like the fix's own generator; t_period <= 0 with seed 0 removes the fix. */
int b200md_fix_langevin(b200md_ctx *ctx, double t_start, double t_stop, double t_period, int seed, double dt);
/* tag[n] = the owned atoms in the reference's current local order (atom->tag[0..nlocal)); sortfreq / userbinsize
 * = atom_modify sort (defaults 1000, 0.0): the library replays Atom::sort (src/atom.cpp:2203-2290) on that
 * order at setup and every sortfreq steps, as Verlet does */
int b200md_langevin_order(b200md_ctx *ctx, int n, const int *tag, int sortfreq, double userbinsize);
/* update->beginstep / endstep of the run the following b200md_run calls are pieces of (the fix langevin
 * temperature ramp, FixLangevin::compute_target :773-787); unset = each b200md_run call is a whole run */
int b200md_run_bounds(b200md_ctx *ctx, int64_t beginstep, int64_t endstep);

/* random numbers of the thermostat: 0 (default) = the reference's RanMars stream replayed on the host in the
 * reference's atom order (one rank; reproduces the reference's trajectory), 1 = counter-based on the device:
 * Philox4x32-10 keyed by the seed with counter (tag, timestep) -- any number of ranks, no per-step host work,
 * results independent of rank count and atom order; statistically equivalent, not stream-compatible. */
int b200md_langevin_rng(b200md_ctx *ctx, int mode);
/* host only (no context, no device): u[3] of atom `tag` on timestep `step` in counter mode; a raw
 * Philox4x32-10 block for known-answer tests */
int b200md_langevin_counter_uniforms(int seed, int tag, int64_t step, double *u);
int b200md_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
/* RanMars(seed)->uniform() x n (src/random_mars.cpp:30-100): the stream b200md_fix_langevin draws from.
 * Host only: needs neither a context nor a device. */
int b200md_ranmars(int seed, int n, double *out);

/* ---- instrumentation: device time per section in ms since the last reset, named like the
 * reference's Timer (src/timer.h:23-42): [0]=Pair [1]=Neigh [2]=Comm [3]=Modify [4]=Other;
 * counts kernels launched by this library since the last reset. */
int b200md_timers(b200md_ctx *ctx, double ms[5], int64_t *kernel_launches, int reset);
/* opt-in timing level: 0 off; 1 = two CUDA-event records around each timed kernel on the launching stream,
 * no host synchronisation (resolved when read); 2 = also the section timers above (blocking). */
int b200md_profile(b200md_ctx *ctx, int level);
/* summed CUDA-event duration (ms) and launch count per kernel class since the last reset:
 * [0] pair force kernel (lj/cut, or the eam force pass) [1] eam rho+fp pass [2] neighbor-build kernel
 * [3] nve kernels [4] halo gather [5] bin sort + permute.  Arrays of 6. */
int b200md_kernel_times(b200md_ctx *ctx, double *ms, int64_t *count, int reset);
/* the cudaStream_t all of this context's work is issued on (for callers that record their own events
 * or order their own kernels against the library's, as KOKKOS styles do with their execution space) */
void *b200md_stream(b200md_ctx *ctx);
/* bins of an x-row one thread block of the last neighbor build handled (about 112 / <atoms per bin>: 6 for in.lj
 * and in.eam, 37 for in.chain) */
int b200md_build_bins_per_block(const b200md_ctx *ctx);
/* execution knobs of this context -- the analogue of the reference's `package` command options
 * (src/GPU/fix_gpu.cpp:43-51 `package gpu ... neigh/split/...`, src/input.cpp package()).  Defaults are the variants
 * that measured fastest on B200; nothing is read from the environment.
 *   "overlap"        0/1  per-step halo exchange on its own stream under the force kernels of the tiles whose lists
 *                         hold no ghost (default 1 on several ranks, 0 on one); applies from the next list build
 *   "host_pipeline"  0/1  b200md_step_host as a three-stream chunk pipeline (default 1) or serial copy-compute-copy
 *   "hs_min_chunk"   n    smallest system the pipeline splits into ranges (default 65536)
 *   "build_gb" n, "build_fp64" 0/1   test knobs of the list build (bins per block; force the all-FP64 fallback kernel)
 *   "build_bond_general" 0/1         test knob: bonded build with the closest-image choice even where a tag is unique
 *   "build_threads"  128|256  threads per block of the list build (default 128, the faster one on B200)
 *   "halo_put"       0/1  per-step halo as a fused pack-and-put kernel over peer memory (after b200md_halo_put_import;
 *                         default 0: NCCL send/recv)
 *   "fuse_nve"       0/1  fix nve inside the force kernel on steps of b200md_run that neither tally nor end the call
 *                         (default 1; 0 = the separate integrate kernels; results are bit-identical)
 * Unknown key: B200MD_ERR_ARG. */
int b200md_tune(b200md_ctx *ctx, const char *key, int value);

/* No counterpart in the reference: added for callers that may not allocate
 * device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * The kinetic-energy sum of b200md_thermo -- sum_i mass[type[i]] * |v_i|^2 over n atoms, the quantity
 * ComputeTemp::compute_scalar multiplies by mvv2e / (dof * boltz) -- computed without a context and without any
 * allocation: no cudaMalloc / cudaFree / allocating library call.  The per-block partial sums go to a fixed
 * file-scope device array of 65536 doubles (512 KiB, held by every GPU that loads the library until the process
 * exits), so n <= b200md_ke_noalloc_max_atoms() = 16 777 216; a larger n returns B200MD_ERR_ARG with a message in
 * b200md_last_error(NULL).  All pointers are DEVICE pointers on `device`, caller-owned: vx, vy, vz [n], type [n]
 * (1-based), mass [ntypes+1], sum_mv2 [1] (written on `stream`, a cudaStream_t or NULL; the call does not synchronise).
 * Every call on a device shares the one array: calls that overlap on the GPU race -- order them on one stream.  Nothing
 * in the array has to start at zero (each call overwrites every element it later reads).  The array's device address
 * is looked up once per device, on the first call for that device. */
int b200md_ke_noalloc(int device, int n, const double *vx, const double *vy, const double *vz, const int *type,
                      const double *mass, double *sum_mv2, void *stream);
int b200md_ke_noalloc_max_atoms(void);

#ifdef __cplusplus
}
#endif
#endif
