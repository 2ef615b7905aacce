/* verlet_b200.cpp -- see verlet_b200.h */
#include "verlet_b200.h"

#include "atom.h"
#include "atom_vec.h"
#include "b200_device.h"
#include "bond_fene_b200.h"
#include "comm.h"
#include "domain.h"
#include "error.h"
#include "fix_langevin_b200.h"
#include "fix_nve_b200.h"
#include "force.h"
#include "modify.h"
#include "neighbor.h"
#include "output.h"
#include "pair.h"
#include "timer.h"
#include "update.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mpi.h>
#include <utility>
#include <vector>

using namespace LAMMPS_NS;

VerletB200::VerletB200(LAMMPS *lmp, int narg, char **arg) :
    Verlet(lmp, narg, arg), dev(nullptr), nve(nullptr), bond(nullptr), langevin(nullptr), device_setup_step(-1)
{
  dev = B200Device::instance(lmp);
}

VerletB200::~VerletB200()
{
  B200Device::release(lmp);
}

void VerletB200::init()
{
  Verlet::init();
  if (domain->triclinic) error->all(FLERR, "Run style verlet/b200 requires an orthogonal box");
  if (!force->pair || !dynamic_cast<B200PairTables *>(force->pair))
    error->all(FLERR, "Run style verlet/b200 requires a pair style with suffix b200 (lj/cut/b200, eam/b200)");
  if (force->kspace || force->angle || force->dihedral || force->improper)
    error->all(FLERR, "Run style verlet/b200: only pair and fene bond interactions run on the device");
  // atom_style atomic, or a bonded style whose only topology is fene/b200 bonds with special_bonds fene (in.chain)
  bond = nullptr;
  if (atom->molecular == Atom::TEMPLATE) error->all(FLERR, "Run style verlet/b200 does not support atom_style template");
  if (atom->molecular == Atom::MOLECULAR) {
    bond = dynamic_cast<BondFENEB200 *>(force->bond);
    if (force->bond && !bond) error->all(FLERR, "Run style verlet/b200 requires bond style fene/b200");
    if (atom->nangles || atom->ndihedrals || atom->nimpropers)
      error->all(FLERR, "Run style verlet/b200: angles, dihedrals and impropers do not run on the device");
    if (atom->nbonds && !bond) error->all(FLERR, "Run style verlet/b200: the system has bonds but no bond style fene/b200");
    if (bond) {
      const double *sl = force->special_lj;
      if (sl[1] != 0.0 || sl[2] != 1.0 || sl[3] != 1.0)
        error->all(FLERR, "Run style verlet/b200 requires special_bonds fene (lj 0.0 1.0 1.0) with bond style fene/b200");
      if (!bond->all_set()) error->all(FLERR, "All bond coeffs are not set");
      if (atom->bond_per_atom > 4) error->all(FLERR, "Run style verlet/b200 holds up to 4 bonds per atom");
    }
  } else if (force->bond)
    error->all(FLERR, "Run style verlet/b200: bond style without a bonded atom style");
  // one time-integration fix, nve/b200 on group all, and optionally one langevin/b200 thermostat; nothing else may
  // touch atoms between output steps
  nve = nullptr;
  langevin = nullptr;
  for (const auto &fix : modify->get_fix_list()) {
    auto *f = dynamic_cast<FixNVEB200 *>(fix);
    auto *l = dynamic_cast<FixLangevinB200 *>(fix);
    if (f && f->igroup == 0 && !nve) nve = f;
    else if (l && !langevin) {
      langevin = l;    // its keywords are checked in upload(): FixLangevin::init has not run yet (Update before Modify)
    } else
      error->all(FLERR, "Run style verlet/b200 supports fix nve/b200 on group all plus one fix langevin/b200 (found fix {} {})", fix->id, fix->style);
  }
  if (!nve) error->all(FLERR, "Run style verlet/b200 requires fix nve/b200 on group all");
  if (bond && comm->nprocs > 1) error->all(FLERR, "Run style verlet/b200: bond style fene/b200 runs on one rank only for now");
  if (comm->nprocs > 1) {
    // the device library places rank r at grid location (r % px, (r / px) % py, r / (px*py)) (include/b200md.h)
    const int *pg = comm->procgrid;
    const int want[3] = {comm->me % pg[0], (comm->me / pg[0]) % pg[1], comm->me / (pg[0] * pg[1])};
    int bad = (want[0] != comm->myloc[0] || want[1] != comm->myloc[1] || want[2] != comm->myloc[2]), anybad = 0;
    MPI_Allreduce(&bad, &anybad, 1, MPI_INT, MPI_MAX, world);
    if (anybad) error->all(FLERR, "Run style verlet/b200 needs x-fastest rank placement: use 'processors * * * map xyz'");
  }
  dev->announce("run_style verlet/b200 (device-resident loop)");
}

void VerletB200::upload()
{
  auto *tables = dynamic_cast<B200PairTables *>(force->pair);
  B200_CHECK(dev, b200md_set_box(dev->ctx, domain->boxlo, domain->boxhi, domain->periodicity, comm->procgrid));
  tables->device_tables(dev);    // also units, masses and neighbor settings
  const int n = atom->nlocal;
  B200_CHECK(dev, b200md_upload_atoms(dev->ctx, n, &atom->x[0][0], &atom->v[0][0], atom->type, atom->tag, atom->image));
  B200_CHECK(dev, b200md_set_timestep(dev->ctx, update->ntimestep));
  if (bond && n > 0) {
    // Atom's bond arrays are contiguous n x bond_per_atom blocks (Memory::create); tagint is int (LAMMPS_SMALLBIG)
    B200_CHECK(dev, b200md_set_bonds(dev->ctx, n, atom->tag, atom->bond_per_atom, atom->num_bond, &atom->bond_type[0][0],
                                     &atom->bond_atom[0][0], force->newton_bond));
    B200_CHECK(dev, b200md_bond_fene(dev->ctx, atom->nbondtypes, bond->dev_k(), bond->dev_r0(), bond->dev_epsilon(), bond->dev_sigma()));
  } else
    B200_CHECK(dev, b200md_set_bonds(dev->ctx, 0, nullptr, 0, nullptr, nullptr, nullptr, 1));
  if (langevin) {
    if (const char *why = langevin->device_unsupported())
      error->all(FLERR, "Fix langevin/b200 under run style verlet/b200 does not support {}", why);
    // one rank: the reference's own RanMars stream in the reference's atom order (same trajectory as `fix langevin`);
    // several ranks, or B200MD_LANGEVIN_RNG=counter: the counter-based device generator (same statistics)
    const char *rng = getenv("B200MD_LANGEVIN_RNG");
    const int counter = comm->nprocs > 1 || (rng && strcmp(rng, "counter") == 0);
    if (counter && comm->nprocs > 1 && comm->me == 0)
      error->warning(FLERR, "Fix langevin/b200 on several ranks uses the counter-based device generator, not the RanMars streams of fix langevin");
    B200_CHECK(dev, b200md_langevin_rng(dev->ctx, counter));
    B200_CHECK(dev, b200md_fix_langevin(dev->ctx, langevin->dev_t_start(), langevin->dev_t_stop(), langevin->dev_t_period(),
                                        langevin->dev_seed(), update->dt));
    if (!langevin->order_sent) {
      B200_CHECK(dev, b200md_langevin_order(dev->ctx, n, atom->tag, atom->sortfreq, atom->userbinsize));
      langevin->order_sent = true;
    }
    B200_CHECK(dev, b200md_run_bounds(dev->ctx, update->beginstep, update->endstep));
  } else
    B200_CHECK(dev, b200md_fix_langevin(dev->ctx, 0.0, 0.0, 0.0, 0, 0.0));
}

// pair (and bond) energy / virial of the device's last tallied force evaluation -> the style objects the
// reference's computes read (compute pe, compute pressure)
void VerletB200::fetch_energies()
{
  double e = 0.0, vir[6] = {0, 0, 0, 0, 0, 0};
  Pair *pair = force->pair;
  // what the computes of this step want (Integrate::ev_set); while the device is resident the pair style's compute()
  // only does its ev_init: sizes and clears eatom / vatom and sets the eflag_* / vflag_* members
  ev_set(update->ntimestep);
  pair->compute(eflag, vflag);
  B200_CHECK(dev, b200md_last_ev(dev->ctx, &e, vir));
  pair->eng_vdwl = e;
  pair->eng_coul = 0.0;
  for (int k = 0; k < 6; k++) pair->virial[k] = vir[k];
  if (bond) {
    B200_CHECK(dev, b200md_thermo_bond(dev->ctx, &e, vir));
    bond->energy = e;
    for (int k = 0; k < 6; k++) bond->virial[k] = vir[k];
  }
  update->eflag_global = update->vflag_global = update->ntimestep;
  // per-atom tallies if a compute wants them on this step: the device pass, scattered by tag
  if ((eflag & ENERGY_ATOM) || (vflag & VIRIAL_ATOM)) {
    if (comm->nprocs > 1) error->all(FLERR, "Run style verlet/b200: per-atom energy/virial on one rank only for now");
    const int n = b200md_nlocal(dev->ctx);
    std::vector<double> e(n), v((size_t) 6 * n);
    std::vector<int> tag(n);
    B200_CHECK(dev, b200md_pair_peratom(dev->ctx));
    B200_CHECK(dev, b200md_peratom_download(dev->ctx, e.data(), v.data()));
    B200_CHECK(dev, b200md_download_atoms(dev->ctx, nullptr, nullptr, nullptr, nullptr, tag.data(), nullptr));
    std::vector<std::pair<tagint, int>> where(atom->nlocal);
    for (int i = 0; i < atom->nlocal; i++) where[i] = {atom->tag[i], i};
    std::sort(where.begin(), where.end());
    for (int s = 0; s < n; s++) {
      auto it = std::lower_bound(where.begin(), where.end(), std::make_pair((tagint) tag[s], -1));
      if (it == where.end() || it->first != tag[s]) error->one(FLERR, "Run style verlet/b200: unknown atom tag {} came back", tag[s]);
      const int i = it->second;
      if (pair->eflag_atom) pair->eatom[i] = e[s];
      if (pair->vflag_atom)
        for (int k = 0; k < 6; k++) pair->vatom[i][k] = v[(size_t) 6 * s + k];
    }
    update->eflag_atom = update->vflag_atom = update->ntimestep;
  }
}

// Verlet::setup (src/verlet.cpp:90-160).  Atomic systems: the stock routine.  Bonded / thermostatted systems: the
// device does the whole setup (pbc, sort, ghosts, build, pair + bond forces, FixLangevin::setup's post_force) and the
// result comes back for the step-0 output.
void VerletB200::setup(int flag)
{
  device_setup_step = -1;
  if (!bond && !langevin) {
    Verlet::setup(flag);
    return;
  }
  if (comm->me == 0 && screen) fputs("Setting up Verlet/b200 run on the device ...\n", screen);
  update->setupflag = 1;
  atom->setup();
  ev_set(update->ntimestep);
  upload();
  dev->resident = 1;
  B200_CHECK(dev, b200md_setup(dev->ctx));
  download();
  fetch_energies();
  neighbor->ncalls = 0;
  modify->setup(vflag);    // fix langevin/b200 and nve/b200 do nothing here while the device is resident
  dev->resident = 0;
  device_setup_step = update->ntimestep;
  output->setup(flag);
  update->setupflag = 0;
}

void VerletB200::download()
{
  const int n = b200md_nlocal(dev->ctx);
  if (comm->nprocs == 1 && n == atom->nlocal && n > 0) {
    // one rank, reasonably dense tags (the usual 1..N): the device permutes into the host's own order and the copies
    // land straight in Atom's arrays (Memory::create makes x, v, f contiguous n x 3 blocks; tagint and imageint are int
    // under LAMMPS_SMALLBIG) -- no staging vectors, no scatter loop on the host
    tagint maxtag = 0;
    for (int i = 0; i < n; i++) maxtag = std::max(maxtag, atom->tag[i]);
    if (maxtag > 0 && (bigint) maxtag <= 4 * (bigint) n + 1024) {
      B200_CHECK(dev, b200md_download_atoms_by_tag(dev->ctx, n, atom->tag, (int) maxtag, &atom->x[0][0], &atom->v[0][0], &atom->f[0][0],
                                                   atom->image));
      return;
    }
  }
  std::vector<double> x((size_t) 3 * n), v((size_t) 3 * n), f((size_t) 3 * n);
  std::vector<int> type(n), tag(n), image(n);
  B200_CHECK(dev, b200md_download_atoms(dev->ctx, x.data(), v.data(), f.data(), type.data(), tag.data(), image.data()));
  if (comm->nprocs == 1) {
    // one rank: every atom is still local; write back by tag so host order, mask and any other per-atom
    // property stay what they were
    if (n != atom->nlocal) error->one(FLERR, "Run style verlet/b200: device holds {} atoms, host {}", n, atom->nlocal);
    // tag -> host index: a direct table when the tags are reasonably dense (the usual 1..N), else a sorted list
    tagint maxtag = 0;
    for (int i = 0; i < n; i++) maxtag = std::max(maxtag, atom->tag[i]);
    const bool dense = maxtag > 0 && (bigint) maxtag <= 4 * (bigint) n + 1024;
    std::vector<int> slot;
    std::vector<std::pair<tagint, int>> where;
    if (dense) {
      slot.assign((size_t) maxtag + 1, -1);
      for (int i = 0; i < n; i++) slot[atom->tag[i]] = i;
    } else {
      where.resize(n);
      for (int i = 0; i < n; i++) where[i] = {atom->tag[i], i};
      std::sort(where.begin(), where.end());
    }
    for (int s = 0; s < n; s++) {
      int i = -1;
      if (dense) {
        if (tag[s] >= 1 && tag[s] <= maxtag) i = slot[tag[s]];
      } else {
        auto it = std::lower_bound(where.begin(), where.end(), std::make_pair((tagint) tag[s], -1));
        if (it != where.end() && it->first == tag[s]) i = it->second;
      }
      if (i < 0) error->one(FLERR, "Run style verlet/b200: unknown atom tag {} came back", tag[s]);
      for (int k = 0; k < 3; k++) {
        atom->x[i][k] = x[3 * s + k];
        atom->v[i][k] = v[3 * s + k];
        atom->f[i][k] = f[3 * s + k];
      }
      atom->image[i] = image[s];
    }
  } else {
    // several ranks: atoms migrated between the devices (Comm::exchange ran there); adopt the device's owned set
    const int mask = atom->nlocal ? atom->mask[0] : 1;
    while (atom->nmax < n) atom->avec->grow(0);
    for (int i = 0; i < n; i++) {
      for (int k = 0; k < 3; k++) {
        atom->x[i][k] = x[3 * i + k];
        atom->v[i][k] = v[3 * i + k];
        atom->f[i][k] = f[3 * i + k];
      }
      atom->type[i] = type[i];
      atom->tag[i] = tag[i];
      atom->image[i] = image[i];
      atom->mask[i] = mask;
    }
    atom->nlocal = n;
    atom->nghost = 0;
    if (atom->map_style != Atom::MAP_NONE) {
      atom->map_init();
      atom->map_set();
    }
  }
}

void VerletB200::run(int n)
{
  if (n <= 0) return;
  if (device_setup_step == update->ntimestep) {
    // setup() above already put this state on the device: do not set up twice (the langevin stream would advance)
    dev->resident = 1;
  } else {
    upload();
    dev->resident = 1;
    B200_CHECK(dev, b200md_setup(dev->ctx));
  }
  device_setup_step = -1;

  bigint done = 0;
  while (done < n) {
    if (timer->check_timeout((int) done)) {
      update->nsteps = (int) done;
      break;
    }
    bigint chunk = n - done;
    if (output->next > update->ntimestep && output->next - update->ntimestep < chunk) chunk = output->next - update->ntimestep;
    timer->stamp();
    // energy and virial are tallied on the last step of the call = the output step (Integrate::ev_set)
    B200_CHECK(dev, b200md_run(dev->ctx, (int) chunk, update->dt, 0));
    timer->stamp(Timer::PAIR);
    done += chunk;
    update->ntimestep += chunk;

    if (update->ntimestep == output->next || done == n) {
      download();
      fetch_energies();
      timer->stamp(Timer::COMM);
      if (update->ntimestep == output->next) {
        output->write(update->ntimestep);
        timer->stamp(Timer::OUTPUT);
      }
    }
  }
  int64_t nbuilds = 0, ndanger = 0;
  b200md_neigh_stats(dev->ctx, &nbuilds, &ndanger, nullptr);
  neighbor->ncalls = nbuilds;
  neighbor->ndanger = ndanger;
  dev->resident = 0;
}
