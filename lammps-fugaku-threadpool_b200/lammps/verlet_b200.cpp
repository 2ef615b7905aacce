/* verlet_b200.cpp -- see verlet_b200.h */
#include "verlet_b200.h"

#include "atom.h"
#include "atom_vec.h"
#include "b200_device.h"
#include "comm.h"
#include "domain.h"
#include "error.h"
#include "fix_nve_b200.h"
#include "force.h"
#include "modify.h"
#include "neighbor.h"
#include "output.h"
#include "pair.h"
#include "timer.h"
#include "update.h"

#include <algorithm>
#include <mpi.h>
#include <utility>
#include <vector>

using namespace LAMMPS_NS;

VerletB200::VerletB200(LAMMPS *lmp, int narg, char **arg) : Verlet(lmp, narg, arg), dev(nullptr), nve(nullptr)
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
  if (atom->molecular != Atom::ATOMIC) error->all(FLERR, "Run style verlet/b200 requires atom_style atomic");
  if (!force->pair || !dynamic_cast<B200PairTables *>(force->pair))
    error->all(FLERR, "Run style verlet/b200 requires a pair style with suffix b200 (lj/cut/b200, eam/b200)");
  if (force->kspace || force->bond || force->angle) error->all(FLERR, "Run style verlet/b200: only pair interactions run on the device");
  // exactly one time-integration fix, nve/b200 on group all; nothing else may touch atoms between output steps
  nve = nullptr;
  for (const auto &fix : modify->get_fix_list()) {
    auto *f = dynamic_cast<FixNVEB200 *>(fix);
    if (f && f->igroup == 0 && !nve) nve = f;
    else error->all(FLERR, "Run style verlet/b200 supports exactly one fix, nve/b200 on group all (found fix {} {})", fix->id, fix->style);
  }
  if (!nve) error->all(FLERR, "Run style verlet/b200 requires fix nve/b200 on group all");
  if (comm->nprocs > 1) {
    // the device library places rank r at grid location (r % px, (r / px) % py, r / (px*py)) (include/b200md.h)
    const int *pg = comm->procgrid;
    const int want[3] = {comm->me % pg[0], (comm->me / pg[0]) % pg[1], comm->me / (pg[0] * pg[1])};
    int bad = (want[0] != comm->myloc[0] || want[1] != comm->myloc[1] || want[2] != comm->myloc[2]), anybad = 0;
    MPI_Allreduce(&bad, &anybad, 1, MPI_INT, MPI_MAX, world);
    if (anybad) error->all(FLERR, "Run style verlet/b200 needs x-fastest rank placement: use 'processors * * * map xyz'");
  }
}

void VerletB200::upload()
{
  auto *tables = dynamic_cast<B200PairTables *>(force->pair);
  B200_CHECK(dev, b200md_set_box(dev->ctx, domain->boxlo, domain->boxhi, domain->periodicity, comm->procgrid));
  tables->device_tables(dev);    // also units, masses and neighbor settings
  const int n = atom->nlocal;
  B200_CHECK(dev, b200md_upload_atoms(dev->ctx, n, &atom->x[0][0], &atom->v[0][0], atom->type, atom->tag, atom->image));
}

void VerletB200::download()
{
  const int n = b200md_nlocal(dev->ctx);
  std::vector<double> x((size_t) 3 * n), v((size_t) 3 * n), f((size_t) 3 * n);
  std::vector<int> type(n), tag(n), image(n);
  B200_CHECK(dev, b200md_download_atoms(dev->ctx, x.data(), v.data(), f.data(), type.data(), tag.data(), image.data()));
  if (comm->nprocs == 1) {
    // one rank: every atom is still local; write back by tag so host order, mask and any other per-atom
    // property stay what they were
    if (n != atom->nlocal) error->one(FLERR, "Run style verlet/b200: device holds {} atoms, host {}", n, atom->nlocal);
    std::vector<std::pair<tagint, int>> where(n);
    for (int i = 0; i < n; i++) where[i] = {atom->tag[i], i};
    std::sort(where.begin(), where.end());
    for (int s = 0; s < n; s++) {
      auto it = std::lower_bound(where.begin(), where.end(), std::make_pair((tagint) tag[s], -1));
      if (it == where.end() || it->first != tag[s]) error->one(FLERR, "Run style verlet/b200: unknown atom tag {} came back", tag[s]);
      const int i = it->second;
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
  upload();
  dev->resident = 1;
  B200_CHECK(dev, b200md_setup(dev->ctx));
  Pair *pair = force->pair;

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
      double e = 0.0, vir[6] = {0, 0, 0, 0, 0, 0};
      download();
      B200_CHECK(dev, b200md_last_ev(dev->ctx, &e, vir));
      pair->eng_vdwl = e;
      pair->eng_coul = 0.0;
      for (int k = 0; k < 6; k++) pair->virial[k] = vir[k];
      update->eflag_global = update->vflag_global = update->ntimestep;
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
