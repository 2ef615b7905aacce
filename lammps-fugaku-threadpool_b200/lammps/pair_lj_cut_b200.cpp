/* pair_lj_cut_b200.cpp -- see pair_lj_cut_b200.h.  Host-ghost mode (stock `run_style verlet`): LAMMPS keeps
   atoms, ghosts, exchange/borders and the rebuild decision; compute() ships x up and f down, the device builds
   its own (full) neighbor list when the host reneighbored -- the contract of PairLJCutGPU::compute
   (src/GPU/pair_lj_cut_gpu.cpp:85-118).  Under `run_style verlet/b200` the atoms are device resident and this
   hook is idle (VerletB200 drives the device loop and publishes eng_vdwl / virial through this object). */
#include "pair_lj_cut_b200.h"

#include "atom.h"
#include "domain.h"
#include "error.h"
#include "force.h"
#include "neighbor.h"
#include "suffix.h"
#include "update.h"

#include <cstring>
#include <vector>

using namespace LAMMPS_NS;

PairLJCutB200::PairLJCutB200(LAMMPS *lmp) : PairLJCut(lmp), dev(nullptr), tables_loaded(0)
{
  respa_enable = 0;
  no_virial_fdotr_compute = 1;    // the device returns the virial; no ghost forces exist to sum f.x over
  dev = B200Device::instance(lmp);
}

PairLJCutB200::~PairLJCutB200()
{
  B200Device::release(lmp);
}

void PairLJCutB200::init_style()
{
  if (!force->newton_pair && comm_forward) {}    // newton on/off both fine: the device list is a full list
  // molecular systems: under the stock loop the device list build applies the host's special-bond lists
  // (b200md_host_special: any special_bonds weights); the device-resident loop serves fene bonds with
  // special_bonds fene only (VerletB200::init checks)
  if (atom->molecular == Atom::TEMPLATE) error->all(FLERR, "Pair style lj/cut/b200 does not support atom_style template");
  cut_respa = nullptr;    // PairLJCut::init_style sets it (its constructor does not); init_one reads it
  // no host NeighList is requested: Neighbor still runs decide/exchange/borders, the list lives on the device.
  // B200MD_HOST_LIST=1 (atomic systems): a full list is requested and npair full/bin/atomonly/b200 fills it from the
  // device list at every rebuild, for lammps_neighlist_element_neighbors and other host-side readers
  if (dev->want_host_list && atom->molecular == Atom::ATOMIC) neighbor->add_request(this, NeighConst::REQ_FULL);
  tables_loaded = 0;
  dev->announce("pair lj/cut/b200");
}

void PairLJCutB200::device_tables(B200Device *d)
{
  const int nt = atom->ntypes, nt1 = nt + 1;
  std::vector<double> t[6];
  double **src[6] = {cutsq, lj1, lj2, lj3, lj4, offset};
  for (int k = 0; k < 6; k++) {
    t[k].assign((size_t) nt1 * nt1, 0.0);
    for (int i = 1; i <= nt; i++)
      for (int j = 1; j <= nt; j++) t[k][(size_t) i * nt1 + j] = src[k][i][j];
  }
  B200_CHECK(d, b200md_lj_init(d->ctx, nt, t[0].data(), t[1].data(), t[2].data(), t[3].data(), t[4].data(), t[5].data(),
                               force->special_lj));
  d->sync_settings();
  tables_loaded = 1;
}

void PairLJCutB200::compute(int eflag, int vflag)
{
  ev_init(eflag, vflag);
  if (dev->resident) return;
  if (!tables_loaded) device_tables(dev);

  const int nlocal = atom->nlocal, nghost = atom->nghost;
  if (nlocal == 0) return;
  if (neighbor->ago == 0 && !dev->host_list_fresh) {
    // the special-bond lists as Neighbor would hand them to NPair::find_special (src/npair.h:125-158); tagint is int
    // (LAMMPS_SMALLBIG) and Atom's 2-d arrays are contiguous (Memory::create)
    if (atom->molecular == Atom::MOLECULAR && atom->maxspecial > 0)
      B200_CHECK(dev, b200md_host_special(dev->ctx, nlocal, nlocal + nghost, atom->tag, atom->maxspecial, &atom->nspecial[0][0],
                                          &atom->special[0][0], neighbor->special_flag));
    else
      B200_CHECK(dev, b200md_host_special(dev->ctx, nlocal, nlocal + nghost, nullptr, 0, nullptr, nullptr, nullptr));
  }
  B200_CHECK(dev, b200md_host_atoms(dev->ctx, nlocal, nghost, &atom->x[0][0], atom->type, domain->sublo, domain->subhi,
                                    neighbor->ago == 0 && !dev->host_list_fresh));
  dev->host_list_fresh = 0;
  double e = 0.0, v[6] = {0, 0, 0, 0, 0, 0};
  B200_CHECK(dev, b200md_pair_compute(dev->ctx, eflag_global, vflag_global, &e, v));
  B200_CHECK(dev, b200md_host_forces(dev->ctx, &atom->f[0][0], 1));
  if (eflag_global) eng_vdwl += e;
  if (vflag_global)
    for (int k = 0; k < 6; k++) virial[k] += v[k];
  if (eflag_atom || vflag_atom) dev->peratom_to_host(nlocal, eflag_atom ? eatom : nullptr, vflag_atom ? vatom : nullptr);
}

double PairLJCutB200::memory_usage()
{
  return PairLJCut::memory_usage() + b200md_memory_usage(dev->ctx);
}
