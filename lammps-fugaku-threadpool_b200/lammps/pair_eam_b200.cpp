/* pair_eam_b200.cpp -- see pair_eam_b200.h.  Host-ghost mode follows PairEAMGPU::compute
   (src/GPU/pair_eam_gpu.cpp:96-139): device passes A+B (rho, fp), fp comes down, comm->forward_comm(this) runs on
   the host with the inherited pack/unpack_forward_comm (src/MANYBODY/pair_eam.cpp:855-876), fp of the ghosts goes
   up, device pass C.  The device list is a full list, so neither of the reference's two reverse_comms exists. */
#include "pair_eam_b200.h"

#include "atom.h"
#include "comm.h"
#include "domain.h"
#include "error.h"
#include "force.h"
#include "memory.h"
#include "neighbor.h"
#include "update.h"

#include <vector>

using namespace LAMMPS_NS;

PairEAMB200::PairEAMB200(LAMMPS *lmp) : PairEAM(lmp), dev(nullptr), tables_loaded(0)
{
  respa_enable = 0;
  no_virial_fdotr_compute = 1;
  dev = B200Device::instance(lmp);
}

PairEAMB200::~PairEAMB200()
{
  B200Device::release(lmp);
}

void PairEAMB200::init_style()
{
  // convert read-in file(s) to arrays and spline them (PairEAM::init_style, src/MANYBODY/pair_eam.cpp:419-431),
  // without requesting a host neighbor list
  file2array();
  array2spline();
  embedstep = -1;
  if (atom->molecular != Atom::ATOMIC) error->all(FLERR, "Pair style eam/b200 requires atom_style atomic");
  if (dev->want_host_list) neighbor->add_request(this, NeighConst::REQ_FULL);    // see PairLJCutB200::init_style
  tables_loaded = 0;
  dev->announce("pair eam/b200");
}

void PairEAMB200::device_tables(B200Device *d)
{
  const int nt = atom->ntypes, nt1 = nt + 1;
  std::vector<int> t2f(nt1, 0), t2r((size_t) nt1 * nt1, 0), t2z((size_t) nt1 * nt1, 0);
  for (int i = 1; i <= nt; i++) {
    t2f[i] = type2frho[i];
    for (int j = 1; j <= nt; j++) {
      t2r[(size_t) i * nt1 + j] = type2rhor[i][j];
      t2z[(size_t) i * nt1 + j] = type2z2r[i][j];
      if (scale[i][j] != 1.0) error->all(FLERR, "Pair style eam/b200 does not support scaled interactions");
    }
  }
  B200_CHECK(d, b200md_eam_init(d->ctx, nt, nrho, nr, rdrho, rdr, rhomax, cutforcesq, nfrho, &frho_spline[0][0][0], nrhor,
                                &rhor_spline[0][0][0], nz2r, &z2r_spline[0][0][0], t2f.data(), t2r.data(), t2z.data()));
  d->sync_settings();
  tables_loaded = 1;
}

void PairEAMB200::compute(int eflag, int vflag)
{
  ev_init(eflag, vflag);
  if (dev->resident) return;
  if (!tables_loaded) device_tables(dev);

  const int nlocal = atom->nlocal, nghost = atom->nghost;
  if (nlocal == 0) return;
  if (atom->nmax > nmax) {    // PairEAM::compute grows rho/fp/numforce the same way (:156-165)
    memory->destroy(rho);
    memory->destroy(fp);
    memory->destroy(numforce);
    nmax = atom->nmax;
    memory->create(rho, nmax, "pair:rho");
    memory->create(fp, nmax, "pair:fp");
    memory->create(numforce, nmax, "pair:numforce");
  }
  B200_CHECK(dev, b200md_host_atoms(dev->ctx, nlocal, nghost, &atom->x[0][0], atom->type, domain->sublo, domain->subhi,
                                    neighbor->ago == 0 && !dev->host_list_fresh));
  dev->host_list_fresh = 0;
  double e_embed = 0.0, e_pair = 0.0, v[6] = {0, 0, 0, 0, 0, 0};
  B200_CHECK(dev, b200md_host_eam_rho(dev->ctx, eflag_global, rho, fp, &e_embed));
  comm->forward_comm(this);    // ghost fp from their owners, on the host (pair_eam.cpp:251)
  embedstep = update->ntimestep;
  B200_CHECK(dev, b200md_host_eam_force(dev->ctx, fp, eflag_global, vflag_global, &e_pair, v));
  B200_CHECK(dev, b200md_host_forces(dev->ctx, &atom->f[0][0], 1));
  if (eflag_global) eng_vdwl += e_embed + e_pair;
  if (vflag_global)
    for (int k = 0; k < 6; k++) virial[k] += v[k];
  if (eflag_atom || vflag_atom) dev->peratom_to_host(nlocal, eflag_atom ? eatom : nullptr, vflag_atom ? vatom : nullptr);
}

double PairEAMB200::memory_usage()
{
  return PairEAM::memory_usage() + b200md_memory_usage(dev->ctx);
}
