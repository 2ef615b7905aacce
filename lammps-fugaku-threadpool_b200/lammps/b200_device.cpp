/* b200_device.cpp -- see b200_device.h */
#include "b200_device.h"

#include "atom.h"
#include "comm.h"
#include "error.h"
#include "force.h"
#include "memory.h"
#include "neighbor.h"
#include "pair.h"
#include "update.h"

#include <cmath>
#include <cstdlib>
#include <map>
#include <mpi.h>
#include <vector>

using namespace LAMMPS_NS;

static std::map<LAMMPS *, B200Device *> &registry()
{
  static std::map<LAMMPS *, B200Device *> r;
  return r;
}

B200Device *B200Device::instance(LAMMPS *lmp)
{
  auto &r = registry();
  auto it = r.find(lmp);
  if (it == r.end()) it = r.insert({lmp, new B200Device(lmp)}).first;
  it->second->refs++;
  return it->second;
}

B200Device *B200Device::peek(LAMMPS *lmp)
{
  auto &r = registry();
  auto it = r.find(lmp);
  return it == r.end() ? nullptr : it->second;
}

void B200Device::release(LAMMPS *lmp)
{
  auto &r = registry();
  auto it = r.find(lmp);
  if (it == r.end()) return;
  if (--it->second->refs <= 0) {
    delete it->second;
    r.erase(it);
  }
}

B200Device::B200Device(LAMMPS *lmp)
    : Pointers(lmp), ctx(nullptr), resident(0), resident_step(-1), want_host_list(0), host_list_fresh(0), refs(0)
{
  if (const char *h = getenv("B200MD_HOST_LIST")) want_host_list = atoi(h) != 0;
  // one context per MPI rank; the NCCL id is minted on rank 0 and broadcast over `world`
  // (what replaces the MPI communicator of Comm on the device side, include/b200md.h)
  int me = comm->me, nprocs = comm->nprocs;
  char id[B200MD_NCCL_ID_BYTES] = {0};
  if (nprocs > 1) {
    int st = (me == 0) ? b200md_nccl_unique_id(id) : 0;
    MPI_Bcast(&st, 1, MPI_INT, 0, world);
    if (st) error->all(FLERR, "b200: could not create the NCCL unique id: {}", b200md_last_error(nullptr));
    MPI_Bcast(id, B200MD_NCCL_ID_BYTES, MPI_CHAR, 0, world);
  }
  int device = 0;
  if (const char *e = getenv("B200MD_DEVICE")) device = atoi(e);
  else if (const char *l = getenv("OMPI_COMM_WORLD_LOCAL_RANK")) device = atoi(l);
  else if (const char *l2 = getenv("SLURM_LOCALID")) device = atoi(l2);
  device_id = device;
  int st = b200md_create(&ctx, device, nprocs, me, nprocs > 1 ? id : nullptr);
  if (st != B200MD_OK)
    error->one(FLERR, "b200: cannot create the device context (there is no CPU fallback): {}", b200md_last_error(ctx));
}

B200Device::~B200Device()
{
  if (ctx) b200md_destroy(ctx);
}

void B200Device::announce(const char *what)
{
  if (comm->me == 0) utils::logmesg(lmp, "b200: {} on CUDA device {} (libb200md, sm_100a)\n", what, device_id);
}

void B200Device::check(int status, const char *file, int line)
{
  if (status == B200MD_OK) return;
  // same texts as the reference raises on its own path (include/b200md.h status table)
  error->one(file, line, "b200 device library: {}", b200md_last_error(ctx));
}

void B200Device::sync_settings()
{
  B200_CHECK(this, b200md_set_units(ctx, force->boltz, force->mvv2e, force->ftm2v, force->nktv2p));
  const int nt = atom->ntypes, nt1 = nt + 1;
  if (nt > 16) error->all(FLERR, "b200 styles support at most 16 atom types");
  // Neighbor::init (src/neighbor.cpp:306-360): cutneighsq[i][j] = (cut_ij + skin)^2
  std::vector<double> cn((size_t) nt1 * nt1, 0.0);
  for (int i = 1; i <= nt; i++)
    for (int j = 1; j <= nt; j++) cn[(size_t) i * nt1 + j] = neighbor->cutneighsq[i][j];
  B200_CHECK(this, b200md_neigh_config(ctx, nt, cn.data(), neighbor->cutneighmax, neighbor->skin, neighbor->every,
                                       neighbor->delay, neighbor->dist_check, 0.0));
  if (atom->mass) {
    std::vector<double> m(nt1, 1.0);
    for (int i = 1; i <= nt; i++) m[i] = atom->mass[i];
    B200_CHECK(this, b200md_set_mass(ctx, nt, m.data()));
  }
}

void B200Device::peratom_to_host(int nlocal, double *eatom, double **vatom)
{
  std::vector<double> e, v;
  if (eatom) e.resize(nlocal);
  if (vatom) v.resize((size_t) 6 * nlocal);
  check(b200md_pair_peratom(ctx), FLERR);
  check(b200md_peratom_download(ctx, eatom ? e.data() : nullptr, vatom ? v.data() : nullptr), FLERR);
  for (int i = 0; i < nlocal; i++) {
    if (eatom) eatom[i] += e[i];
    if (vatom)
      for (int k = 0; k < 6; k++) vatom[i][k] += v[(size_t) 6 * i + k];
  }
}
