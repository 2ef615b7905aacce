/* npair_b200.cpp -- see npair_b200.h */
#include "npair_b200.h"

#include "atom.h"
#include "b200_device.h"
#include "domain.h"
#include "error.h"
#include "my_page.h"
#include "neigh_list.h"
#include "pair.h"

#include <cstring>
#include <vector>

using namespace LAMMPS_NS;

NPairFullBinAtomonlyB200::NPairFullBinAtomonlyB200(LAMMPS *lmp) : NPairFullBinAtomonly(lmp), announced(0) {}

void NPairFullBinAtomonlyB200::build(NeighList *list)
{
  B200Device *dev = B200Device::peek(lmp);
  B200PairTables *pt = nullptr;
  if (dev && !dev->resident && list->requestor_type == NeighList::PAIR)
    pt = dynamic_cast<B200PairTables *>(static_cast<Pair *>(list->requestor));
  if (!pt) {    // anybody else's list: the reference's own algorithm
    NPairFullBinAtomonly::build(list);
    return;
  }
  if (!announced) {
    dev->announce("npair full/bin/atomonly/b200 (host NeighList filled from the device list)");
    announced = 1;
  }
  const int nlocal = atom->nlocal, nghost = atom->nghost;
  list->inum = 0;
  list->gnum = 0;
  if (nlocal == 0) return;
  pt->device_tables_once(dev);
  // what the pair style's compute() would do on this step (neighbor->ago == 0); it finds the list current and skips it
  B200_CHECK(dev, b200md_host_special(dev->ctx, nlocal, nlocal + nghost, nullptr, 0, nullptr, nullptr, nullptr));
  B200_CHECK(dev, b200md_host_atoms(dev->ctx, nlocal, nghost, &atom->x[0][0], atom->type, domain->sublo, domain->subhi, 1));
  dev->host_list_fresh = 1;

  // host order == device order in host-ghost mode: entries are local indices of owned and ghost atoms as they are
  std::vector<int> nn((size_t) nlocal);
  int64_t total = 0;
  B200_CHECK(dev, b200md_neigh_export(dev->ctx, nn.data(), nullptr, &total));
  std::vector<int> nb((size_t) (total > 0 ? total : 1));
  B200_CHECK(dev, b200md_neigh_export(dev->ctx, nn.data(), nb.data(), &total));

  int *ilist = list->ilist, *numneigh = list->numneigh;
  int **firstneigh = list->firstneigh;
  MyPage<int> *ipage = list->ipage;
  ipage->reset();
  size_t off = 0;
  for (int i = 0; i < nlocal; i++) {
    int *neighptr = ipage->vget();
    const int n = nn[i];
    if (n > 0) memcpy(neighptr, nb.data() + off, (size_t) n * sizeof(int));
    off += (size_t) n;
    ilist[i] = i;
    firstneigh[i] = neighptr;
    numneigh[i] = n;
    ipage->vgot(n);
    if (ipage->status()) error->one(FLERR, "Neighbor list overflow, boost neigh_modify one");
  }
  list->inum = nlocal;
}
