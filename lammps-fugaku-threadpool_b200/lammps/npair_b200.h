/* npair style full/bin/atomonly/b200 -- the device neighbor list as a host-visible NeighList (src/neigh_list.h:52-61).
   Role of NPairKokkos / NPairIntel (src/KOKKOS/npair_kokkos.h, src/INTEL/npair_intel.h): when a b200 pair style asks for a
   host list (B200MD_HOST_LIST=1; off by default, the device needs none), Neighbor::build reaches this class, which has
   the device build ITS list from the host's atoms and ghosts (b200md_host_atoms, rebuild) and copies it down
   (b200md_neigh_export) into ilist / numneigh / firstneigh -- so lammps_neighlist_element_neighbors (src/library.h:224),
   compute pair/local-like tools and python's get_neighlist see the list the forces are computed from.

   Selection without a core patch: NPair classes are picked by Neighbor::choose_pair (src/neighbor.cpp:2050-2179) as the
   FIRST class whose mask fits the request; there is no suffix lookup.  This file's name sorts before every npair_*.h
   of the reference, so the generated style_npair.h (src/Make.sh:15-40) lists it first and it receives every
   full / bin / atomonly / orthogonal request; for any requestor that is not a b200 pair style in host-ghost mode
   it is exactly its base class NPairFullBinAtomonly (src/npair_full_bin_atomonly.cpp).  INTEGRATION.md also spells
   out the flag-and-mask patch that would select it the way NP_INTEL / NP_KOKKOS_DEVICE classes are selected. */
#ifdef NPAIR_CLASS
// clang-format off
NPairStyle(full/bin/atomonly/b200,
           NPairFullBinAtomonlyB200,
           NP_FULL | NP_BIN | NP_ATOMONLY |
           NP_NEWTON | NP_NEWTOFF | NP_ORTHO);
// clang-format on
#else

#ifndef LMP_NPAIR_B200_H
#define LMP_NPAIR_B200_H

#include "npair_full_bin_atomonly.h"

namespace LAMMPS_NS {

class NPairFullBinAtomonlyB200 : public NPairFullBinAtomonly {
 public:
  NPairFullBinAtomonlyB200(class LAMMPS *);
  void build(class NeighList *) override;

 private:
  int announced;
};

}    // namespace LAMMPS_NS

#endif
#endif
