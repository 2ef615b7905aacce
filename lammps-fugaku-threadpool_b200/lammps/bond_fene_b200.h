/* bond_style fene/b200 -- BondFENE (src/MOLECULE/bond_fene.h).  Under the stock Verlet loop the atoms live on the
   host and this is BondFENE itself (bonds stay on the host there, as with the GPU package); under
   `run_style verlet/b200` VerletB200 hands the topology and these coefficients to the device (b200md_set_bonds,
   b200md_bond_fene) and the bond forces come from k_bond_fene inside the resident loop. */
#ifdef BOND_CLASS
// clang-format off
BondStyle(fene/b200,BondFENEB200);
// clang-format on
#else

#ifndef LMP_BOND_FENE_B200_H
#define LMP_BOND_FENE_B200_H

#include "bond_fene.h"

namespace LAMMPS_NS {

class BondFENEB200 : public BondFENE {
 public:
  BondFENEB200(class LAMMPS *lmp) : BondFENE(lmp) {}
  void init_style() override;
  // bond_coeff arrays, indexed 1..nbondtypes (what b200md_bond_fene takes)
  const double *dev_k() const { return k; }
  const double *dev_r0() const { return r0; }
  const double *dev_epsilon() const { return epsilon; }
  const double *dev_sigma() const { return sigma; }
  bool all_set() const;
};

}    // namespace LAMMPS_NS
#endif
#endif
