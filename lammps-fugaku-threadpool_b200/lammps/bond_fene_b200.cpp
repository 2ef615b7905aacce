/* bond_fene_b200.cpp -- see bond_fene_b200.h */
#include "bond_fene_b200.h"

#include "atom.h"
#include "error.h"

using namespace LAMMPS_NS;

void BondFENEB200::init_style()
{
  BondFENE::init_style();    // the special_bonds warning of the reference
}

bool BondFENEB200::all_set() const
{
  if (!allocated) return false;
  for (int i = 1; i <= atom->nbondtypes; i++)
    if (!setflag[i]) return false;
  return true;
}
