/* fix_nve_b200.cpp -- see fix_nve_b200.h */
#include "fix_nve_b200.h"

#include "atom.h"
#include "b200_device.h"
#include "error.h"
#include "group.h"

using namespace LAMMPS_NS;

FixNVEB200::FixNVEB200(LAMMPS *lmp, int narg, char **arg) : FixNVE(lmp, narg, arg) {}

void FixNVEB200::init()
{
  FixNVE::init();
  if (atom->rmass) error->all(FLERR, "Fix nve/b200 requires per-type masses");
}

void FixNVEB200::initial_integrate(int vflag)
{
  FixNVE::initial_integrate(vflag);
}

void FixNVEB200::final_integrate()
{
  FixNVE::final_integrate();
}
