/* fix_langevin_b200.cpp -- see fix_langevin_b200.h */
#include "fix_langevin_b200.h"

#include "atom.h"
#include "b200_device.h"

using namespace LAMMPS_NS;

FixLangevinB200::FixLangevinB200(LAMMPS *lmp, int narg, char **arg) : FixLangevin(lmp, narg, arg) {}

FixLangevinB200::~FixLangevinB200()
{
  // unfix: the device stops applying the thermostat and forgets the stream (a new fix starts a new one)
  B200Device *dev = B200Device::peek(lmp);
  if (dev && dev->ctx) b200md_fix_langevin(dev->ctx, 0.0, 0.0, 0.0, 0, 0.0);
}

// while VerletB200 owns the atoms the thermostat force is applied on the device (k_langevin); the host hooks are
// only reached under the stock Verlet loop, where this fix is FixLangevin itself
void FixLangevinB200::setup(int vflag)
{
  B200Device *dev = B200Device::peek(lmp);
  if (dev && dev->resident) return;
  FixLangevin::setup(vflag);
}

void FixLangevinB200::post_force(int vflag)
{
  B200Device *dev = B200Device::peek(lmp);
  if (dev && dev->resident) return;
  FixLangevin::post_force(vflag);
}

const char *FixLangevinB200::device_unsupported() const
{
  if (tstyle != 0) return "a variable target temperature";    // enum {CONSTANT, EQUAL, ATOM} in fix_langevin.cpp
  if (gjfflag) return "the gjf keyword";
  if (tallyflag) return "the tally keyword";
  if (zeroflag) return "the zero keyword";
  if (oflag) return "the omega keyword";
  if (ascale != 0.0) return "the angmom keyword";
  if (tbiasflag) return "a temperature compute with a bias";
  if (igroup != 0) return "a group other than all";
  if (atom->rmass) return "per-atom masses";
  for (int i = 1; i <= atom->ntypes; i++)
    if (ratio[i] != 1.0) return "the scale keyword";
  return nullptr;
}
