/* fix langevin/b200 -- FixLangevin (src/fix_langevin.h).  Under the stock Verlet loop this is FixLangevin itself;
   under `run_style verlet/b200` VerletB200 reads the thermostat's parameters from here and the device applies the
   same force (k_langevin) with the same RanMars stream, consumed in this rank's local atom order
   (b200md_fix_langevin, b200md_langevin_order).  Only the plain form runs on the device: constant target or a
   linear ramp, no tally / zero / gjf / omega / angmom / scale keywords, per-type masses. */
#ifdef FIX_CLASS
// clang-format off
FixStyle(langevin/b200,FixLangevinB200);
// clang-format on
#else

#ifndef LMP_FIX_LANGEVIN_B200_H
#define LMP_FIX_LANGEVIN_B200_H

#include "fix_langevin.h"

namespace LAMMPS_NS {

class FixLangevinB200 : public FixLangevin {
 public:
  FixLangevinB200(class LAMMPS *, int, char **);
  ~FixLangevinB200() override;
  void setup(int) override;
  void post_force(int) override;
  // null when the device form applies, else the reason it does not
  const char *device_unsupported() const;
  double dev_t_start() const { return t_start; }
  double dev_t_stop() const { return t_stop; }
  double dev_t_period() const { return t_period; }
  int dev_seed() const { return seed; }
  bool order_sent = false;    // b200md_langevin_order is given once: the device then follows Atom::sort itself
};

}    // namespace LAMMPS_NS
#endif
#endif
