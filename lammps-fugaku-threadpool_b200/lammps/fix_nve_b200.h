/* fix nve/b200 -- FixNVE (src/fix_nve.h).  Under the stock Verlet loop the atoms live on the host and this is
   FixNVE itself; under `run_style verlet/b200` VerletB200 recognises it and runs initial/final_integrate as the
   device kernels k_nve_initial / k_nve_final (b200md_nve_initial/final), so these host hooks are never reached. */
#ifdef FIX_CLASS
// clang-format off
FixStyle(nve/b200,FixNVEB200);
// clang-format on
#else

#ifndef LMP_FIX_NVE_B200_H
#define LMP_FIX_NVE_B200_H

#include "fix_nve.h"

namespace LAMMPS_NS {

class FixNVEB200 : public FixNVE {
 public:
  FixNVEB200(class LAMMPS *, int, char **);
  void init() override;
  void initial_integrate(int) override;
  void final_integrate() override;
  double device_dtv() const { return dtv; }
  double device_dtf() const { return dtf; }
};

}    // namespace LAMMPS_NS
#endif
#endif
