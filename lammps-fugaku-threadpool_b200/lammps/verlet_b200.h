/* run_style verlet/b200 -- the device-resident time loop.  Precedent: `run_style verlet/kk`
   (src/KOKKOS/verlet_kokkos.h:16-18), which also replaces Verlet::run so that atoms stay on the device between
   the style calls.  For atomic systems setup() is the stock Verlet::setup (the b200 pair style serves it in
   host-ghost mode, so the step-0 thermo line goes through the reference's own machinery); run(n) uploads the atoms once, advances them
   with b200md_run between output steps (initial_integrate / decide / exchange+borders+build | forward_comm /
   pair / final_integrate all on the device, src/verlet.cpp:229-360) and brings x, v, f, image and the pair
   energy/virial back whenever Output needs them. */
#ifdef INTEGRATE_CLASS
// clang-format off
IntegrateStyle(verlet/b200,VerletB200);
// clang-format on
#else

#ifndef LMP_VERLET_B200_H
#define LMP_VERLET_B200_H

#include "verlet.h"

namespace LAMMPS_NS {

class VerletB200 : public Verlet {
 public:
  VerletB200(class LAMMPS *, int, char **);
  ~VerletB200() override;
  void init() override;
  void setup(int) override;
  void run(int) override;

 private:
  class B200Device *dev;
  class FixNVEB200 *nve;
  // bench/in.chain: bond_style fene/b200 and fix langevin/b200 also run inside the resident loop (one rank).  Such a
  // system is set up on the device as well (setup() below), because the host-ghost pair path serves atomic
  // systems only (no special-bond exclusion).
  class BondFENEB200 *bond;
  class FixLangevinB200 *langevin;
  bigint device_setup_step;    // update->ntimestep the device was last set up at by setup(), -1 = not
  void upload();
  void download();
  void fetch_energies();
};

}    // namespace LAMMPS_NS
#endif
#endif
