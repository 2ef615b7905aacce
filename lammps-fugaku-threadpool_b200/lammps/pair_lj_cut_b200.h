/* pair_style lj/cut/b200 -- PairLJCut (src/pair_lj_cut.h) with compute() on the B200 device library.
   Same role and shape as PairLJCutGPU (src/GPU/pair_lj_cut_gpu.h:14-45): settings/coeff/init_one/restart are
   inherited unchanged, so input scripts, mixing rules, pair_modify shift and restart files behave as for lj/cut. */
#ifdef PAIR_CLASS
// clang-format off
PairStyle(lj/cut/b200,PairLJCutB200);
// clang-format on
#else

#ifndef LMP_PAIR_LJ_CUT_B200_H
#define LMP_PAIR_LJ_CUT_B200_H

#include "b200_device.h"
#include "pair_lj_cut.h"

namespace LAMMPS_NS {

class PairLJCutB200 : public PairLJCut, public B200PairTables {
 public:
  PairLJCutB200(LAMMPS *lmp);
  ~PairLJCutB200() override;
  void compute(int, int) override;
  void init_style() override;
  double memory_usage() override;
  void device_tables(B200Device *) override;
  void device_tables_once(B200Device *d) override { if (!tables_loaded) device_tables(d); }

 private:
  B200Device *dev;
  int tables_loaded;
};

}    // namespace LAMMPS_NS
#endif
#endif
