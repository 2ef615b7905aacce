/* pair_style eam/b200 -- PairEAM (src/MANYBODY/pair_eam.h) with compute() on the B200 device library.
   Same shape as PairEAMGPU (src/GPU/pair_eam_gpu.h): coeff/read_file/file2array/array2spline are inherited,
   the spline tables they produce are handed to the device as they are (b200md_eam_init). */
#ifdef PAIR_CLASS
// clang-format off
PairStyle(eam/b200,PairEAMB200);
// clang-format on
#else

#ifndef LMP_PAIR_EAM_B200_H
#define LMP_PAIR_EAM_B200_H

#include "b200_device.h"
#include "pair_eam.h"

namespace LAMMPS_NS {

// `virtual`: PairEAMAlloy / PairEAMFS derive virtually from PairEAM, and eam/alloy/b200, eam/fs/b200 combine them with this class
class PairEAMB200 : virtual public PairEAM, public B200PairTables {
 public:
  PairEAMB200(LAMMPS *lmp);
  ~PairEAMB200() override;
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
