/* pair_style eam/alloy/b200 and eam/fs/b200 -- the reference's setfl / Finnis-Sinclair readers (PairEAMAlloy,
   PairEAMFS: coeff, read_file, file2array; src/MANYBODY/pair_eam_alloy.h, pair_eam_fs.h) combined with the device
   compute of PairEAMB200, the way PairEAMAlloyGPU / PairEAMAlloyOMP combine them with their compute
   (src/OPENMP/pair_eam_alloy_omp.h:29-38).  Nothing is re-implemented: the spline tables the reference builds from
   the file go to the device as they are. */
#ifdef PAIR_CLASS
// clang-format off
PairStyle(eam/alloy/b200,PairEAMAlloyB200);
PairStyle(eam/fs/b200,PairEAMFSB200);
// clang-format on
#else

#ifndef LMP_PAIR_EAM_ALLOY_B200_H
#define LMP_PAIR_EAM_ALLOY_B200_H

#include "pair_eam_alloy.h"
#include "pair_eam_b200.h"
#include "pair_eam_fs.h"

namespace LAMMPS_NS {

class PairEAMAlloyB200 : public PairEAMAlloy, public PairEAMB200 {
 public:
  PairEAMAlloyB200(LAMMPS *lmp) : PairEAM(lmp), PairEAMAlloy(lmp), PairEAMB200(lmp) {}
};

class PairEAMFSB200 : public PairEAMFS, public PairEAMB200 {
 public:
  PairEAMFSB200(LAMMPS *lmp) : PairEAM(lmp), PairEAMFS(lmp), PairEAMB200(lmp) {}
};

}    // namespace LAMMPS_NS
#endif
#endif
