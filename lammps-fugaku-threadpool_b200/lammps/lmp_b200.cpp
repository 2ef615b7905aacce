/* lmp_b200.cpp -- the LAMMPS executable (src/main.cpp) with the b200 styles registered: the unmodified reference
   library + the style classes of this directory + libb200md.so.  `lmp_b200 -sf b200 -in in.lj` runs an unchanged
   input script with pair lj/cut/b200, fix nve/b200 and (with `run_style verlet/b200` or -var/-echo as usual) the
   device-resident loop. */
#include "b200_register.h"
#include "input.h"
#include "lammps.h"
#include "exceptions.h"

#include <cstdio>
#include <mpi.h>

using namespace LAMMPS_NS;

int main(int argc, char **argv)
{
  MPI_Init(&argc, &argv);
  int rc = 0;
  try {
    auto *lammps = new LAMMPS(argc, argv, MPI_COMM_WORLD);
    b200_register_styles(lammps);
    lammps->input->file();
    delete lammps;
  } catch (LAMMPSAbortException &ae) {
    fprintf(stderr, "%s\n", ae.what());
    rc = 1;
  } catch (LAMMPSException &) {
    rc = 1;
  }
  MPI_Finalize();
  return rc;
}
