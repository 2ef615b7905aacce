/* b200_register.cpp -- see b200_register.h */
#include "b200_register.h"

#include "bond_fene_b200.h"
#include "fix_langevin_b200.h"
#include "fix_nve_b200.h"
#include "force.h"
#include "input.h"
#include "lammps.h"
#include "modify.h"
#include "pair_eam_alloy_b200.h"
#include "pair_eam_b200.h"
#include "pair_lj_cut_b200.h"
#include "update.h"
#include "verlet_b200.h"

#include <cstring>

namespace LAMMPS_NS {

static Pair *make_lj(LAMMPS *lmp) { return new PairLJCutB200(lmp); }
static Pair *make_eam(LAMMPS *lmp) { return new PairEAMB200(lmp); }
static Pair *make_eam_alloy(LAMMPS *lmp) { return new PairEAMAlloyB200(lmp); }
static Pair *make_eam_fs(LAMMPS *lmp) { return new PairEAMFSB200(lmp); }
static Fix *make_nve(LAMMPS *lmp, int narg, char **arg) { return new FixNVEB200(lmp, narg, arg); }
static Fix *make_langevin(LAMMPS *lmp, int narg, char **arg) { return new FixLangevinB200(lmp, narg, arg); }
static Bond *make_fene(LAMMPS *lmp) { return new BondFENEB200(lmp); }
static Integrate *make_verlet(LAMMPS *lmp, int narg, char **arg) { return new VerletB200(lmp, narg, arg); }

void b200_register_styles(LAMMPS *lmp)
{
  (*lmp->force->pair_map)["lj/cut/b200"] = &make_lj;
  (*lmp->force->pair_map)["eam/b200"] = &make_eam;
  (*lmp->force->pair_map)["eam/alloy/b200"] = &make_eam_alloy;
  (*lmp->force->pair_map)["eam/fs/b200"] = &make_eam_fs;
  (*lmp->modify->fix_map)["nve/b200"] = &make_nve;
  (*lmp->modify->fix_map)["langevin/b200"] = &make_langevin;
  (*lmp->force->bond_map)["fene/b200"] = &make_fene;
  (*lmp->update->integrate_map)["verlet/b200"] = &make_verlet;
  // Update's constructor created the default `verlet` before these entries existed; with `-sf b200` re-create
  // it so the suffix logic of Update::new_integrate (src/update.cpp:360-400) now finds verlet/b200
  if (lmp->suffix_enable && lmp->suffix && strcmp(lmp->suffix, "b200") == 0) {
    char style[] = "verlet";
    char *args[1] = {style};
    lmp->update->create_integrate(1, args, 1);
  }
}

}    // namespace LAMMPS_NS
