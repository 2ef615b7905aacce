/* b200plugin.cpp -- the PLUGIN-package route (src/lammpsplugin.h:25-44, src/PLUGIN/plugin.cpp:74-125;
   template: examples/plugins/morse2plugin.cpp, nve2plugin.cpp): `plugin load b200plugin.so` registers
   pair lj/cut/b200, pair eam/b200 and fix nve/b200 into an unmodified, shared-library LAMMPS.  The PLUGIN package
   has no integrate-style slot, so under this route the stock Verlet loop drives the pair styles in host-ghost
   mode; the device-resident loop (run_style verlet/b200) needs the compile-in route or b200_register_styles(). */
#include "lammpsplugin.h"

#include "fix_nve_b200.h"
#include "lammps.h"
#include "pair_eam_alloy_b200.h"
#include "pair_eam_b200.h"
#include "pair_lj_cut_b200.h"
#include "version.h"

using namespace LAMMPS_NS;

static Pair *lj_creator(LAMMPS *lmp) { return new PairLJCutB200(lmp); }
static Pair *eam_creator(LAMMPS *lmp) { return new PairEAMB200(lmp); }
static Pair *eam_alloy_creator(LAMMPS *lmp) { return new PairEAMAlloyB200(lmp); }
static Pair *eam_fs_creator(LAMMPS *lmp) { return new PairEAMFSB200(lmp); }
static Fix *nve_creator(LAMMPS *lmp, int argc, char **argv) { return new FixNVEB200(lmp, argc, argv); }

extern "C" void lammpsplugin_init(void *lmp, void *handle, void *regfunc)
{
  lammpsplugin_t plugin;
  lammpsplugin_regfunc register_plugin = (lammpsplugin_regfunc) regfunc;
  plugin.version = LAMMPS_VERSION;
  plugin.author = "b200md";
  plugin.handle = handle;

  plugin.style = "pair";
  plugin.name = "lj/cut/b200";
  plugin.info = "lj/cut on the B200 device library (libb200md.so)";
  plugin.creator.v1 = (lammpsplugin_factory1 *) &lj_creator;
  (*register_plugin)(&plugin, lmp);

  plugin.name = "eam/b200";
  plugin.info = "eam (funcfl) on the B200 device library (libb200md.so)";
  plugin.creator.v1 = (lammpsplugin_factory1 *) &eam_creator;
  (*register_plugin)(&plugin, lmp);

  plugin.name = "eam/alloy/b200";
  plugin.info = "eam/alloy (setfl) on the B200 device library (libb200md.so)";
  plugin.creator.v1 = (lammpsplugin_factory1 *) &eam_alloy_creator;
  (*register_plugin)(&plugin, lmp);

  plugin.name = "eam/fs/b200";
  plugin.info = "eam/fs (Finnis-Sinclair) on the B200 device library (libb200md.so)";
  plugin.creator.v1 = (lammpsplugin_factory1 *) &eam_fs_creator;
  (*register_plugin)(&plugin, lmp);

  plugin.style = "fix";
  plugin.name = "nve/b200";
  plugin.info = "nve companion of the b200 pair styles";
  plugin.creator.v2 = (lammpsplugin_factory2 *) &nve_creator;
  (*register_plugin)(&plugin, lmp);
}
