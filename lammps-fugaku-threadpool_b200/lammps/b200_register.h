/* b200_register.h -- run-time registration of the b200 styles into a LAMMPS instance, for hosts that link the
   style classes without regenerating style_*.h: inserts the creators into Force::pair_map, Modify::fix_map and
   Update::integrate_map, which is what the PLUGIN package does for pair/fix styles (src/PLUGIN/plugin.cpp:141-280).
   With the entries present, `-sf b200` resolves lj/cut -> lj/cut/b200 etc. through the stock suffix logic
   (Force::new_pair src/force.cpp:247-278, Modify::add_fix, Update::new_integrate). */
#ifndef LMP_B200_REGISTER_H
#define LMP_B200_REGISTER_H
namespace LAMMPS_NS {
class LAMMPS;
void b200_register_styles(LAMMPS *lmp);
}
#endif
