/* b200_device.h -- the one device context a LAMMPS instance (one MPI rank, one GPU) owns, shared by the
   `b200` styles: PairLJCutB200, PairEAMB200, FixNVEB200, BondFENEB200, FixLangevinB200, VerletB200.  Plays the role FixGPU plays for the GPU
   package (src/GPU/fix_gpu.cpp:43-51, lmp_init_device / lmp_clear_device) and the KokkosLMP object plays for
   KOKKOS: create the device library context once, translate its status codes into error->one like
   GPU_EXTRA::check_flag (src/GPU/gpu_extra.h) does. */
#ifndef LMP_B200_DEVICE_H
#define LMP_B200_DEVICE_H

#include "pointers.h"

#include "../../include/b200md.h"

namespace LAMMPS_NS {

class B200Device : protected Pointers {
 public:
  // the context of `lmp` (created on first use; device = local rank modulo visible GPUs, as `package gpu` picks)
  static B200Device *instance(LAMMPS *lmp);
  static void release(LAMMPS *lmp);    // drop one reference; the last one destroys the context
  static B200Device *peek(LAMMPS *lmp);    // the context if one exists, without creating or referencing it

  b200md_ctx *ctx;
  int device_id;      // CUDA device of this rank
  int resident;       // 1 while VerletB200::run owns the atoms on the device (pair/fix host hooks are idle)
  bigint resident_step;
  int want_host_list;     // B200MD_HOST_LIST=1: the pair styles request a host NeighList, served by npair full/bin/atomonly/b200
  int host_list_fresh;    // set by that NPair class when it rebuilt the device list on this step; compute() then does not

  // status -> error->one(FLERR, text), like GPU_EXTRA::check_flag
  void check(int status, const char *file, int line);
  // units, neighbor settings and per-type cutoffs from the LAMMPS objects (Update, Neighbor, Force)
  void sync_settings();
  // eflag_atom / vflag_atom under the stock loop: run the per-atom pass and ADD the owned atoms' tallies to the pair
  // style's eatom / vatom (host order == device order in host-ghost mode)
  void peratom_to_host(int nlocal, double *eatom, double **vatom);
  // one log line per style per init ("b200: pair lj/cut/b200 on CUDA device 0"): the evidence in a run's output that
  // the suffix resolved to the device style and not to the reference's own class (LAMMPS falls back silently)
  void announce(const char *what);

 private:
  B200Device(LAMMPS *);
  ~B200Device() override;
  int refs;
};

// implemented by the pair styles so VerletB200 can (re)load their coefficient tables into the context
class B200PairTables {
 public:
  virtual ~B200PairTables() noexcept(false) {}
  virtual void device_tables(B200Device *) = 0;
  virtual void device_tables_once(B200Device *) = 0;    // load them unless they are loaded
};

}    // namespace LAMMPS_NS

#define B200_CHECK(dev, call) (dev)->check((call), FLERR)

#endif
