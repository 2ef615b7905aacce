// TEST INFRASTRUCTURE (oracle side) -- never linked into the product.
//
// Small driver over the reference's own C library interface (src/library.h:125-224).
// It runs an input script through the UNMODIFIED reference (oracle/_ref/liblammps_ref.a)
// and dumps, at full precision, what the parity tests compare against:
//   <out>.meta        text: natoms nlocal nghost boxlo/hi pe ke press npairs
//   <out>.atoms.bin   int32 tag[n], int32 type[n], f64 x[3n], f64 v[3n], f64 f[3n]   (owned atoms, local order)
//   <out>.neigh.bin   int32 triples (tag_i, tag_j, code) for every stored pair of the pair style's
//                     neighbor list; code = (sx+1) | (sy+1)<<2 | (sz+1)<<4 | special<<6, where s* is the
//                     periodic image shift of j relative to the owned atom with the same tag.
//
// usage: ref_extract <out-prefix> <pair-style> <lammps command line args ...>
//   e.g. ref_extract /tmp/lj lj/cut -in in.lj -var x 1 -log none -screen none
#include "library.h"
#ifdef B200_STYLES   // built a second time as oracle/_ref/b200_extract (tests/lammps_host/Makefile): the same reader with the
#include "b200_register.h"   // b200 style classes registered, i.e. the list a library user sees under -sf b200
#endif

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

int main(int argc, char **argv)
{
  if (argc < 4) {
    fprintf(stderr, "usage: %s out-prefix pair-style <lammps args>\n", argv[0]);
    return 1;
  }
  std::string out = argv[1];
  const char *pstyle = argv[2];
  std::vector<char *> largs;
  largs.push_back(argv[0]);
  for (int i = 3; i < argc; i++) largs.push_back(argv[i]);
  void *lmp = lammps_open_no_mpi((int) largs.size(), largs.data(), nullptr);
  if (!lmp) return 2;
#ifdef B200_STYLES
  LAMMPS_NS::b200_register_styles((LAMMPS_NS::LAMMPS *) lmp);
#endif
  // the library constructor only records "-in file"; main() is what normally runs it (src/main.cpp:77)
  for (int i = 3; i + 1 < argc; i++)
    if (std::string(argv[i]) == "-in" || std::string(argv[i]) == "-i") lammps_file(lmp, argv[i + 1]);

  int nlocal = lammps_extract_setting(lmp, "nlocal");
  int nghost = lammps_extract_setting(lmp, "nghost");
  double natoms = lammps_get_natoms(lmp);
  double lo[3], hi[3], xy, yz, xz;
  int pflags[3], boxflag;
  lammps_extract_box(lmp, lo, hi, &xy, &yz, &xz, pflags, &boxflag);
  int *tag = (int *) lammps_extract_atom(lmp, "id");
  int *type = (int *) lammps_extract_atom(lmp, "type");
  double **x = (double **) lammps_extract_atom(lmp, "x");
  double **v = (double **) lammps_extract_atom(lmp, "v");
  double **f = (double **) lammps_extract_atom(lmp, "f");

  FILE *fa = fopen((out + ".atoms.bin").c_str(), "wb");
  fwrite(tag, sizeof(int), nlocal, fa);
  fwrite(type, sizeof(int), nlocal, fa);
  fwrite(&x[0][0], sizeof(double), 3 * (size_t) nlocal, fa);
  fwrite(&v[0][0], sizeof(double), 3 * (size_t) nlocal, fa);
  fwrite(&f[0][0], sizeof(double), 3 * (size_t) nlocal, fa);
  fclose(fa);

  // tag -> owned local index
  int maxtag = 0;
  for (int i = 0; i < nlocal; i++)
    if (tag[i] > maxtag) maxtag = tag[i];
  std::vector<int> owner(maxtag + 1, -1);
  for (int i = 0; i < nlocal; i++) owner[tag[i]] = i;
  double prd[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]};

  long npairs = 0;
  int idx = lammps_find_pair_neighlist(lmp, pstyle, 1, 0, 0);
  if (idx >= 0) {
    FILE *fn = fopen((out + ".neigh.bin").c_str(), "wb");
    int nel = lammps_neighlist_num_elements(lmp, idx);
    std::vector<int32_t> buf;
    for (int e = 0; e < nel; e++) {
      int i, nn, *nb;
      lammps_neighlist_element_neighbors(lmp, idx, e, &i, &nn, &nb);
      buf.clear();
      for (int k = 0; k < nn; k++) {
        int j = nb[k] & 0x1FFFFFFF;                 // NEIGHMASK, src/lmptype.h:61-65
        int special = (nb[k] >> 30) & 3;            // SBBITS
        int o = owner[tag[j]];
        int code = special << 6;
        for (int d = 0; d < 3; d++) {
          int s = (int) std::lround((x[j][d] - x[o][d]) / prd[d]);
          code |= (s + 1) << (2 * d);
        }
        buf.push_back(tag[i]);
        buf.push_back(tag[j]);
        buf.push_back(code);
      }
      fwrite(buf.data(), sizeof(int32_t), buf.size(), fn);
      npairs += nn;
    }
    fclose(fn);
  }

  FILE *fm = fopen((out + ".meta").c_str(), "w");
  fprintf(fm, "natoms %.0f\nnlocal %d\nnghost %d\n", natoms, nlocal, nghost);
  fprintf(fm, "boxlo %.17g %.17g %.17g\nboxhi %.17g %.17g %.17g\n", lo[0], lo[1], lo[2], hi[0], hi[1], hi[2]);
  fprintf(fm, "pe %.17g\nke %.17g\npress %.17g\ntemp %.17g\n", lammps_get_thermo(lmp, "pe"),
          lammps_get_thermo(lmp, "ke"), lammps_get_thermo(lmp, "press"), lammps_get_thermo(lmp, "temp"));
  fprintf(fm, "pxx %.17g\npyy %.17g\npzz %.17g\npxy %.17g\npxz %.17g\npyz %.17g\n",
          lammps_get_thermo(lmp, "pxx"), lammps_get_thermo(lmp, "pyy"), lammps_get_thermo(lmp, "pzz"),
          lammps_get_thermo(lmp, "pxy"), lammps_get_thermo(lmp, "pxz"), lammps_get_thermo(lmp, "pyz"));
  fprintf(fm, "evdwl %.17g\nebond %.17g\n", lammps_get_thermo(lmp, "evdwl"), lammps_get_thermo(lmp, "ebond"));
  fprintf(fm, "npairs %ld\n", npairs);
  fclose(fm);
  lammps_close(lmp);
  return 0;
}
