// eam_host.cu -- host side of pair_style eam for funcfl files: what PairEAM::coeff / read_file /
// file2array / array2spline / init_one do before the first compute (src/MANYBODY/pair_eam.cpp:372-412,
// 460-526, 533-765, 435-456).  Produces the spline tables and hands them to b200md_eam_init.  In a LAMMPS
// build PairEAMB200 inherits these arrays from PairEAM instead; this entry point serves stand-alone hosts.
#include <math.h>
#include <stdio.h>

#include <algorithm>
#include <string>
#include <vector>

#include "ctx.cuh"

namespace {

struct FuncflFile {          // Funcfl (src/MANYBODY/pair_eam.h:78-85); tables are 0-based here
  double mass = 0, drho = 0, dr = 0, cut = 0;
  std::vector<double> frho, zr, rhor;
};

bool read_numbers(FILE *fp, std::vector<double> &dst, int n)
{
  dst.resize(n);
  for (int k = 0; k < n; k++)
    if (fscanf(fp, "%lg", &dst[k]) != 1) return false;
  return true;
}

// line 1 comment, line 2 "Z mass ...", line 3 "nrho drho nr dr cut", then frho[nrho], zr[nr], rhor[nr]
// `system_real`: the run uses `units real`.  A file whose first line carries "UNITS: metal" is then converted
// like PairEAM::read_file does through PotentialFileReader (pair_eam.cpp:466-505, utils.cpp:1488-1506:
// energy x 23.060549 for frho, x sqrt(23.060549) for zr); a "UNITS: real" file under metal gets the inverse.
int read_funcfl(const char *name, FuncflFile &f, std::string &why, bool system_real)
{
  FILE *fp = fopen(name, "r");
  if (!fp) { why = std::string("Cannot open EAM potential file ") + name; return 1; }
  char line[2048], first[2048] = {0};
  int iz = 0, nrho = 0, nr = 0;
  bool ok = fgets(first, sizeof first, fp) && fgets(line, sizeof line, fp) && sscanf(line, "%d %lg", &iz, &f.mass) == 2 &&
            fgets(line, sizeof line, fp) && sscanf(line, "%d %lg %d %lg %lg", &nrho, &f.drho, &nr, &f.dr, &f.cut) == 5 &&
            nrho > 4 && nr > 4 && f.drho > 0 && f.dr > 0 && read_numbers(fp, f.frho, nrho) && read_numbers(fp, f.zr, nr) &&
            read_numbers(fp, f.rhor, nr);
  fclose(fp);
  if (!ok) { why = std::string("Invalid EAM potential file ") + name; return 1; }
  double conv = 1.0;
  const std::string head(first);
  if (head.find("UNITS: metal") != std::string::npos && system_real) conv = 23.060549;
  else if (head.find("UNITS: real") != std::string::npos && !system_real) conv = 1.0 / 23.060549;
  if (conv != 1.0) {
    const double sq = sqrt(conv);
    for (auto &v : f.frho) v *= conv;
    for (auto &v : f.zr) v *= sq;
  }
  return 0;
}

// value of a tabulated function (spacing d, first knot at 0) at r by the reference's 4-point Lagrange
// formula (pair_eam.cpp:580-596); knots are 1-based there, hence the index shifts
double lagrange4(const std::vector<double> &tab, double d, double r)
{
  const int n = (int) tab.size();
  double p = r / d + 1.0;
  int k = (int) p;
  k = std::max(2, std::min(k, n - 2));
  p -= k;
  p = std::min(p, 2.0);
  const double sixth = 1.0 / 6.0;
  const double w1 = -sixth * p * (p - 1.0) * (p - 2.0);
  const double w2 = 0.5 * (p * p - 1.0) * (p - 2.0);
  const double w3 = -0.5 * p * (p + 1.0) * (p - 2.0);
  const double w4 = sixth * p * (p * p - 1.0);
  return w1 * tab[k - 2] + w2 * tab[k - 1] + w3 * tab[k] + w4 * tab[k + 1];
}

// PairEAM::interpolate (pair_eam.cpp:769-797): rows 1..n of a [n+1][7] table, row 0 unused
void cubic_rows(int n, double delta, const double *f1, double *row0)
{
  auto S = [&](int m, int k) -> double & { return row0[(size_t) m * 7 + k]; };
  for (int m = 1; m <= n; m++) S(m, 6) = f1[m];
  S(1, 5) = S(2, 6) - S(1, 6);
  S(2, 5) = 0.5 * (S(3, 6) - S(1, 6));
  S(n - 1, 5) = 0.5 * (S(n, 6) - S(n - 2, 6));
  S(n, 5) = S(n, 6) - S(n - 1, 6);
  for (int m = 3; m <= n - 2; m++) S(m, 5) = ((S(m - 2, 6) - S(m + 2, 6)) + 8.0 * (S(m + 1, 6) - S(m - 1, 6))) / 12.0;
  for (int m = 1; m < n; m++) {
    const double d = S(m + 1, 6) - S(m, 6);
    S(m, 4) = 3.0 * d - 2.0 * S(m, 5) - S(m + 1, 5);
    S(m, 3) = S(m, 5) + S(m + 1, 5) - 2.0 * d;
  }
  S(n, 4) = 0.0;
  S(n, 3) = 0.0;
  for (int m = 1; m <= n; m++) {
    S(m, 2) = S(m, 5) / delta;
    S(m, 1) = 2.0 * S(m, 4) / delta;
    S(m, 0) = 3.0 * S(m, 3) / delta;
  }
}

}  // namespace

extern "C" int b200md_eam_init_funcfl(b200md_ctx *c, int ntypes, int nfiles, const char *const *files, const int *map,
                                      double *mass_out, double *cutmax_out)
{
  if (!c) return b200md_fail(nullptr, B200MD_ERR_ARG, "NULL context");
  if (ntypes < 1 || ntypes > B200MD_MAXTYPES || nfiles < 1) return b200md_fail(c, B200MD_ERR_ARG, "eam: bad ntypes/nfiles");
  std::vector<FuncflFile> F(nfiles);
  std::string why;
  // `units real` is recognised by its Boltzmann constant (b200md_set_units carries the constants, src/update.cpp:166-185)
  const bool system_real = fabs(c->boltz - 0.0019872067) < 1e-12;
  for (int k = 0; k < nfiles; k++)
    if (read_funcfl(files[k], F[k], why, system_real)) return b200md_fail(c, B200MD_ERR_ARG, "%s", why.c_str());
  for (int i = 1; i <= ntypes; i++)
    if (map[i] < -1 || map[i] >= nfiles) return b200md_fail(c, B200MD_ERR_ARG, "eam: bad map[%d]", i);

  // common grid: largest spacing, largest extent among the files in use (pair_eam.cpp:540-565)
  double dr = 0, drho = 0, rmax = 0, rhomax = 0, cutmax = 0;
  for (int k = 0; k < nfiles; k++) {
    bool used = false;
    for (int i = 1; i <= ntypes; i++) used |= (map[i] == k);
    cutmax = std::max(cutmax, F[k].cut);  // init_one: max over ALL files read
    if (!used) continue;
    dr = std::max(dr, F[k].dr);
    drho = std::max(drho, F[k].drho);
    rmax = std::max(rmax, ((int) F[k].rhor.size() - 1) * F[k].dr);
    rhomax = std::max(rhomax, ((int) F[k].frho.size() - 1) * F[k].drho);
  }
  if (dr <= 0 || drho <= 0) return b200md_fail(c, B200MD_ERR_ARG, "eam: no atom type is mapped to a funcfl file");
  const int nr = (int) (rmax / dr + 0.5), nrho = (int) (rhomax / drho + 0.5);
  const int nfrho = nfiles + 1, nrhor = nfiles, nz2r = nfiles * (nfiles + 1) / 2, nt1 = ntypes + 1;

  std::vector<double> g((size_t) std::max(nr, nrho) + 1);
  std::vector<double> frho_s((size_t) nfrho * (nrho + 1) * 7, 0.0), rhor_s((size_t) nrhor * (nr + 1) * 7, 0.0),
      z2r_s((size_t) nz2r * (nr + 1) * 7, 0.0);
  for (int k = 0; k < nfrho; k++) {  // last table stays zero: types without an EAM file (pair_eam.cpp:600-611)
    for (int m = 1; m <= nrho; m++) g[m] = (k < nfiles) ? lagrange4(F[k].frho, F[k].drho, (m - 1) * drho) : 0.0;
    cubic_rows(nrho, drho, g.data(), &frho_s[(size_t) k * (nrho + 1) * 7]);
  }
  for (int k = 0; k < nrhor; k++) {
    for (int m = 1; m <= nr; m++) g[m] = lagrange4(F[k].rhor, F[k].dr, (m - 1) * dr);
    cubic_rows(nr, dr, g.data(), &rhor_s[(size_t) k * (nr + 1) * 7]);
  }
  int t = 0;
  for (int a = 0; a < nfiles; a++)
    for (int b = 0; b <= a; b++, t++) {  // z2r = 27.2 * 0.529 * Z_a(r) * Z_b(r)   (pair_eam.cpp:690-712)
      for (int m = 1; m <= nr; m++) {
        const double r = (m - 1) * dr;
        g[m] = 27.2 * 0.529 * lagrange4(F[a].zr, F[a].dr, r) * lagrange4(F[b].zr, F[b].dr, r);
      }
      cubic_rows(nr, dr, g.data(), &z2r_s[(size_t) t * (nr + 1) * 7]);
    }

  std::vector<int> t2f(nt1, 0), t2r((size_t) nt1 * nt1, 0), t2z((size_t) nt1 * nt1, 0);
  for (int i = 1; i <= ntypes; i++) {
    t2f[i] = map[i] >= 0 ? map[i] : nfrho - 1;
    for (int j = 1; j <= ntypes; j++) {
      t2r[(size_t) i * nt1 + j] = map[i];
      int hi = std::max(map[i], map[j]), lo = std::min(map[i], map[j]);
      t2z[(size_t) i * nt1 + j] = (lo < 0) ? 0 : hi * (hi + 1) / 2 + lo;
    }
  }
  if (mass_out)
    for (int i = 1; i <= ntypes; i++) mass_out[i] = map[i] >= 0 ? F[map[i]].mass : 0.0;
  if (cutmax_out) *cutmax_out = cutmax;
  return b200md_eam_init(c, ntypes, nrho, nr, 1.0 / drho, 1.0 / dr, rhomax, cutmax * cutmax, nfrho, frho_s.data(),
                         nrhor, rhor_s.data(), nz2r, z2r_s.data(), t2f.data(), t2r.data(), t2z.data());
}

// ---- pair_style eam/alloy (DYNAMO setfl) and eam/fs (Finnis-Sinclair) -----------------------------------------
// pair_coeff * * file El1 El2 ...: PairEAMAlloy::coeff / read_file / file2array (src/MANYBODY/pair_eam_alloy.cpp:55-337)
// and PairEAMFS (src/MANYBODY/pair_eam_fs.cpp:55-345).  The tables are used on the file's own grid; splines and
// cutoff as PairEAM (array2spline, init_one).  The device kernels are the multi-type ones the funcfl path uses:
// type2rhor is a full [type][type] map, which is all eam/fs needs.
namespace {

bool content_line(FILE *fp, char *line, int n)
{
  while (fgets(line, n, fp)) {
    const char *q = line;
    while (*q == ' ' || *q == '\t' || *q == '\r' || *q == '\n') q++;
    if (*q) return true;
  }
  return false;
}

bool read_into(FILE *fp, double *dst, int n)
{
  for (int k = 0; k < n; k++)
    if (fscanf(fp, "%lg", &dst[k]) != 1) return false;
  return true;
}

}  // namespace

extern "C" int b200md_eam_init_setfl(b200md_ctx *c, int ntypes, const char *file, int fs, const char *const *elements,
                                     double *mass_out, double *cutmax_out)
{
  if (!c) return b200md_fail(nullptr, B200MD_ERR_ARG, "NULL context");
  if (ntypes < 1 || ntypes > B200MD_MAXTYPES || !file || !elements) return b200md_fail(c, B200MD_ERR_ARG, "eam: bad ntypes / file / elements");
  FILE *fp = fopen(file, "r");
  if (!fp) return b200md_fail(c, B200MD_ERR_ARG, "Cannot open EAM potential file %s", file);
  struct Closer { FILE *f; ~Closer() { fclose(f); } } closer{fp};
  char first[4096], line[4096];
  if (!fgets(first, sizeof first, fp) || !fgets(line, sizeof line, fp) || !fgets(line, sizeof line, fp) || !content_line(fp, line, sizeof line))
    return b200md_fail(c, B200MD_ERR_ARG, "Invalid EAM potential file %s", file);
  // PotentialFileReader's transparent metal <-> real conversion of energies (src/utils.cpp:1488-1506)
  const bool system_real = fabs(c->boltz - 0.0019872067) < 1e-12;
  double conv = 1.0;
  if (strstr(first, "UNITS: metal") && system_real) conv = 23.060549;
  else if (strstr(first, "UNITS: real") && !system_real) conv = 1.0 / 23.060549;
  int nel = 0, off = 0;
  if (sscanf(line, "%d%n", &nel, &off) != 1 || nel < 1 || nel > B200MD_MAXTYPES) return b200md_fail(c, B200MD_ERR_ARG, "Invalid EAM potential file %s", file);
  std::vector<std::string> names;
  for (const char *q = line + off; (int) names.size() < nel;) {
    char buf[64];
    int used = 0;
    if (sscanf(q, "%63s%n", buf, &used) != 1) return b200md_fail(c, B200MD_ERR_ARG, "Incorrect element names in EAM potential file");
    names.push_back(buf);
    q += used;
  }
  int nrho = 0, nr = 0;
  double drho = 0, dr = 0, cut = 0;
  if (!content_line(fp, line, sizeof line) || sscanf(line, "%d %lg %d %lg %lg", &nrho, &drho, &nr, &dr, &cut) != 5 || nrho <= 4 || nr <= 4 ||
      dr <= 0.0 || drho <= 0.0)
    return b200md_fail(c, B200MD_ERR_ARG, "Invalid EAM potential file %s", file);
  const int nfrho = nel + 1, nrhor = fs ? nel * nel : nel, nz2r = nel * (nel + 1) / 2, nt1 = ntypes + 1;
  std::vector<double> mass(nel), frho((size_t) nfrho * (nrho + 1), 0.0), rhor((size_t) nrhor * (nr + 1), 0.0), z2r((size_t) nz2r * (nr + 1), 0.0);
  for (int e = 0; e < nel; e++) {
    int iz = 0;
    if (!content_line(fp, line, sizeof line) || sscanf(line, "%d %lg", &iz, &mass[e]) != 2 || !read_into(fp, &frho[(size_t) e * (nrho + 1) + 1], nrho))
      return b200md_fail(c, B200MD_ERR_ARG, "Invalid EAM potential file %s", file);
    for (int j = 0; j < (fs ? nel : 1); j++)
      if (!read_into(fp, &rhor[(size_t) (fs ? e * nel + j : e) * (nr + 1) + 1], nr)) return b200md_fail(c, B200MD_ERR_ARG, "Invalid EAM potential file %s", file);
    // the setfl reader converts knots 1..nrho-1 (pair_eam_alloy.cpp:181), the fs reader all of them (pair_eam_fs.cpp:165)
    if (conv != 1.0)
      for (int m = 1; m < nrho + (fs ? 1 : 0); m++) frho[(size_t) e * (nrho + 1) + m] *= conv;
  }
  for (int t = 0; t < nz2r; t++) {
    if (!read_into(fp, &z2r[(size_t) t * (nr + 1) + 1], nr)) return b200md_fail(c, B200MD_ERR_ARG, "Invalid EAM potential file %s", file);
    if (conv != 1.0)
      for (int m = 1; m < nr + (fs ? 1 : 0); m++) z2r[(size_t) t * (nr + 1) + m] *= conv;   // pair_eam_alloy.cpp:189 / pair_eam_fs.cpp:177
  }
  std::vector<int> map(nt1, -1);
  int count = 0;
  for (int i = 1; i <= ntypes; i++) {
    if (!elements[i] || strcmp(elements[i], "NULL") == 0) continue;
    for (int e = 0; e < nel; e++)
      if (names[e] == elements[i]) map[i] = e;
    if (map[i] < 0) return b200md_fail(c, B200MD_ERR_ARG, "No matching element in EAM potential file");
    count++;
  }
  if (!count) return b200md_fail(c, B200MD_ERR_ARG, "Incorrect args for pair coefficients");

  std::vector<double> frho_s((size_t) nfrho * (nrho + 1) * 7, 0.0), rhor_s((size_t) nrhor * (nr + 1) * 7, 0.0), z2r_s((size_t) nz2r * (nr + 1) * 7, 0.0);
  for (int k = 0; k < nfrho; k++) cubic_rows(nrho, drho, &frho[(size_t) k * (nrho + 1)], &frho_s[(size_t) k * (nrho + 1) * 7]);
  for (int k = 0; k < nrhor; k++) cubic_rows(nr, dr, &rhor[(size_t) k * (nr + 1)], &rhor_s[(size_t) k * (nr + 1) * 7]);
  for (int k = 0; k < nz2r; k++) cubic_rows(nr, dr, &z2r[(size_t) k * (nr + 1)], &z2r_s[(size_t) k * (nr + 1) * 7]);
  std::vector<int> t2f(nt1, 0), t2r((size_t) nt1 * nt1, 0), t2z((size_t) nt1 * nt1, 0);
  for (int i = 1; i <= ntypes; i++) {
    t2f[i] = map[i] >= 0 ? map[i] : nfrho - 1;
    for (int j = 1; j <= ntypes; j++) {
      if (map[i] < 0 || map[j] < 0) continue;   // pair hybrid's non-EAM types: never looked up
      t2r[(size_t) i * nt1 + j] = fs ? map[i] * nel + map[j] : map[i];
      const int hi = std::max(map[i], map[j]), lo = std::min(map[i], map[j]);
      t2z[(size_t) i * nt1 + j] = hi * (hi + 1) / 2 + lo;
    }
  }
  if (mass_out)
    for (int i = 1; i <= ntypes; i++) mass_out[i] = map[i] >= 0 ? mass[map[i]] : 0.0;
  if (cutmax_out) *cutmax_out = cut;
  return b200md_eam_init(c, ntypes, nrho, nr, 1.0 / drho, 1.0 / dr, (nrho - 1) * drho, cut * cut, nfrho, frho_s.data(), nrhor, rhor_s.data(),
                         nz2r, z2r_s.data(), t2f.data(), t2r.data(), t2z.data());
}
