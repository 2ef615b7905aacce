/* ======================================================================================
 * oracle.c -- TEST INFRASTRUCTURE ONLY.  Plain-C CPU restatement of the reference's
 * short-range MD hot path (single rank, orthogonal periodic box, newton on).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
 * may load this.  The product (libb200md.so) never does; it has no CPU fallback.
 *
 * Parity status: PINNED.  tests/test_oracle_*.py check this file against
 *   (1) the reference's own golden logs bench/log.6Oct16.{lj,eam}.fixed.icc.1 (thermo),
 *   (2) fixtures under tests/golden/ minted from the unmodified reference built by
 *       oracle/Makefile.ref (per-atom forces at 17 digits, neighbor pair sets, thermo at %.17g),
 *   (3) the live oracle/_ref/ref_extract binary when it is present.
 * Known, documented difference: the reference spatially re-sorts owned atoms at setup
 * (Atom::sort, src/atom.cpp:2246) which changes floating-point summation order only;
 * this restatement keeps creation order, so agreement with the reference is ~1e-13
 * relative on forces/thermo, and exact on neighbor pair sets.
 *
 * Every function cites the reference file:line (relative to /root/reference/) it follows.
 * Arrays are the reference's layout: double x[n][3] contiguous (src/memory.h:144-160),
 * 1-based type tables.
 * ====================================================================================== */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define BIG 1.0e20
#define NEIGHMASK 0x1FFFFFFF /* src/lmptype.h:61-65 (LAMMPS_SMALLBIG) */
#define SBBITS 30
#define IMGMASK 1023
#define IMGBITS 10
#define IMG2BITS 20
#define IMGMAX 512

typedef struct {
  /* ---- units (src/update.cpp:145-207) ---- */
  double boltz, mvv2e, ftm2v, nktv2p, dt;
  /* ---- box (src/domain.cpp) ---- */
  double boxlo[3], boxhi[3], prd[3];
  /* ---- atoms (src/atom.h:70-73) ---- */
  int nlocal, nghost, nmax, ntypes;
  double *x, *v, *f;          /* [nmax][3] */
  int *type, *tag, *image;
  double *mass;               /* [ntypes+1] */
  /* ---- neighbor settings (src/neighbor.cpp:306-695) ---- */
  double skin, cutneighmax, cutforce, triggersq;
  double *cutneighsq;         /* [(ntypes+1)^2] */
  int every, delay, dist_check;
  int ago, ncalls, ndanger;
  double *xhold;              /* [nlocal][3] */
  /* ---- bins (src/nbin_standard.cpp) ---- */
  int nbinx, nbiny, nbinz, mbinx, mbiny, mbinz, mbinxlo, mbinylo, mbinzlo, mbins;
  double binsizex, binsizey, binsizez, bininvx, bininvy, bininvz;
  int *binhead, *bins, *atom2bin;
  int nstencil, *stencil;
  /* ---- half neighbor list (src/neigh_list.h:52-61), CSR instead of pages ---- */
  int *numneigh;
  int64_t *firstneigh;        /* offsets into neighbors */
  int *neighbors;
  int64_t maxneighbors;
  /* ---- comm: per swap send lists (src/comm_brick.cpp:771-979) ---- */
  int nswap;
  int *sendnum, *firstrecv, **sendlist, *maxsendlist;
  int *swapdim;
  double *swapshift;          /* pbc*prd for the swap's dim */
  double cutghost;
  /* ---- pair style: 0 none, 1 lj/cut, 2 eam ---- */
  int pair_style;
  double *cutsq, *lj1, *lj2, *lj3, *lj4, *offset; /* [(ntypes+1)^2] */
  double special_lj[4];
  /* eam (src/MANYBODY/pair_eam.h) */
  int nfuncfl;
  struct funcfl {
    double mass, drho, dr, cut;
    int nrho, nr;
    double *frho, *rhor, *zr; /* 1-based */
  } *funcfl;
  int *map;                   /* [ntypes+1] type -> funcfl */
  int nrho, nr, nfrho, nrhor, nz2r;
  double drho, dr, rdr, rdrho, rhomax, cutforcesq, cutmax;
  double *frho_spline, *rhor_spline, *z2r_spline; /* [ntab][n+1][7] */
  int *type2frho, *type2rhor, *type2z2r;
  double *rho, *fp;
  /* special bonds (for the unit-test molecular fixture): CSR of (tag, which) per owned atom */
  int have_special;
  int *nspecial;              /* [nlocal][3] cumulative counts like atom->nspecial */
  int *special;               /* [nlocal][maxspecial] partner tags */
  int maxspecial;
  int special_flag[4];        /* src/neighbor.cpp: 0 = exclude, 1 = keep plain, 2 = keep with bits */
  /* ---- in.chain additions: Atom::sort, bonds, bond_style fene, fix langevin ---- */
  int sort_enabled;           /* spatial re-ordering of owned atoms at setup (src/atom.cpp:2246-2345) */
  int bond_per_atom;          /* atom->bond_per_atom */
  int *num_bond, *bond_type, *bond_atom; /* [nlocal], [nlocal][bond_per_atom] (partner tags), like src/atom.h */
  int nbondtypes;
  double *fene_k, *fene_r0, *fene_eps, *fene_sigma; /* [nbondtypes+1] src/MOLECULE/bond_fene.h */
  int nbondlist, maxbondlist, *bondlist; /* [nbondlist][3]: i1, i2, type (src/ntopo_bond_all.cpp) */
  int *tag2local, maxtag, *sametag;      /* Atom::map_set (array style) for Domain::closest_image */
  double eng_bond, virial_bond[6];
  int langevin;               /* fix langevin (src/fix_langevin.cpp), constant target, per-type mass */
  double lang_tstart, lang_tstop, lang_period;
  double mars_u[98], mars_c, mars_cd, mars_cm; int mars_i97, mars_j97; /* RanMars (src/random_mars.cpp) */
  int64_t beginstep, endstep;
  /* ---- results ---- */
  double eng_vdwl, virial[6];
  int eflag, vflag;
} orc_sys;

/* ------------------------------------------------------------------------------------ */
static void *xmalloc(size_t n) { void *p = malloc(n ? n : 1); if (!p) { fprintf(stderr, "oracle: out of memory\n"); abort(); } return p; }
static void *xcalloc(size_t n, size_t s) { void *p = calloc(n ? n : 1, s); if (!p) { fprintf(stderr, "oracle: out of memory\n"); abort(); } return p; }
static int build_bondlist(orc_sys *s);
static int bond_fene_compute(orc_sys *s);
static void langevin_post_force(orc_sys *s, int64_t ntimestep);
#define MIN(a,b) ((a) < (b) ? (a) : (b))
#define MAX(a,b) ((a) > (b) ? (a) : (b))

/* =====================================================================================
 * Input generation: lattice, create_atoms, velocity
 * ===================================================================================== */

/* src/lattice.cpp:255-262 : lj units, scale = (nbasis/(V*rho*))^(1/dim); fcc nbasis = 4 */
double orc_lattice_scale_lj(double rhostar) { return pow(4 / 1.0 / rhostar, 1.0 / 3); }

/* src/create_atoms.cpp:1301-1370 (loop k,j,i,basis), src/lattice.cpp:90-94 (fcc basis),
 * src/lattice.cpp:469-486 (lattice2box with identity primitive/rotation: x = (i+b)*scale).
 * Atoms with coord >= boxhi are skipped (:1338-1340).  Returns number created; tags 1..n
 * in creation order (src/atom.cpp:899-930, one rank). */
int orc_create_fcc(int nx, int ny, int nz, double scale, double *x, int *tag)
{
  static const double basis[4][3] = {{0, 0, 0}, {0.5, 0.5, 0}, {0.5, 0, 0.5}, {0, 0.5, 0.5}};
  double hi[3] = {nx * scale, ny * scale, nz * scale};
  int n = 0;
  for (int k = -1; k <= nz; k++)
    for (int j = -1; j <= ny; j++)
      for (int i = -1; i <= nx; i++)
        for (int m = 0; m < 4; m++) {
          double c[3];
          c[0] = (i + basis[m][0]) * scale;
          c[1] = (j + basis[m][1]) * scale;
          c[2] = (k + basis[m][2]) * scale;
          if (c[0] < 0.0 || c[0] >= hi[0] || c[1] < 0.0 || c[1] >= hi[1] || c[2] < 0.0 || c[2] >= hi[2]) continue;
          x[3 * n] = c[0]; x[3 * n + 1] = c[1]; x[3 * n + 2] = c[2];
          tag[n] = n + 1;
          n++;
        }
  return n;
}

/* src/random_park.cpp:23-49 Park-Miller minimal standard */
static double park_uniform(int *seed)
{
  int k = *seed / 127773;
  *seed = 16807 * (*seed - k * 127773) - 2836 * k;
  if (*seed < 0) *seed += 2147483647;
  return (1.0 / 2147483647) * *seed;
}

/* src/random_park.cpp:92-128 : Jenkins one-at-a-time hash over the (signed) bytes of the
 * caller's seed and the 24 coordinate bytes, keep 27 bits, 5 warm-up draws */
static int park_reset_geom(int ibase, const double *coord)
{
  unsigned int hash = 0;
  const char *str = (const char *) &ibase;
  for (int i = 0; i < (int) sizeof(int); i++) { hash += str[i]; hash += (hash << 10); hash ^= (hash >> 6); }
  str = (const char *) coord;
  for (int i = 0; i < (int) (3 * sizeof(double)); i++) { hash += str[i]; hash += (hash << 10); hash ^= (hash >> 6); }
  hash += (hash << 3); hash ^= (hash >> 11); hash += (hash << 15);
  int seed = hash & 0x7ffffff;
  if (!seed) seed = 1;
  for (int i = 0; i < 5; i++) park_uniform(&seed);
  return seed;
}

/* `velocity all create T seed loop geom` : src/velocity.cpp:329-369 (uniform dist, momentum yes,
 * rotation no), Group::vcm src/group.cpp:1121-1159, ComputeTemp src/compute_temp.cpp:55-100
 * (dof = 3N - 3), Velocity::rescale src/velocity.cpp:717-733 */
void orc_velocity_create_geom(int n, const double *x, const int *type, const double *mass, double t_desired,
                              int seed, double mvv2e, double boltz, double *v)
{
  for (int i = 0; i < n; i++) {
    int s = park_reset_geom(seed, &x[3 * i]);
    double vx = park_uniform(&s) - 0.5;
    double vy = park_uniform(&s) - 0.5;
    double vz = park_uniform(&s) - 0.5;
    double factor = 1.0 / sqrt(mass[type[i]]);
    v[3 * i] = vx * factor; v[3 * i + 1] = vy * factor; v[3 * i + 2] = vz * factor;
  }
  double p[3] = {0, 0, 0}, masstotal = 0.0;
  for (int i = 0; i < n; i++) masstotal += mass[type[i]];
  for (int i = 0; i < n; i++) {
    double m = mass[type[i]];
    p[0] += v[3 * i] * m; p[1] += v[3 * i + 1] * m; p[2] += v[3 * i + 2] * m;
  }
  if (masstotal > 0.0) { p[0] /= masstotal; p[1] /= masstotal; p[2] /= masstotal; }
  for (int i = 0; i < n; i++) { v[3 * i] -= p[0]; v[3 * i + 1] -= p[1]; v[3 * i + 2] -= p[2]; }
  double t = 0.0;
  for (int i = 0; i < n; i++)
    t += (v[3 * i] * v[3 * i] + v[3 * i + 1] * v[3 * i + 1] + v[3 * i + 2] * v[3 * i + 2]) * mass[type[i]];
  double dof = 3.0 * n - 3.0;
  double tfactor = mvv2e / (dof * boltz);
  t *= tfactor;
  double factor = sqrt(t_desired / t);
  for (int i = 0; i < 3 * n; i++) v[i] *= factor;
}

/* =====================================================================================
 * System object
 * ===================================================================================== */
orc_sys *orc_create(int ntypes)
{
  orc_sys *s = (orc_sys *) xcalloc(1, sizeof(orc_sys));
  s->ntypes = ntypes;
  int nt2 = (ntypes + 1) * (ntypes + 1);
  s->mass = (double *) xcalloc(ntypes + 1, sizeof(double));
  s->cutneighsq = (double *) xcalloc(nt2, sizeof(double));
  s->cutsq = (double *) xcalloc(nt2, sizeof(double));
  s->lj1 = (double *) xcalloc(nt2, sizeof(double));
  s->lj2 = (double *) xcalloc(nt2, sizeof(double));
  s->lj3 = (double *) xcalloc(nt2, sizeof(double));
  s->lj4 = (double *) xcalloc(nt2, sizeof(double));
  s->offset = (double *) xcalloc(nt2, sizeof(double));
  s->map = (int *) xcalloc(ntypes + 1, sizeof(int));
  for (int i = 0; i <= ntypes; i++) s->map[i] = -1;
  s->special_lj[0] = 1.0; s->special_lj[1] = s->special_lj[2] = s->special_lj[3] = 0.0; /* src/force.cpp defaults */
  s->special_flag[0] = 1; s->special_flag[1] = s->special_flag[2] = s->special_flag[3] = 0;
  /* default = lj units */
  s->boltz = 1.0; s->mvv2e = 1.0; s->ftm2v = 1.0; s->nktv2p = 1.0; s->dt = 0.005; s->skin = 0.3;
  s->every = 1; s->delay = 0; s->dist_check = 1; /* src/neighbor.cpp:62-64 defaults */
  return s;
}

static void free_swaps(orc_sys *s)
{
  if (s->sendlist) for (int i = 0; i < s->nswap; i++) free(s->sendlist[i]);
  free(s->sendlist); free(s->sendnum); free(s->firstrecv); free(s->maxsendlist); free(s->swapdim); free(s->swapshift);
  s->sendlist = NULL; s->sendnum = s->firstrecv = s->maxsendlist = s->swapdim = NULL; s->swapshift = NULL; s->nswap = 0;
}

void orc_destroy(orc_sys *s)
{
  if (!s) return;
  free(s->x); free(s->v); free(s->f); free(s->type); free(s->tag); free(s->image); free(s->mass);
  free(s->cutneighsq); free(s->xhold); free(s->binhead); free(s->bins); free(s->atom2bin); free(s->stencil);
  free(s->numneigh); free(s->firstneigh); free(s->neighbors);
  free_swaps(s);
  free(s->cutsq); free(s->lj1); free(s->lj2); free(s->lj3); free(s->lj4); free(s->offset);
  for (int i = 0; i < s->nfuncfl; i++) { free(s->funcfl[i].frho); free(s->funcfl[i].rhor); free(s->funcfl[i].zr); }
  free(s->funcfl); free(s->map);
  free(s->frho_spline); free(s->rhor_spline); free(s->z2r_spline);
  free(s->type2frho); free(s->type2rhor); free(s->type2z2r); free(s->rho); free(s->fp);
  free(s->nspecial); free(s->special);
  free(s);
}

/* units: src/update.cpp:145-164 (lj), :187-207 (metal), :166-185 (real) */
void orc_set_units(orc_sys *s, const char *style)
{
  if (!strcmp(style, "lj")) {
    s->boltz = 1.0; s->mvv2e = 1.0; s->ftm2v = 1.0; s->nktv2p = 1.0; s->dt = 0.005; s->skin = 0.3;
  } else if (!strcmp(style, "metal")) {
    s->boltz = 8.617343e-5; s->mvv2e = 1.0364269e-4; s->ftm2v = 1.0 / 1.0364269e-4; s->nktv2p = 1.6021765e6;
    s->dt = 0.001; s->skin = 2.0;
  } else if (!strcmp(style, "real")) {
    s->boltz = 0.0019872067; s->mvv2e = 48.88821291 * 48.88821291; s->ftm2v = 1.0 / 48.88821291 / 48.88821291;
    s->nktv2p = 68568.415; s->dt = 1.0; s->skin = 2.0;
  }
}
void orc_set_dt(orc_sys *s, double dt) { s->dt = dt; }
void orc_set_box(orc_sys *s, const double *lo, const double *hi)
{
  for (int d = 0; d < 3; d++) { s->boxlo[d] = lo[d]; s->boxhi[d] = hi[d]; s->prd[d] = hi[d] - lo[d]; }
}
void orc_set_mass(orc_sys *s, int itype, double m) { s->mass[itype] = m; }
void orc_set_neighbor(orc_sys *s, double skin, int every, int delay, int check)
{
  s->skin = skin; s->every = every; s->delay = delay; s->dist_check = check;
}

static void grow_atoms(orc_sys *s, int nmax)
{
  if (nmax <= s->nmax) return;
  nmax = nmax + nmax / 4 + 1024;
  s->x = (double *) realloc(s->x, sizeof(double) * 3 * nmax);
  s->v = (double *) realloc(s->v, sizeof(double) * 3 * nmax);
  s->f = (double *) realloc(s->f, sizeof(double) * 3 * nmax);
  s->type = (int *) realloc(s->type, sizeof(int) * nmax);
  s->tag = (int *) realloc(s->tag, sizeof(int) * nmax);
  s->image = (int *) realloc(s->image, sizeof(int) * nmax);
  s->bins = (int *) realloc(s->bins, sizeof(int) * nmax);
  s->atom2bin = (int *) realloc(s->atom2bin, sizeof(int) * nmax);
  s->rho = (double *) realloc(s->rho, sizeof(double) * nmax);
  s->fp = (double *) realloc(s->fp, sizeof(double) * nmax);
  s->nmax = nmax;
}

void orc_set_atoms(orc_sys *s, int n, const double *x, const double *v, const int *type, const int *tag)
{
  grow_atoms(s, n);
  s->nlocal = n; s->nghost = 0;
  memcpy(s->x, x, sizeof(double) * 3 * n);
  if (v) memcpy(s->v, v, sizeof(double) * 3 * n); else memset(s->v, 0, sizeof(double) * 3 * n);
  memcpy(s->type, type, sizeof(int) * n);
  memcpy(s->tag, tag, sizeof(int) * n);
  for (int i = 0; i < n; i++) s->image[i] = (IMGMAX << IMG2BITS) | (IMGMAX << IMGBITS) | IMGMAX;
  free(s->xhold); s->xhold = (double *) xmalloc(sizeof(double) * 3 * n);
  free(s->numneigh); s->numneigh = (int *) xcalloc(n, sizeof(int));
  free(s->firstneigh); s->firstneigh = (int64_t *) xcalloc(n + 1, sizeof(int64_t));
}

/* special-bond partners per owned atom, in the reference's layout (atom->nspecial[i][0..2] cumulative,
 * atom->special[i][k] partner tags), and special_lj[1..3] -> special_flag as in
 * src/neighbor.cpp:523-541 (0.0 -> 0 exclude; 1.0 -> 1 plain; else 2 with bits) */
void orc_set_special(orc_sys *s, int maxspecial, const int *nspecial, const int *special, const double *special_lj)
{
  int n = s->nlocal;
  s->have_special = 1; s->maxspecial = maxspecial;
  free(s->nspecial); free(s->special);
  s->nspecial = (int *) xmalloc(sizeof(int) * 3 * n); memcpy(s->nspecial, nspecial, sizeof(int) * 3 * n);
  s->special = (int *) xmalloc(sizeof(int) * (size_t) maxspecial * n); memcpy(s->special, special, sizeof(int) * (size_t) maxspecial * n);
  s->special_lj[0] = 1.0;
  for (int k = 1; k <= 3; k++) {
    s->special_lj[k] = special_lj[k];
    if (special_lj[k] == 0.0) s->special_flag[k] = 0;
    else if (special_lj[k] == 1.0) s->special_flag[k] = 1;
    else s->special_flag[k] = 2;
  }
}

/* =====================================================================================
 * pair_style lj/cut : coefficients  (src/pair_lj_cut.cpp:501-524, mixing src/pair.cpp:693-729)
 * epsilon/sigma/cut are full (ntypes+1)^2 tables already mixed by the caller OR -1 to mix.
 * ===================================================================================== */
void orc_pair_lj_cut(orc_sys *s, const double *epsilon, const double *sigma, const double *cut, int offset_flag)
{
  int nt = s->ntypes, ld = nt + 1;
  s->pair_style = 1;
  double cutmax = 0.0;
  for (int i = 1; i <= nt; i++)
    for (int j = i; j <= nt; j++) {
      double e = epsilon[i * ld + j], sg = sigma[i * ld + j], c = cut[i * ld + j];
      s->lj1[i * ld + j] = 48.0 * e * pow(sg, 12.0);
      s->lj2[i * ld + j] = 24.0 * e * pow(sg, 6.0);
      s->lj3[i * ld + j] = 4.0 * e * pow(sg, 12.0);
      s->lj4[i * ld + j] = 4.0 * e * pow(sg, 6.0);
      if (offset_flag && c > 0.0) {
        double ratio = sg / c;
        s->offset[i * ld + j] = 4.0 * e * (pow(ratio, 12.0) - pow(ratio, 6.0));
      } else s->offset[i * ld + j] = 0.0;
      s->cutsq[i * ld + j] = c * c;                        /* src/pair.cpp:258-259 */
      s->lj1[j * ld + i] = s->lj1[i * ld + j]; s->lj2[j * ld + i] = s->lj2[i * ld + j];
      s->lj3[j * ld + i] = s->lj3[i * ld + j]; s->lj4[j * ld + i] = s->lj4[i * ld + j];
      s->offset[j * ld + i] = s->offset[i * ld + j]; s->cutsq[j * ld + i] = s->cutsq[i * ld + j];
      cutmax = MAX(cutmax, c);
    }
  s->cutforce = cutmax;
  /* src/neighbor.cpp:330-344: cutneighsq[i][j] = (cut_ij + skin)^2 */
  double skin = s->skin; s->cutneighmax = 0.0;
  for (int i = 1; i <= nt; i++)
    for (int j = 1; j <= nt; j++) {
      double c = sqrt(s->cutsq[i * ld + j]) + skin;
      s->cutneighsq[i * ld + j] = c * c;
      s->cutneighmax = MAX(s->cutneighmax, c);
    }
}

/* =====================================================================================
 * pair_style eam (funcfl) : file reader, re-gridding, splines
 * ===================================================================================== */

/* src/MANYBODY/pair_eam.cpp:460-526 : line 1 comment; line 2 "Z mass ..."; line 3 "nrho drho nr dr cut";
 * then nrho frho values, nr zr values, nr rhor values, free format */
int orc_eam_read_funcfl(orc_sys *s, const char *filename)
{
  FILE *fp = fopen(filename, "r");
  if (!fp) return -1;
  char line[1024];
  s->funcfl = (struct funcfl *) realloc(s->funcfl, sizeof(struct funcfl) * (s->nfuncfl + 1));
  struct funcfl *file = &s->funcfl[s->nfuncfl];
  if (!fgets(line, sizeof line, fp)) { fclose(fp); return -2; }
  /* PotentialFileReader unit conversion (src/MANYBODY/pair_eam.cpp:466-505, src/utils.cpp:1488-1506) */
  double conv = 1.0;
  {
    int sys_real = fabs(s->boltz - 0.0019872067) < 1e-12, sys_metal = fabs(s->boltz - 8.617343e-5) < 1e-15;
    if (strstr(line, "UNITS: metal") && sys_real) conv = 23.060549;
    else if (strstr(line, "UNITS: real") && sys_metal) conv = 1.0 / 23.060549;
  }
  int iz;
  if (!fgets(line, sizeof line, fp) || sscanf(line, "%d %lg", &iz, &file->mass) != 2) { fclose(fp); return -2; }
  if (!fgets(line, sizeof line, fp) ||
      sscanf(line, "%d %lg %d %lg %lg", &file->nrho, &file->drho, &file->nr, &file->dr, &file->cut) != 5) { fclose(fp); return -2; }
  file->frho = (double *) xcalloc(file->nrho + 1, sizeof(double));
  file->rhor = (double *) xcalloc(file->nr + 1, sizeof(double));
  file->zr = (double *) xcalloc(file->nr + 1, sizeof(double));
  for (int i = 1; i <= file->nrho; i++) if (fscanf(fp, "%lg", &file->frho[i]) != 1) { fclose(fp); return -3; }
  for (int i = 1; i <= file->nr; i++) if (fscanf(fp, "%lg", &file->zr[i]) != 1) { fclose(fp); return -3; }
  for (int i = 1; i <= file->nr; i++) if (fscanf(fp, "%lg", &file->rhor[i]) != 1) { fclose(fp); return -3; }
  fclose(fp);
  if (conv != 1.0) {
    double sq = sqrt(conv);
    for (int i = 1; i <= file->nrho; i++) file->frho[i] *= conv;
    for (int i = 1; i <= file->nr; i++) file->zr[i] *= sq;
  }
  return s->nfuncfl++;
}

/* pair_coeff i i file : src/MANYBODY/pair_eam.cpp:372-412 (sets map and the type's mass) */
void orc_eam_map_type(orc_sys *s, int itype, int ifuncfl)
{
  s->map[itype] = ifuncfl;
  s->mass[itype] = s->funcfl[ifuncfl].mass;
}

/* 4-point Lagrange re-gridding, identical for frho/rhor/zr: src/MANYBODY/pair_eam.cpp:580-596 */
static double regrid(const double *tab, int ntab, double dtab, double r)
{
  double sixth = 1.0 / 6.0;
  double p = r / dtab + 1.0;
  int k = (int) p;
  k = MIN(k, ntab - 2);
  k = MAX(k, 2);
  p -= k;
  p = MIN(p, 2.0);
  double cof1 = -sixth * p * (p - 1.0) * (p - 2.0);
  double cof2 = 0.5 * (p * p - 1.0) * (p - 2.0);
  double cof3 = -0.5 * p * (p + 1.0) * (p - 2.0);
  double cof4 = sixth * p * (p * p - 1.0);
  return cof1 * tab[k - 1] + cof2 * tab[k] + cof3 * tab[k + 1] + cof4 * tab[k + 2];
}

/* src/MANYBODY/pair_eam.cpp:769-797 ; spline is [n+1][7], rows 1..n used */
static void interpolate(int n, double delta, const double *f, double *spline)
{
#define SP(m, k) spline[(m) * 7 + (k)]
  for (int m = 1; m <= n; m++) SP(m, 6) = f[m];
  SP(1, 5) = SP(2, 6) - SP(1, 6);
  SP(2, 5) = 0.5 * (SP(3, 6) - SP(1, 6));
  SP(n - 1, 5) = 0.5 * (SP(n, 6) - SP(n - 2, 6));
  SP(n, 5) = SP(n, 6) - SP(n - 1, 6);
  for (int m = 3; m <= n - 2; m++)
    SP(m, 5) = ((SP(m - 2, 6) - SP(m + 2, 6)) + 8.0 * (SP(m + 1, 6) - SP(m - 1, 6))) / 12.0;
  for (int m = 1; m <= n - 1; m++) {
    SP(m, 4) = 3.0 * (SP(m + 1, 6) - SP(m, 6)) - 2.0 * SP(m, 5) - SP(m + 1, 5);
    SP(m, 3) = SP(m, 5) + SP(m + 1, 5) - 2.0 * (SP(m + 1, 6) - SP(m, 6));
  }
  SP(n, 4) = 0.0;
  SP(n, 3) = 0.0;
  for (int m = 1; m <= n; m++) {
    SP(m, 2) = SP(m, 5) / delta;
    SP(m, 1) = 2.0 * SP(m, 4) / delta;
    SP(m, 0) = 3.0 * SP(m, 3) / delta;
  }
#undef SP
}

/* file2array + array2spline + init_one : src/MANYBODY/pair_eam.cpp:533-765, 435-456 */
void orc_pair_eam_init(orc_sys *s)
{
  int nt = s->ntypes, ld = nt + 1;
  s->pair_style = 2;
  double rmax = 0.0;
  s->dr = s->drho = s->rhomax = 0.0;
  for (int i = 0; i < s->nfuncfl; i++) {
    int active = 0;
    for (int j = 1; j <= nt; j++) if (s->map[j] == i) active = 1;
    if (!active) continue;
    struct funcfl *file = &s->funcfl[i];
    s->dr = MAX(s->dr, file->dr);
    s->drho = MAX(s->drho, file->drho);
    rmax = MAX(rmax, (file->nr - 1) * file->dr);
    s->rhomax = MAX(s->rhomax, (file->nrho - 1) * file->drho);
  }
  s->nr = (int) (rmax / s->dr + 0.5);
  s->nrho = (int) (s->rhomax / s->drho + 0.5);
  int nr = s->nr, nrho = s->nrho;

  s->nfrho = s->nfuncfl + 1;
  double *frho = (double *) xcalloc((size_t) s->nfrho * (nrho + 1), sizeof(double));
  for (int i = 0; i < s->nfuncfl; i++) {
    struct funcfl *file = &s->funcfl[i];
    for (int m = 1; m <= nrho; m++) frho[i * (nrho + 1) + m] = regrid(file->frho, file->nrho, file->drho, (m - 1) * s->drho);
  }
  free(s->type2frho); s->type2frho = (int *) xcalloc(ld, sizeof(int));
  for (int i = 1; i <= nt; i++) s->type2frho[i] = (s->map[i] >= 0) ? s->map[i] : s->nfrho - 1;

  s->nrhor = s->nfuncfl;
  double *rhor = (double *) xcalloc((size_t) s->nrhor * (nr + 1), sizeof(double));
  for (int i = 0; i < s->nfuncfl; i++) {
    struct funcfl *file = &s->funcfl[i];
    for (int m = 1; m <= nr; m++) rhor[i * (nr + 1) + m] = regrid(file->rhor, file->nr, file->dr, (m - 1) * s->dr);
  }
  free(s->type2rhor); s->type2rhor = (int *) xcalloc(ld * ld, sizeof(int));
  for (int i = 1; i <= nt; i++) for (int j = 1; j <= nt; j++) s->type2rhor[i * ld + j] = s->map[i];

  s->nz2r = s->nfuncfl * (s->nfuncfl + 1) / 2;
  double *z2r = (double *) xcalloc((size_t) s->nz2r * (nr + 1), sizeof(double));
  int n = 0;
  for (int i = 0; i < s->nfuncfl; i++)
    for (int j = 0; j <= i; j++) {
      struct funcfl *ifile = &s->funcfl[i], *jfile = &s->funcfl[j];
      for (int m = 1; m <= nr; m++) {
        double r = (m - 1) * s->dr;
        double zri = regrid(ifile->zr, ifile->nr, ifile->dr, r);
        double zrj = regrid(jfile->zr, jfile->nr, jfile->dr, r);
        z2r[n * (nr + 1) + m] = 27.2 * 0.529 * zri * zrj;
      }
      n++;
    }
  free(s->type2z2r); s->type2z2r = (int *) xcalloc(ld * ld, sizeof(int));
  for (int i = 1; i <= nt; i++)
    for (int j = 1; j <= nt; j++) {
      int irow = s->map[i], icol = s->map[j];
      if (irow == -1 || icol == -1) { s->type2z2r[i * ld + j] = 0; continue; }
      if (irow < icol) { irow = s->map[j]; icol = s->map[i]; }
      n = 0;
      for (int m = 0; m < irow; m++) n += m + 1;
      n += icol;
      s->type2z2r[i * ld + j] = n;
    }

  s->rdr = 1.0 / s->dr;
  s->rdrho = 1.0 / s->drho;
  free(s->frho_spline); free(s->rhor_spline); free(s->z2r_spline);
  s->frho_spline = (double *) xcalloc((size_t) s->nfrho * (nrho + 1) * 7, sizeof(double));
  s->rhor_spline = (double *) xcalloc((size_t) s->nrhor * (nr + 1) * 7, sizeof(double));
  s->z2r_spline = (double *) xcalloc((size_t) s->nz2r * (nr + 1) * 7, sizeof(double));
  for (int i = 0; i < s->nfrho; i++) interpolate(nrho, s->drho, &frho[i * (nrho + 1)], &s->frho_spline[(size_t) i * (nrho + 1) * 7]);
  for (int i = 0; i < s->nrhor; i++) interpolate(nr, s->dr, &rhor[i * (nr + 1)], &s->rhor_spline[(size_t) i * (nr + 1) * 7]);
  for (int i = 0; i < s->nz2r; i++) interpolate(nr, s->dr, &z2r[i * (nr + 1)], &s->z2r_spline[(size_t) i * (nr + 1) * 7]);
  free(frho); free(rhor); free(z2r);

  s->cutmax = 0.0;
  for (int m = 0; m < s->nfuncfl; m++) s->cutmax = MAX(s->cutmax, s->funcfl[m].cut);
  s->cutforcesq = s->cutmax * s->cutmax;
  s->cutforce = s->cutmax;
  double c = s->cutmax + s->skin;
  for (int i = 1; i <= nt; i++) for (int j = 1; j <= nt; j++) { s->cutsq[i * ld + j] = s->cutforcesq; s->cutneighsq[i * ld + j] = c * c; }
  s->cutneighmax = c;
}

/* -------------------------------------------------------------------------------------
 * pair_style eam/alloy (DYNAMO setfl) and eam/fs (Finnis-Sinclair):
 *   pair_coeff * * file El1 El2 ...   PairEAMAlloy::coeff / read_file / file2array
 *   (src/MANYBODY/pair_eam_alloy.cpp:55-337), PairEAMFS (src/MANYBODY/pair_eam_fs.cpp:55-345);
 *   array2spline and init_one are PairEAM's (src/MANYBODY/pair_eam.cpp:720-765, 435-456).
 * The tables are used on the file's own grid (no re-gridding).  elements[i] (i = 1..ntypes) = element name of atom
 * type i, or "NULL".  Unit conversion as the reference does it, INCLUDING its loop bounds: the setfl reader converts
 * frho[1..nrho-1] and z2r[1..nr-1] (the last knot keeps the file's units, pair_eam_alloy.cpp:181,189), the fs reader
 * converts all knots (pair_eam_fs.cpp:165,177).
 * ------------------------------------------------------------------------------------- */
static int next_content_line(FILE *fp, char *line, int n)
{
  while (fgets(line, n, fp)) {
    const char *q = line;
    while (*q == ' ' || *q == '\t' || *q == '\r' || *q == '\n') q++;
    if (*q) return 1;
  }
  return 0;
}
static int read_dvector(FILE *fp, double *dst, int n)
{
  for (int i = 0; i < n; i++) if (fscanf(fp, "%lg", &dst[i]) != 1) return 0;
  return 1;
}
int orc_pair_eam_setfl(orc_sys *s, const char *filename, int fs, const char *const *elements)
{
  FILE *fp = fopen(filename, "r");
  if (!fp) return -1;
  char line[4096], first[4096];
  if (!fgets(first, sizeof first, fp) || !fgets(line, sizeof line, fp) || !fgets(line, sizeof line, fp)) { fclose(fp); return -2; }
  double conv = 1.0;
  {
    int sys_real = fabs(s->boltz - 0.0019872067) < 1e-12, sys_metal = fabs(s->boltz - 8.617343e-5) < 1e-15;
    if (strstr(first, "UNITS: metal") && sys_real) conv = 23.060549;
    else if (strstr(first, "UNITS: real") && sys_metal) conv = 1.0 / 23.060549;
  }
  if (!next_content_line(fp, line, sizeof line)) { fclose(fp); return -2; }
  int nel = 0, off = 0;
  char names[16][32];
  if (sscanf(line, "%d%n", &nel, &off) != 1 || nel < 1 || nel > 16) { fclose(fp); return -2; }
  {
    const char *q = line + off;
    for (int e = 0; e < nel; e++) {
      int used = 0;
      if (sscanf(q, "%31s%n", names[e], &used) != 1) { fclose(fp); return -2; }   /* "Incorrect element names" */
      q += used;
    }
  }
  int nrho, nr; double drho, dr, cut;
  if (!next_content_line(fp, line, sizeof line) || sscanf(line, "%d %lg %d %lg %lg", &nrho, &drho, &nr, &dr, &cut) != 5 ||
      nrho <= 0 || nr <= 0 || dr <= 0.0) { fclose(fp); return -2; }
  int nrhor_file = fs ? nel * nel : nel, npair = nel * (nel + 1) / 2;
  double *mass = (double *) xcalloc(nel, sizeof(double));
  double *frho = (double *) xcalloc((size_t) (nel + 1) * (nrho + 1), sizeof(double));    /* + the all-zero table */
  double *rhor = (double *) xcalloc((size_t) nrhor_file * (nr + 1), sizeof(double));
  double *z2r = (double *) xcalloc((size_t) npair * (nr + 1), sizeof(double));
  int ok = 1;
  for (int e = 0; e < nel && ok; e++) {
    int iz;
    ok = next_content_line(fp, line, sizeof line) && sscanf(line, "%d %lg", &iz, &mass[e]) == 2 &&
         read_dvector(fp, &frho[(size_t) e * (nrho + 1) + 1], nrho);
    if (!ok) break;
    if (fs) {
      for (int j = 0; j < nel && ok; j++) ok = read_dvector(fp, &rhor[(size_t) (e * nel + j) * (nr + 1) + 1], nr);
      if (conv != 1.0) for (int m = 1; m <= nrho; m++) frho[(size_t) e * (nrho + 1) + m] *= conv;
    } else {
      ok = read_dvector(fp, &rhor[(size_t) e * (nr + 1) + 1], nr);
      if (conv != 1.0) for (int m = 1; m < nrho; m++) frho[(size_t) e * (nrho + 1) + m] *= conv;
    }
  }
  int n = 0;
  for (int i = 0; i < nel && ok; i++)
    for (int j = 0; j <= i && ok; j++, n++) {
      ok = read_dvector(fp, &z2r[(size_t) n * (nr + 1) + 1], nr);
      if (conv != 1.0) for (int m = 1; m < nr + (fs ? 1 : 0); m++) z2r[(size_t) n * (nr + 1) + m] *= conv;
    }
  fclose(fp);
  if (!ok) { free(mass); free(frho); free(rhor); free(z2r); return -3; }

  int nt = s->ntypes, ld = nt + 1, count = 0;
  for (int i = 1; i <= nt; i++) {
    s->map[i] = -1;
    if (!elements[i] || strcmp(elements[i], "NULL") == 0) continue;
    for (int e = 0; e < nel; e++) if (strcmp(elements[i], names[e]) == 0) s->map[i] = e;
    if (s->map[i] < 0) { free(mass); free(frho); free(rhor); free(z2r); return -4; }   /* "No matching element in EAM potential file" */
    s->mass[i] = mass[s->map[i]];
    count++;
  }
  free(mass);
  if (!count) { free(frho); free(rhor); free(z2r); return -5; }
  s->pair_style = 2;
  s->nrho = nrho; s->nr = nr; s->drho = drho; s->dr = dr; s->rhomax = (nrho - 1) * drho;
  s->nfrho = nel + 1; s->nrhor = nrhor_file; s->nz2r = npair;
  free(s->type2frho); s->type2frho = (int *) xcalloc(ld, sizeof(int));
  free(s->type2rhor); s->type2rhor = (int *) xcalloc(ld * ld, sizeof(int));
  free(s->type2z2r); s->type2z2r = (int *) xcalloc(ld * ld, sizeof(int));
  for (int i = 1; i <= nt; i++) {
    s->type2frho[i] = s->map[i] >= 0 ? s->map[i] : s->nfrho - 1;
    for (int j = 1; j <= nt; j++) {
      s->type2rhor[i * ld + j] = fs ? s->map[i] * nel + s->map[j] : s->map[i];
      int irow = s->map[i], icol = s->map[j];
      if (irow == -1 || icol == -1) { s->type2z2r[i * ld + j] = 0; if (fs) s->type2rhor[i * ld + j] = 0; continue; }
      if (irow < icol) { irow = s->map[j]; icol = s->map[i]; }
      int k = 0;
      for (int m = 0; m < irow; m++) k += m + 1;
      s->type2z2r[i * ld + j] = k + icol;
    }
  }
  s->rdr = 1.0 / dr; s->rdrho = 1.0 / drho;
  free(s->frho_spline); free(s->rhor_spline); free(s->z2r_spline);
  s->frho_spline = (double *) xcalloc((size_t) s->nfrho * (nrho + 1) * 7, sizeof(double));
  s->rhor_spline = (double *) xcalloc((size_t) s->nrhor * (nr + 1) * 7, sizeof(double));
  s->z2r_spline = (double *) xcalloc((size_t) s->nz2r * (nr + 1) * 7, sizeof(double));
  for (int i = 0; i < s->nfrho; i++) interpolate(nrho, drho, &frho[(size_t) i * (nrho + 1)], &s->frho_spline[(size_t) i * (nrho + 1) * 7]);
  for (int i = 0; i < s->nrhor; i++) interpolate(nr, dr, &rhor[(size_t) i * (nr + 1)], &s->rhor_spline[(size_t) i * (nr + 1) * 7]);
  for (int i = 0; i < s->nz2r; i++) interpolate(nr, dr, &z2r[(size_t) i * (nr + 1)], &s->z2r_spline[(size_t) i * (nr + 1) * 7]);
  free(frho); free(rhor); free(z2r);
  s->cutmax = cut;                                   /* init_one: cutmax = setfl->cut (pair_eam.cpp:446-450) */
  s->cutforcesq = cut * cut; s->cutforce = cut;
  double c = cut + s->skin;
  for (int i = 1; i <= nt; i++) for (int j = 1; j <= nt; j++) { s->cutsq[i * ld + j] = s->cutforcesq; s->cutneighsq[i * ld + j] = c * c; }
  s->cutneighmax = c;
  return 0;
}

/* table accessors for tests (sizes, then raw copies) */
void orc_eam_dims(orc_sys *s, int *out) { out[0] = s->nrho; out[1] = s->nr; out[2] = s->nfrho; out[3] = s->nrhor; out[4] = s->nz2r; }
void orc_eam_scalars(orc_sys *s, double *out) { out[0] = s->rdr; out[1] = s->rdrho; out[2] = s->rhomax; out[3] = s->cutforcesq; out[4] = s->cutmax; }
void orc_eam_tables(orc_sys *s, double *frho, double *rhor, double *z2r, int *t2frho, int *t2rhor, int *t2z2r)
{
  int ld = s->ntypes + 1;
  memcpy(frho, s->frho_spline, sizeof(double) * (size_t) s->nfrho * (s->nrho + 1) * 7);
  memcpy(rhor, s->rhor_spline, sizeof(double) * (size_t) s->nrhor * (s->nr + 1) * 7);
  memcpy(z2r, s->z2r_spline, sizeof(double) * (size_t) s->nz2r * (s->nr + 1) * 7);
  memcpy(t2frho, s->type2frho, sizeof(int) * ld);
  memcpy(t2rhor, s->type2rhor, sizeof(int) * ld * ld);
  memcpy(t2z2r, s->type2z2r, sizeof(int) * ld * ld);
}
double orc_get_mass(orc_sys *s, int itype) { return s->mass[itype]; }

/* =====================================================================================
 * Domain::pbc  (src/domain.cpp:530-660), orthogonal, all dims periodic
 * ===================================================================================== */
void orc_pbc(orc_sys *s)
{
  for (int i = 0; i < s->nlocal; i++) {
    for (int d = 0; d < 3; d++) {
      double *xi = &s->x[3 * i + d];
      int shiftbits = d * IMGBITS;
      if (*xi < s->boxlo[d]) {
        *xi += s->prd[d];
        int idim = (s->image[i] >> shiftbits) & IMGMASK;
        int other = s->image[i] ^ (idim << shiftbits);
        idim--; idim &= IMGMASK;
        s->image[i] = other | (idim << shiftbits);
      }
      if (*xi >= s->boxhi[d]) {
        *xi -= s->prd[d];
        *xi = MAX(*xi, s->boxlo[d]);
        int idim = (s->image[i] >> shiftbits) & IMGMASK;
        int other = s->image[i] ^ (idim << shiftbits);
        idim++; idim &= IMGMASK;
        s->image[i] = other | (idim << shiftbits);
      }
    }
  }
}

/* =====================================================================================
 * CommBrick::setup + borders + forward/reverse, one rank, periodic
 * (src/comm_brick.cpp:191-483, 771-979, 538-639; AtomVec::pack_border src/atom_vec.cpp:791-827)
 * ===================================================================================== */
static void comm_setup(orc_sys *s)
{
  free_swaps(s);
  s->cutghost = s->cutneighmax;                 /* src/comm.cpp:681-... (no user cutoff) */
  int maxneed[3];
  for (int d = 0; d < 3; d++) maxneed[d] = (int) (s->cutghost * 1 / s->prd[d]) + 1;
  s->nswap = 2 * (maxneed[0] + maxneed[1] + maxneed[2]);
  s->sendnum = (int *) xcalloc(s->nswap, sizeof(int));
  s->firstrecv = (int *) xcalloc(s->nswap, sizeof(int));
  s->maxsendlist = (int *) xcalloc(s->nswap, sizeof(int));
  s->sendlist = (int **) xcalloc(s->nswap, sizeof(int *));
  s->swapdim = (int *) xcalloc(s->nswap, sizeof(int));
  s->swapshift = (double *) xcalloc(s->nswap, sizeof(double));
}

void orc_borders(orc_sys *s)
{
  comm_setup(s);
  s->nghost = 0;
  int iswap = 0, nfirst = 0, nlast = 0;
  for (int dim = 0; dim < 3; dim++) {
    int maxneed = (int) (s->cutghost / s->prd[dim]) + 1;
    nlast = 0;
    for (int ineed = 0; ineed < 2 * maxneed; ineed++) {
      double lo, hi, shift;
      if (ineed % 2 == 0) {
        lo = (ineed < 2) ? -BIG : 0.5 * (s->boxlo[dim] + s->boxhi[dim]);
        hi = s->boxlo[dim] + s->cutghost;
        shift = 1 * s->prd[dim];                /* myloc == 0: pbc = +1 */
        nfirst = nlast;
        nlast = s->nlocal + s->nghost;
      } else {
        lo = s->boxhi[dim] - s->cutghost;
        hi = (ineed < 2) ? BIG : 0.5 * (s->boxlo[dim] + s->boxhi[dim]);
        shift = -1 * s->prd[dim];               /* myloc == procgrid-1: pbc = -1 */
      }
      int nsend = 0, cap = 1024;
      int *list = (int *) xmalloc(sizeof(int) * cap);
      for (int i = nfirst; i < nlast; i++)
        if (s->x[3 * i + dim] >= lo && s->x[3 * i + dim] <= hi) {
          if (nsend == cap) { cap *= 2; list = (int *) realloc(list, sizeof(int) * cap); }
          list[nsend++] = i;
        }
      grow_atoms(s, s->nlocal + s->nghost + nsend);
      int first = s->nlocal + s->nghost;
      for (int k = 0; k < nsend; k++) {
        int j = list[k], g = first + k;
        s->x[3 * g] = s->x[3 * j]; s->x[3 * g + 1] = s->x[3 * j + 1]; s->x[3 * g + 2] = s->x[3 * j + 2];
        s->x[3 * g + dim] = s->x[3 * j + dim] + shift;
        s->type[g] = s->type[j]; s->tag[g] = s->tag[j];
      }
      s->sendlist[iswap] = list; s->sendnum[iswap] = nsend; s->maxsendlist[iswap] = cap;
      s->firstrecv[iswap] = first; s->swapdim[iswap] = dim; s->swapshift[iswap] = shift;
      s->nghost += nsend;
      iswap++;
    }
  }
}

/* src/comm_brick.cpp:538-591 + AtomVec::pack_comm src/atom_vec.cpp:349-379 */
void orc_forward_comm(orc_sys *s)
{
  for (int iswap = 0; iswap < s->nswap; iswap++) {
    int dim = s->swapdim[iswap]; double shift = s->swapshift[iswap];
    for (int k = 0; k < s->sendnum[iswap]; k++) {
      int j = s->sendlist[iswap][k], g = s->firstrecv[iswap] + k;
      s->x[3 * g] = s->x[3 * j]; s->x[3 * g + 1] = s->x[3 * j + 1]; s->x[3 * g + 2] = s->x[3 * j + 2];
      s->x[3 * g + dim] = s->x[3 * j + dim] + shift;
    }
  }
}

/* src/comm_brick.cpp:598-639 + AtomVec::unpack_reverse src/atom_vec.cpp:724-735 */
void orc_reverse_comm(orc_sys *s)
{
  for (int iswap = s->nswap - 1; iswap >= 0; iswap--)
    for (int k = 0; k < s->sendnum[iswap]; k++) {
      int j = s->sendlist[iswap][k], g = s->firstrecv[iswap] + k;
      s->f[3 * j] += s->f[3 * g]; s->f[3 * j + 1] += s->f[3 * g + 1]; s->f[3 * j + 2] += s->f[3 * g + 2];
    }
}

/* PairEAM::pack/unpack_forward/reverse_comm src/MANYBODY/pair_eam.cpp:855-901 through
 * CommBrick::forward_comm(Pair*) / reverse_comm(Pair*) src/comm_brick.cpp:986-1053 */
static void forward_comm_scalar(orc_sys *s, double *a)
{
  for (int iswap = 0; iswap < s->nswap; iswap++)
    for (int k = 0; k < s->sendnum[iswap]; k++) a[s->firstrecv[iswap] + k] = a[s->sendlist[iswap][k]];
}
static void reverse_comm_scalar(orc_sys *s, double *a)
{
  for (int iswap = s->nswap - 1; iswap >= 0; iswap--)
    for (int k = 0; k < s->sendnum[iswap]; k++) a[s->sendlist[iswap][k]] += a[s->firstrecv[iswap] + k];
}

/* =====================================================================================
 * NBinStandard::setup_bins / bin_atoms, NBin::coord2bin, NStencilHalfBin3d
 * ===================================================================================== */
void orc_setup_bins(orc_sys *s)  /* src/nbin_standard.cpp:82-215 */
{
  const double SMALL = 1.0e-6;
  double bbox[3], bsubboxlo[3], bsubboxhi[3];
  for (int d = 0; d < 3; d++) {
    bsubboxlo[d] = s->boxlo[d] - s->cutghost;
    bsubboxhi[d] = s->boxhi[d] + s->cutghost;
    bbox[d] = s->boxhi[d] - s->boxlo[d];
  }
  double binsize_optimal = 0.5 * s->cutneighmax;
  if (binsize_optimal == 0.0) binsize_optimal = bbox[0];
  double binsizeinv = 1.0 / binsize_optimal;
  s->nbinx = (int) (bbox[0] * binsizeinv);
  s->nbiny = (int) (bbox[1] * binsizeinv);
  s->nbinz = (int) (bbox[2] * binsizeinv);
  if (s->nbinx == 0) s->nbinx = 1;
  if (s->nbiny == 0) s->nbiny = 1;
  if (s->nbinz == 0) s->nbinz = 1;
  s->binsizex = bbox[0] / s->nbinx; s->binsizey = bbox[1] / s->nbiny; s->binsizez = bbox[2] / s->nbinz;
  s->bininvx = 1.0 / s->binsizex; s->bininvy = 1.0 / s->binsizey; s->bininvz = 1.0 / s->binsizez;
  int mhi[3], mlo[3];
  double inv[3] = {s->bininvx, s->bininvy, s->bininvz};
  for (int d = 0; d < 3; d++) {
    double coord = bsubboxlo[d] - SMALL * bbox[d];
    mlo[d] = (int) ((coord - s->boxlo[d]) * inv[d]);
    if (coord < s->boxlo[d]) mlo[d] = mlo[d] - 1;
    coord = bsubboxhi[d] + SMALL * bbox[d];
    mhi[d] = (int) ((coord - s->boxlo[d]) * inv[d]);
    mlo[d] -= 1; mhi[d] += 1;
  }
  s->mbinxlo = mlo[0]; s->mbinylo = mlo[1]; s->mbinzlo = mlo[2];
  s->mbinx = mhi[0] - mlo[0] + 1; s->mbiny = mhi[1] - mlo[1] + 1; s->mbinz = mhi[2] - mlo[2] + 1;
  s->mbins = s->mbinx * s->mbiny * s->mbinz + 1;
  free(s->binhead); s->binhead = (int *) xmalloc(sizeof(int) * s->mbins);

  /* NStencil::create_setup src/nstencil.cpp:213-246 ; NStencilHalfBin3d::create src/nstencil_half_bin_3d.cpp:26-38 ;
   * bin_distance src/nstencil.cpp:406-423 */
  int sx = (int) (s->cutneighmax * s->bininvx); if (sx * s->binsizex < s->cutneighmax) sx++;
  int sy = (int) (s->cutneighmax * s->bininvy); if (sy * s->binsizey < s->cutneighmax) sy++;
  int sz = (int) (s->cutneighmax * s->bininvz); if (sz * s->binsizez < s->cutneighmax) sz++;
  free(s->stencil); s->stencil = (int *) xmalloc(sizeof(int) * (2 * sx + 1) * (2 * sy + 1) * (2 * sz + 1));
  s->nstencil = 0;
  double cutneighmaxsq = s->cutneighmax * s->cutneighmax;
  for (int k = 0; k <= sz; k++)
    for (int j = -sy; j <= sy; j++)
      for (int i = -sx; i <= sx; i++)
        if (k > 0 || j > 0 || (j == 0 && i > 0)) {
          double delx = (i > 0) ? (i - 1) * s->binsizex : (i == 0 ? 0.0 : (i + 1) * s->binsizex);
          double dely = (j > 0) ? (j - 1) * s->binsizey : (j == 0 ? 0.0 : (j + 1) * s->binsizey);
          double delz = (k > 0) ? (k - 1) * s->binsizez : (k == 0 ? 0.0 : (k + 1) * s->binsizez);
          if (delx * delx + dely * dely + delz * delz < cutneighmaxsq)
            s->stencil[s->nstencil++] = k * s->mbiny * s->mbinx + j * s->mbinx + i;
        }
}

int orc_coord2bin(orc_sys *s, const double *x)  /* src/nbin.cpp:139-171 */
{
  int ix, iy, iz;
  if (x[0] >= s->boxhi[0]) ix = (int) ((x[0] - s->boxhi[0]) * s->bininvx) + s->nbinx;
  else if (x[0] >= s->boxlo[0]) { ix = (int) ((x[0] - s->boxlo[0]) * s->bininvx); ix = MIN(ix, s->nbinx - 1); }
  else ix = (int) ((x[0] - s->boxlo[0]) * s->bininvx) - 1;
  if (x[1] >= s->boxhi[1]) iy = (int) ((x[1] - s->boxhi[1]) * s->bininvy) + s->nbiny;
  else if (x[1] >= s->boxlo[1]) { iy = (int) ((x[1] - s->boxlo[1]) * s->bininvy); iy = MIN(iy, s->nbiny - 1); }
  else iy = (int) ((x[1] - s->boxlo[1]) * s->bininvy) - 1;
  if (x[2] >= s->boxhi[2]) iz = (int) ((x[2] - s->boxhi[2]) * s->bininvz) + s->nbinz;
  else if (x[2] >= s->boxlo[2]) { iz = (int) ((x[2] - s->boxlo[2]) * s->bininvz); iz = MIN(iz, s->nbinz - 1); }
  else iz = (int) ((x[2] - s->boxlo[2]) * s->bininvz) - 1;
  return (iz - s->mbinzlo) * s->mbiny * s->mbinx + (iy - s->mbinylo) * s->mbinx + (ix - s->mbinxlo);
}

void orc_bin_atoms(orc_sys *s)  /* src/nbin_standard.cpp:221-261 */
{
  for (int i = 0; i < s->mbins; i++) s->binhead[i] = -1;
  int nall = s->nlocal + s->nghost;
  for (int i = nall - 1; i >= 0; i--) {
    int ibin = orc_coord2bin(s, &s->x[3 * i]);
    s->atom2bin[i] = ibin;
    s->bins[i] = s->binhead[ibin];
    s->binhead[ibin] = i;
  }
}
void orc_bin_info(orc_sys *s, int *out)
{
  out[0] = s->nbinx; out[1] = s->nbiny; out[2] = s->nbinz; out[3] = s->mbinx; out[4] = s->mbiny; out[5] = s->mbinz;
  out[6] = s->mbinxlo; out[7] = s->mbinylo; out[8] = s->mbinzlo; out[9] = s->nstencil;
}

/* =====================================================================================
 * NPairHalfBinAtomonlyNewton::build (src/npair_half_bin_atomonly_newton.cpp:33-113) and, when
 * special partners were given, NPairHalfBinNewton::build (src/npair_half_bin_newton.cpp:36-161)
 * with NPair::find_special (src/npair.h:125-158)
 * ===================================================================================== */
static int find_special(orc_sys *s, int i, int tagj)
{
  const int *list = &s->special[(size_t) i * s->maxspecial];
  const int *n = &s->nspecial[3 * i];
  int n1 = n[0], n2 = n[1], n3 = n[2];
  for (int k = 0; k < n3; k++) {
    if (list[k] == tagj) {
      int which = (k < n1) ? 1 : (k < n2 ? 2 : 3);
      if (s->special_flag[which] == 0) return -1;
      else if (s->special_flag[which] == 1) return 0;
      else return which;
    }
  }
  return 0;
}

static void push_neigh(orc_sys *s, int64_t *n, int j)
{
  if (*n == s->maxneighbors) {
    s->maxneighbors = s->maxneighbors ? 2 * s->maxneighbors : (int64_t) 1 << 20;
    s->neighbors = (int *) realloc(s->neighbors, sizeof(int) * s->maxneighbors);
  }
  s->neighbors[(*n)++] = j;
}

void orc_neigh_build(orc_sys *s)
{
  int nlocal = s->nlocal, ld = s->ntypes + 1;
  const double *x = s->x;
  int64_t ntot = 0;
  for (int i = 0; i < nlocal; i++) {
    s->firstneigh[i] = ntot;
    int itype = s->type[i];
    double xtmp = x[3 * i], ytmp = x[3 * i + 1], ztmp = x[3 * i + 2];
    for (int j = s->bins[i]; j >= 0; j = s->bins[j]) {
      if (j >= nlocal) {
        if (x[3 * j + 2] < ztmp) continue;
        if (x[3 * j + 2] == ztmp) {
          if (x[3 * j + 1] < ytmp) continue;
          if (x[3 * j + 1] == ytmp && x[3 * j] < xtmp) continue;
        }
      }
      int jtype = s->type[j];
      double delx = xtmp - x[3 * j], dely = ytmp - x[3 * j + 1], delz = ztmp - x[3 * j + 2];
      double rsq = delx * delx + dely * dely + delz * delz;
      if (rsq <= s->cutneighsq[itype * ld + jtype]) {
        if (s->have_special) {
          int which = find_special(s, i, s->tag[j]);
          if (which == 0) push_neigh(s, &ntot, j);
          else if (which > 0) push_neigh(s, &ntot, j ^ (which << SBBITS));
        } else push_neigh(s, &ntot, j);
      }
    }
    int ibin = s->atom2bin[i];
    for (int k = 0; k < s->nstencil; k++) {
      for (int j = s->binhead[ibin + s->stencil[k]]; j >= 0; j = s->bins[j]) {
        int jtype = s->type[j];
        double delx = xtmp - x[3 * j], dely = ytmp - x[3 * j + 1], delz = ztmp - x[3 * j + 2];
        double rsq = delx * delx + dely * dely + delz * delz;
        if (rsq <= s->cutneighsq[itype * ld + jtype]) {
          if (s->have_special) {
            int which = find_special(s, i, s->tag[j]);
            if (which == 0) push_neigh(s, &ntot, j);
            else if (which > 0) push_neigh(s, &ntot, j ^ (which << SBBITS));
          } else push_neigh(s, &ntot, j);
        }
      }
    }
    s->numneigh[i] = (int) (ntot - s->firstneigh[i]);
  }
  s->firstneigh[nlocal] = ntot;
}

/* Neighbor::build src/neighbor.cpp:2382-2464 : snapshot xhold, bin, build */
void orc_neighbor_build(orc_sys *s)
{
  s->ago = 0; s->ncalls++;
  memcpy(s->xhold, s->x, sizeof(double) * 3 * s->nlocal);
  orc_bin_atoms(s);
  orc_neigh_build(s);
  if (build_bondlist(s) > 0) { fprintf(stderr, "oracle: Bond atoms missing\n"); abort(); }  /* Neighbor::build_topology */
}

/* Neighbor::decide / check_distance src/neighbor.cpp:2293-2374 */
int orc_neighbor_decide(orc_sys *s)
{
  s->ago++;
  if (s->ago >= s->delay && s->ago % s->every == 0) {
    if (s->dist_check == 0) return 1;
    int flag = 0;
    for (int i = 0; i < s->nlocal; i++) {
      double delx = s->x[3 * i] - s->xhold[3 * i];
      double dely = s->x[3 * i + 1] - s->xhold[3 * i + 1];
      double delz = s->x[3 * i + 2] - s->xhold[3 * i + 2];
      double rsq = delx * delx + dely * dely + delz * delz;
      if (rsq > s->triggersq) flag = 1;
    }
    if (flag && s->ago == MAX(s->every, s->delay)) s->ndanger++;
    return flag;
  }
  return 0;
}

/* =====================================================================================
 * Pair::ev_tally (src/pair.cpp:1013-1108, global terms only) and virial_fdotr_compute (:1735-1751)
 * ===================================================================================== */
static void ev_tally(orc_sys *s, int i, int j, double evdwl, double fpair, double delx, double dely, double delz)
{
  (void) i; (void) j;
  if (s->eflag) s->eng_vdwl += evdwl;     /* newton_pair: full contribution */
}

static void virial_fdotr(orc_sys *s)
{
  int nall = s->nlocal + s->nghost;
  const double *x = s->x, *f = s->f;
  for (int i = 0; i < nall; i++) {
    s->virial[0] += f[3 * i] * x[3 * i];
    s->virial[1] += f[3 * i + 1] * x[3 * i + 1];
    s->virial[2] += f[3 * i + 2] * x[3 * i + 2];
    s->virial[3] += f[3 * i + 1] * x[3 * i];
    s->virial[4] += f[3 * i + 2] * x[3 * i];
    s->virial[5] += f[3 * i + 2] * x[3 * i + 1];
  }
}

/* PairLJCut::compute src/pair_lj_cut.cpp:69-139 (newton_pair on) */
static void pair_lj_compute(orc_sys *s)
{
  int ld = s->ntypes + 1;
  const double *x = s->x; double *f = s->f;
  for (int i = 0; i < s->nlocal; i++) {
    double xtmp = x[3 * i], ytmp = x[3 * i + 1], ztmp = x[3 * i + 2];
    int itype = s->type[i];
    const int *jlist = &s->neighbors[s->firstneigh[i]];
    int jnum = s->numneigh[i];
    for (int jj = 0; jj < jnum; jj++) {
      int j = jlist[jj];
      double factor_lj = s->special_lj[(j >> SBBITS) & 3];
      j &= NEIGHMASK;
      double delx = xtmp - x[3 * j], dely = ytmp - x[3 * j + 1], delz = ztmp - x[3 * j + 2];
      double rsq = delx * delx + dely * dely + delz * delz;
      int jtype = s->type[j];
      if (rsq < s->cutsq[itype * ld + jtype]) {
        double r2inv = 1.0 / rsq;
        double r6inv = r2inv * r2inv * r2inv;
        double forcelj = r6inv * (s->lj1[itype * ld + jtype] * r6inv - s->lj2[itype * ld + jtype]);
        double fpair = factor_lj * forcelj * r2inv;
        f[3 * i] += delx * fpair; f[3 * i + 1] += dely * fpair; f[3 * i + 2] += delz * fpair;
        f[3 * j] -= delx * fpair; f[3 * j + 1] -= dely * fpair; f[3 * j + 2] -= delz * fpair;
        double evdwl = 0.0;
        if (s->eflag) {
          evdwl = r6inv * (s->lj3[itype * ld + jtype] * r6inv - s->lj4[itype * ld + jtype]) - s->offset[itype * ld + jtype];
          evdwl *= factor_lj;
        }
        if (s->eflag || s->vflag) ev_tally(s, i, j, evdwl, fpair, delx, dely, delz);
      }
    }
  }
  if (s->vflag) virial_fdotr(s);
}

/* PairEAM::compute src/MANYBODY/pair_eam.cpp:142-328 (newton_pair on, scale = 1) */
static void pair_eam_compute(orc_sys *s)
{
  int ld = s->ntypes + 1, nlocal = s->nlocal, nall = s->nlocal + s->nghost;
  int nr = s->nr, nrho = s->nrho;
  const double *x = s->x; double *f = s->f, *rho = s->rho, *fp = s->fp;
  for (int i = 0; i < nall; i++) rho[i] = 0.0;
  for (int i = 0; i < nlocal; i++) {
    double xtmp = x[3 * i], ytmp = x[3 * i + 1], ztmp = x[3 * i + 2];
    int itype = s->type[i];
    const int *jlist = &s->neighbors[s->firstneigh[i]];
    for (int jj = 0; jj < s->numneigh[i]; jj++) {
      int j = jlist[jj] & NEIGHMASK;
      double delx = xtmp - x[3 * j], dely = ytmp - x[3 * j + 1], delz = ztmp - x[3 * j + 2];
      double rsq = delx * delx + dely * dely + delz * delz;
      if (rsq < s->cutforcesq) {
        int jtype = s->type[j];
        double p = sqrt(rsq) * s->rdr + 1.0;
        int m = (int) p;
        m = MIN(m, nr - 1);
        p -= m;
        p = MIN(p, 1.0);
        const double *coeff = &s->rhor_spline[((size_t) s->type2rhor[jtype * ld + itype] * (nr + 1) + m) * 7];
        rho[i] += ((coeff[3] * p + coeff[4]) * p + coeff[5]) * p + coeff[6];
        coeff = &s->rhor_spline[((size_t) s->type2rhor[itype * ld + jtype] * (nr + 1) + m) * 7];
        rho[j] += ((coeff[3] * p + coeff[4]) * p + coeff[5]) * p + coeff[6];
      }
    }
  }
  reverse_comm_scalar(s, rho);
  for (int i = 0; i < nlocal; i++) {
    double p = rho[i] * s->rdrho + 1.0;
    int m = (int) p;
    m = MAX(1, MIN(m, nrho - 1));
    p -= m;
    p = MIN(p, 1.0);
    const double *coeff = &s->frho_spline[((size_t) s->type2frho[s->type[i]] * (nrho + 1) + m) * 7];
    fp[i] = (coeff[0] * p + coeff[1]) * p + coeff[2];
    if (s->eflag) {
      double phi = ((coeff[3] * p + coeff[4]) * p + coeff[5]) * p + coeff[6];
      if (rho[i] > s->rhomax) phi += fp[i] * (rho[i] - s->rhomax);
      s->eng_vdwl += phi;
    }
  }
  forward_comm_scalar(s, fp);
  for (int i = 0; i < nlocal; i++) {
    double xtmp = x[3 * i], ytmp = x[3 * i + 1], ztmp = x[3 * i + 2];
    int itype = s->type[i];
    const int *jlist = &s->neighbors[s->firstneigh[i]];
    for (int jj = 0; jj < s->numneigh[i]; jj++) {
      int j = jlist[jj] & NEIGHMASK;
      double delx = xtmp - x[3 * j], dely = ytmp - x[3 * j + 1], delz = ztmp - x[3 * j + 2];
      double rsq = delx * delx + dely * dely + delz * delz;
      if (rsq < s->cutforcesq) {
        int jtype = s->type[j];
        double r = sqrt(rsq);
        double p = r * s->rdr + 1.0;
        int m = (int) p;
        m = MIN(m, nr - 1);
        p -= m;
        p = MIN(p, 1.0);
        const double *coeff = &s->rhor_spline[((size_t) s->type2rhor[itype * ld + jtype] * (nr + 1) + m) * 7];
        double rhoip = (coeff[0] * p + coeff[1]) * p + coeff[2];
        coeff = &s->rhor_spline[((size_t) s->type2rhor[jtype * ld + itype] * (nr + 1) + m) * 7];
        double rhojp = (coeff[0] * p + coeff[1]) * p + coeff[2];
        coeff = &s->z2r_spline[((size_t) s->type2z2r[itype * ld + jtype] * (nr + 1) + m) * 7];
        double z2p = (coeff[0] * p + coeff[1]) * p + coeff[2];
        double z2 = ((coeff[3] * p + coeff[4]) * p + coeff[5]) * p + coeff[6];
        double recip = 1.0 / r;
        double phi = z2 * recip;
        double phip = z2p * recip - phi * recip;
        double psip = fp[i] * rhojp + fp[j] * rhoip + phip;
        double fpair = -1.0 * psip * recip;
        f[3 * i] += delx * fpair; f[3 * i + 1] += dely * fpair; f[3 * i + 2] += delz * fpair;
        f[3 * j] -= delx * fpair; f[3 * j + 1] -= dely * fpair; f[3 * j + 2] -= delz * fpair;
        if (s->eflag) s->eng_vdwl += 1.0 * phi;
      }
    }
  }
  if (s->vflag) virial_fdotr(s);
}

/* Verlet::force_clear (src/verlet.cpp:376-421) + Pair::ev_init + compute + CommBrick::reverse_comm */
void orc_compute_forces(orc_sys *s, int eflag, int vflag)
{
  int nall = s->nlocal + s->nghost;
  memset(s->f, 0, sizeof(double) * 3 * nall);
  s->eflag = eflag; s->vflag = vflag;
  s->eng_vdwl = 0.0;
  for (int k = 0; k < 6; k++) s->virial[k] = 0.0;
  s->eng_bond = 0.0;
  for (int k = 0; k < 6; k++) s->virial_bond[k] = 0.0;
  if (s->pair_style == 1) pair_lj_compute(s);
  else if (s->pair_style == 2) pair_eam_compute(s);
  if (s->nbondtypes && bond_fene_compute(s) < 0) { fprintf(stderr, "oracle: Bad FENE bond\n"); abort(); }  /* src/verlet.cpp:327-332 */
  orc_reverse_comm(s);
}

/* =====================================================================================
 * in.chain additions.  Atom::sort (src/atom.cpp:2246-2345, setup_sort_bins :2351-2405):
 * stable counting sort of owned atoms by sort bin (binsize = 0.5*cutneighmax), x fastest.
 * The reference does this in Verlet::setup; it fixes the local atom order, which matters
 * here because fix langevin draws its random numbers in local-index order.
 * ===================================================================================== */
void orc_enable_sort(orc_sys *s, int flag) { s->sort_enabled = flag; }

static void permute_int(int *a, const int *perm, int n, int w)
{
  int *t = (int *) xmalloc(sizeof(int) * (size_t) n * w);
  for (int i = 0; i < n; i++) memcpy(t + (size_t) i * w, a + (size_t) perm[i] * w, sizeof(int) * w);
  memcpy(a, t, sizeof(int) * (size_t) n * w);
  free(t);
}
static void permute_dbl(double *a, const int *perm, int n, int w)
{
  double *t = (double *) xmalloc(sizeof(double) * (size_t) n * w);
  for (int i = 0; i < n; i++) memcpy(t + (size_t) i * w, a + (size_t) perm[i] * w, sizeof(double) * w);
  memcpy(a, t, sizeof(double) * (size_t) n * w);
  free(t);
}

void orc_sort_atoms(orc_sys *s)
{
  double binsize = 0.5 * s->cutneighmax;
  if (binsize <= 0.0) return;
  double bininv = 1.0 / binsize;
  int nb[3]; double inv[3];
  for (int d = 0; d < 3; d++) {
    nb[d] = (int) ((s->boxhi[d] - s->boxlo[d]) * bininv);
    if (nb[d] == 0) nb[d] = 1;
    inv[d] = nb[d] / (s->boxhi[d] - s->boxlo[d]);
  }
  int nbins = nb[0] * nb[1] * nb[2];
  if (nbins == 1) return;
  int n = s->nlocal;
  int *binhead = (int *) xmalloc(sizeof(int) * nbins), *next = (int *) xmalloc(sizeof(int) * n);
  int *perm = (int *) xmalloc(sizeof(int) * n);
  for (int b = 0; b < nbins; b++) binhead[b] = -1;
  for (int i = n - 1; i >= 0; i--) {
    int ix = (int) ((s->x[3 * i] - s->boxlo[0]) * inv[0]);
    int iy = (int) ((s->x[3 * i + 1] - s->boxlo[1]) * inv[1]);
    int iz = (int) ((s->x[3 * i + 2] - s->boxlo[2]) * inv[2]);
    ix = MIN(MAX(ix, 0), nb[0] - 1); iy = MIN(MAX(iy, 0), nb[1] - 1); iz = MIN(MAX(iz, 0), nb[2] - 1);
    int ibin = iz * nb[1] * nb[0] + iy * nb[0] + ix;
    next[i] = binhead[ibin];
    binhead[ibin] = i;
  }
  int m = 0;
  for (int b = 0; b < nbins; b++)
    for (int i = binhead[b]; i >= 0; i = next[i]) perm[m++] = i;   /* perm[I] = J: new atom I is old atom J */
  permute_dbl(s->x, perm, n, 3); permute_dbl(s->v, perm, n, 3);
  permute_int(s->type, perm, n, 1); permute_int(s->tag, perm, n, 1); permute_int(s->image, perm, n, 1);
  if (s->have_special) { permute_int(s->nspecial, perm, n, 3); permute_int(s->special, perm, n, s->maxspecial); }
  if (s->bond_per_atom) {
    permute_int(s->num_bond, perm, n, 1);
    permute_int(s->bond_type, perm, n, s->bond_per_atom); permute_int(s->bond_atom, perm, n, s->bond_per_atom);
  }
  free(binhead); free(next); free(perm);
}

/* bonds as Atom::data_bonds stores them with newton_bond on (src/atom.cpp: each bond lives on its first atom):
 * num_bond[i], bond_type[i][m], bond_atom[i][m] = partner TAG */
void orc_set_bonds(orc_sys *s, int bond_per_atom, const int *num_bond, const int *bond_type, const int *bond_atom)
{
  int n = s->nlocal;
  s->bond_per_atom = bond_per_atom;
  free(s->num_bond); free(s->bond_type); free(s->bond_atom);
  s->num_bond = (int *) xmalloc(sizeof(int) * n); memcpy(s->num_bond, num_bond, sizeof(int) * n);
  s->bond_type = (int *) xmalloc(sizeof(int) * (size_t) n * bond_per_atom);
  s->bond_atom = (int *) xmalloc(sizeof(int) * (size_t) n * bond_per_atom);
  memcpy(s->bond_type, bond_type, sizeof(int) * (size_t) n * bond_per_atom);
  memcpy(s->bond_atom, bond_atom, sizeof(int) * (size_t) n * bond_per_atom);
}

/* bond_style fene + bond_coeff (src/MOLECULE/bond_fene.cpp:142-165); arrays are 1-based [nbondtypes+1] */
void orc_bond_fene(orc_sys *s, int nbondtypes, const double *k, const double *r0, const double *eps, const double *sigma)
{
  s->nbondtypes = nbondtypes;
  size_t b = sizeof(double) * (nbondtypes + 1);
  free(s->fene_k); free(s->fene_r0); free(s->fene_eps); free(s->fene_sigma);
  s->fene_k = (double *) xmalloc(b); s->fene_r0 = (double *) xmalloc(b); s->fene_eps = (double *) xmalloc(b); s->fene_sigma = (double *) xmalloc(b);
  memcpy(s->fene_k, k, b); memcpy(s->fene_r0, r0, b); memcpy(s->fene_eps, eps, b); memcpy(s->fene_sigma, sigma, b);
}

/* Atom::map_set, array style (src/atom_map.cpp:167-199): map[tag] = lowest local index carrying the tag (the
 * owned atom), sametag[i] = next higher index with the same tag (its ghost images) */
static void map_set(orc_sys *s)
{
  int nall = s->nlocal + s->nghost, maxtag = 0;
  for (int i = 0; i < nall; i++) maxtag = MAX(maxtag, s->tag[i]);
  if (maxtag > s->maxtag || !s->tag2local) {
    s->maxtag = maxtag;
    s->tag2local = (int *) realloc(s->tag2local, sizeof(int) * ((size_t) maxtag + 1));
  }
  for (int t = 0; t <= s->maxtag; t++) s->tag2local[t] = -1;
  s->sametag = (int *) realloc(s->sametag, sizeof(int) * (size_t) (nall ? nall : 1));
  for (int i = nall - 1; i >= 0; i--) {
    s->sametag[i] = s->tag2local[s->tag[i]];
    s->tag2local[s->tag[i]] = i;
  }
}

/* Domain::closest_image (src/domain.cpp:1054-1082) */
static int closest_image(orc_sys *s, int i, int j)
{
  if (j < 0) return j;
  const double *x = s->x, *xi = &s->x[3 * i];
  int closest = j;
  double delx = xi[0] - x[3 * j], dely = xi[1] - x[3 * j + 1], delz = xi[2] - x[3 * j + 2];
  double rsqmin = delx * delx + dely * dely + delz * delz;
  while (s->sametag[j] >= 0) {
    j = s->sametag[j];
    delx = xi[0] - x[3 * j]; dely = xi[1] - x[3 * j + 1]; delz = xi[2] - x[3 * j + 2];
    double rsq = delx * delx + dely * dely + delz * delz;
    if (rsq < rsqmin) { rsqmin = rsq; closest = j; }
  }
  return closest;
}

/* NTopoBondAll::build (src/ntopo_bond_all.cpp:37-76), newton_bond on; returns the number of missing partners */
static int build_bondlist(orc_sys *s)
{
  int nmissing = 0;
  s->nbondlist = 0;
  if (!s->bond_per_atom) return 0;
  map_set(s);
  for (int i = 0; i < s->nlocal; i++)
    for (int m = 0; m < s->num_bond[i]; m++) {
      int t = s->bond_atom[i * s->bond_per_atom + m];
      int atom1 = (t >= 0 && t <= s->maxtag) ? s->tag2local[t] : -1;
      if (atom1 == -1) { nmissing++; continue; }
      atom1 = closest_image(s, i, atom1);
      if (s->nbondlist == s->maxbondlist) {
        s->maxbondlist = s->maxbondlist ? 2 * s->maxbondlist : 16384;
        s->bondlist = (int *) realloc(s->bondlist, sizeof(int) * 3 * (size_t) s->maxbondlist);
      }
      int *b = &s->bondlist[3 * s->nbondlist++];
      b[0] = i; b[1] = atom1; b[2] = s->bond_type[i * s->bond_per_atom + m];
    }
  return nmissing;
}

/* BondFENE::compute (src/MOLECULE/bond_fene.cpp:52-127) with Bond::ev_tally (src/bond.cpp:220-282), newton_bond on.
 * Returns -1 on "Bad FENE bond". */
static int bond_fene_compute(orc_sys *s)
{
  const double CUBEROOT2 = 1.25992104989487316476; /* MY_CUBEROOT2, src/math_const.h */
  const double *x = s->x; double *f = s->f;
  for (int n = 0; n < s->nbondlist; n++) {
    int i1 = s->bondlist[3 * n], i2 = s->bondlist[3 * n + 1], type = s->bondlist[3 * n + 2];
    double delx = x[3 * i1] - x[3 * i2], dely = x[3 * i1 + 1] - x[3 * i2 + 1], delz = x[3 * i1 + 2] - x[3 * i2 + 2];
    double rsq = delx * delx + dely * dely + delz * delz;
    double r0sq = s->fene_r0[type] * s->fene_r0[type];
    double rlogarg = 1.0 - rsq / r0sq;
    if (rlogarg < 0.1) {
      if (rlogarg <= -3.0) return -1;
      rlogarg = 0.1;
    }
    double fbond = -s->fene_k[type] / rlogarg, sr6 = 0.0, ebond = 0.0;
    if (rsq < CUBEROOT2 * s->fene_sigma[type] * s->fene_sigma[type]) {
      double sr2 = s->fene_sigma[type] * s->fene_sigma[type] / rsq;
      sr6 = sr2 * sr2 * sr2;
      fbond += 48.0 * s->fene_eps[type] * sr6 * (sr6 - 0.5) / rsq;
    }
    if (s->eflag) {
      ebond = -0.5 * s->fene_k[type] * r0sq * log(rlogarg);
      if (rsq < CUBEROOT2 * s->fene_sigma[type] * s->fene_sigma[type])
        ebond += 4.0 * s->fene_eps[type] * sr6 * (sr6 - 1.0) + s->fene_eps[type];
    }
    f[3 * i1] += delx * fbond; f[3 * i1 + 1] += dely * fbond; f[3 * i1 + 2] += delz * fbond;
    f[3 * i2] -= delx * fbond; f[3 * i2 + 1] -= dely * fbond; f[3 * i2 + 2] -= delz * fbond;
    if (s->eflag) s->eng_bond += ebond;
    if (s->vflag) {
      s->virial_bond[0] += delx * delx * fbond; s->virial_bond[1] += dely * dely * fbond; s->virial_bond[2] += delz * delz * fbond;
      s->virial_bond[3] += delx * dely * fbond; s->virial_bond[4] += delx * delz * fbond; s->virial_bond[5] += dely * delz * fbond;
    }
  }
  return 0;
}

/* RanMars (src/random_mars.cpp:30-100): constructor and uniform() */
static double mars_uniform(orc_sys *s)
{
  double uni = s->mars_u[s->mars_i97] - s->mars_u[s->mars_j97];
  if (uni < 0.0) uni += 1.0;
  s->mars_u[s->mars_i97] = uni;
  s->mars_i97--; if (s->mars_i97 == 0) s->mars_i97 = 97;
  s->mars_j97--; if (s->mars_j97 == 0) s->mars_j97 = 97;
  s->mars_c -= s->mars_cd;
  if (s->mars_c < 0.0) s->mars_c += s->mars_cm;
  uni -= s->mars_c;
  if (uni < 0.0) uni += 1.0;
  return uni;
}
static void mars_init(orc_sys *s, int seed)
{
  memset(s->mars_u, 0, sizeof s->mars_u);
  int ij = (seed - 1) / 30082, kl = (seed - 1) - 30082 * ij;
  int i = (ij / 177) % 177 + 2, j = ij % 177 + 2, k = (kl / 169) % 178 + 1, l = kl % 169;
  for (int ii = 1; ii <= 97; ii++) {
    double sum = 0.0, t = 0.5;
    for (int jj = 1; jj <= 24; jj++) {
      int m = ((i * j) % 179) * k % 179;
      i = j; j = k; k = m;
      l = (53 * l + 1) % 169;
      if ((l * m) % 64 >= 32) sum = sum + t;
      t = 0.5 * t;
    }
    s->mars_u[ii] = sum;
  }
  s->mars_c = 362436.0 / 16777216.0; s->mars_cd = 7654321.0 / 16777216.0; s->mars_cm = 16777213.0 / 16777216.0;
  s->mars_i97 = 97; s->mars_j97 = 33;
  mars_uniform(s);
}

/* the first n numbers of RanMars(seed)->uniform(), for checking another restatement of the generator */
void orc_ranmars(int seed, int n, double *out)
{
  orc_sys *s = (orc_sys *) calloc(1, sizeof(orc_sys));
  mars_init(s, seed);
  for (int i = 0; i < n; i++) out[i] = mars_uniform(s);
  free(s);
}

/* fix ID all langevin Tstart Tstop damp seed (src/fix_langevin.cpp:55-190): RanMars(seed + me), me = 0 */
void orc_fix_langevin(orc_sys *s, double t_start, double t_stop, double t_period, int seed)
{
  s->langevin = 1; s->lang_tstart = t_start; s->lang_tstop = t_stop; s->lang_period = t_period;
  mars_init(s, seed);
}

/* FixLangevin::post_force, the plain template instance (no gjf/tally/bias/rmass/zero; src/fix_langevin.cpp:583-716,
 * gfactor1/2 :284-298, compute_target :773-787) */
static void langevin_post_force(orc_sys *s, int64_t ntimestep)
{
  double delta = (double) (ntimestep - s->beginstep);
  if (delta != 0.0) delta /= (double) (s->endstep - s->beginstep);
  double t_target = s->lang_tstart + delta * (s->lang_tstop - s->lang_tstart);
  double tsqrt = sqrt(t_target);
  for (int i = 0; i < s->nlocal; i++) {
    double m = s->mass[s->type[i]];
    double gamma1 = -m / s->lang_period / s->ftm2v;
    double gamma2 = sqrt(m) / s->ftm2v;
    gamma2 *= sqrt(24.0 * s->boltz / s->lang_period / s->dt / s->mvv2e);
    gamma1 *= 1.0 / 1.0;            /* ratio[type] = 1 */
    gamma2 *= 1.0 / sqrt(1.0);
    gamma2 = gamma2 * tsqrt;
    double fran0 = gamma2 * (mars_uniform(s) - 0.5);
    double fran1 = gamma2 * (mars_uniform(s) - 0.5);
    double fran2 = gamma2 * (mars_uniform(s) - 0.5);
    double fdrag0 = gamma1 * s->v[3 * i], fdrag1 = gamma1 * s->v[3 * i + 1], fdrag2 = gamma1 * s->v[3 * i + 2];
    s->f[3 * i] += fdrag0 + fran0; s->f[3 * i + 1] += fdrag1 + fran1; s->f[3 * i + 2] += fdrag2 + fran2;
  }
}

/* =====================================================================================
 * FixNVE (src/fix_nve.cpp:51-141)
 * ===================================================================================== */
void orc_nve_initial(orc_sys *s)
{
  double dtv = s->dt, dtf = 0.5 * s->dt * s->ftm2v;
  for (int i = 0; i < s->nlocal; i++) {
    double dtfm = dtf / s->mass[s->type[i]];
    s->v[3 * i] += dtfm * s->f[3 * i]; s->v[3 * i + 1] += dtfm * s->f[3 * i + 1]; s->v[3 * i + 2] += dtfm * s->f[3 * i + 2];
    s->x[3 * i] += dtv * s->v[3 * i]; s->x[3 * i + 1] += dtv * s->v[3 * i + 1]; s->x[3 * i + 2] += dtv * s->v[3 * i + 2];
  }
}
void orc_nve_final(orc_sys *s)
{
  double dtf = 0.5 * s->dt * s->ftm2v;
  for (int i = 0; i < s->nlocal; i++) {
    double dtfm = dtf / s->mass[s->type[i]];
    s->v[3 * i] += dtfm * s->f[3 * i]; s->v[3 * i + 1] += dtfm * s->f[3 * i + 1]; s->v[3 * i + 2] += dtfm * s->f[3 * i + 2];
  }
}

/* =====================================================================================
 * Verlet::setup (src/verlet.cpp:93-162) and Verlet::run (src/verlet.cpp:229-360)
 * ===================================================================================== */
void orc_setup(orc_sys *s)
{
  s->triggersq = 0.25 * s->skin * s->skin;      /* src/neighbor.cpp:344 */
  orc_pbc(s);
  /* cutghost needed by setup_bins */
  s->cutghost = s->cutneighmax;
  orc_setup_bins(s);
  if (s->sort_enabled) orc_sort_atoms(s);       /* Verlet::setup: if (atom->sortfreq > 0) atom->sort() (src/verlet.cpp:121) */
  orc_borders(s);
  orc_neighbor_build(s);
  s->ncalls = 0; s->ndanger = 0;                /* build at setup is not counted (src/neighbor.cpp:2382 called with 1 -> ncalls reset in init) */
  orc_compute_forces(s, 1, 1);
  if (s->langevin) langevin_post_force(s, s->beginstep);  /* modify->setup -> FixLangevin::setup -> post_force (src/fix_langevin.cpp:359) */
}

/* eflag/vflag: energy+virial only on steps in thermo_every multiples and the last step
 * (Integrate::ev_set src/integrate.cpp:138-183) */
void orc_run(orc_sys *s, int nsteps, int thermo_every, int step0)
{
  s->beginstep = step0; s->endstep = (int64_t) step0 + nsteps;   /* Run::command (src/run.cpp:172-176) */
  for (int istep = 1; istep <= nsteps; istep++) {
    int ntimestep = step0 + istep;
    int ev = (istep == nsteps) || (thermo_every > 0 && ntimestep % thermo_every == 0);
    orc_nve_initial(s);
    int nflag = orc_neighbor_decide(s);
    if (nflag == 0) orc_forward_comm(s);
    else {
      orc_pbc(s);
      orc_borders(s);
      orc_neighbor_build(s);
    }
    orc_compute_forces(s, ev, ev);
    if (s->langevin) langevin_post_force(s, ntimestep);  /* modify->post_force, before final_integrate (src/verlet.cpp:345-347) */
    orc_nve_final(s);
  }
}

/* thermo: ComputeTemp src/compute_temp.cpp:55-100, ComputePressure src/compute_pressure.cpp:228-260,
 * KE = 0.5*dof*boltz*T (src/thermo.cpp:1712-1717); out = pe, ke, temp, press, virial[6] (pressure tensor pieces) */
void orc_thermo(orc_sys *s, double *out)
{
  double t = 0.0;
  for (int i = 0; i < s->nlocal; i++)
    t += (s->v[3 * i] * s->v[3 * i] + s->v[3 * i + 1] * s->v[3 * i + 1] + s->v[3 * i + 2] * s->v[3 * i + 2]) * s->mass[s->type[i]];
  double dof = 3.0 * s->nlocal - 3.0;
  double tfactor = s->mvv2e / (dof * s->boltz);
  double temp = t * tfactor;
  double inv_volume = 1.0 / (s->prd[0] * s->prd[1] * s->prd[2]);
  /* ComputePressure::virial_compute sums pair and bond virials (src/compute_pressure.cpp:314-341) */
  double press = (dof * s->boltz * temp + (s->virial[0] + s->virial_bond[0]) + (s->virial[1] + s->virial_bond[1]) +
                  (s->virial[2] + s->virial_bond[2])) / 3.0 * inv_volume * s->nktv2p;
  out[0] = s->eng_vdwl;
  out[1] = 0.5 * dof * s->boltz * temp;
  out[2] = temp;
  out[3] = press;
  for (int k = 0; k < 6; k++) out[4 + k] = s->virial[k];
}

/* bond energy (thermo E_mol) and bond virial */
void orc_thermo_bond(orc_sys *s, double *out)
{
  out[0] = s->eng_bond;
  for (int k = 0; k < 6; k++) out[1 + k] = s->virial_bond[k];
}
int orc_nbondlist(orc_sys *s) { return s->nbondlist; }

/* ---- getters ---- */
int orc_nlocal(orc_sys *s) { return s->nlocal; }
int orc_nghost(orc_sys *s) { return s->nghost; }
int orc_ncalls(orc_sys *s) { return s->ncalls; }
int orc_ndanger(orc_sys *s) { return s->ndanger; }
int64_t orc_npairs(orc_sys *s) { return s->firstneigh[s->nlocal]; }
void orc_get_atoms(orc_sys *s, double *x, double *v, double *f, int *type, int *tag, int *image)
{
  int n = s->nlocal;
  if (x) memcpy(x, s->x, sizeof(double) * 3 * n);
  if (v) memcpy(v, s->v, sizeof(double) * 3 * n);
  if (f) memcpy(f, s->f, sizeof(double) * 3 * n);
  if (type) memcpy(type, s->type, sizeof(int) * n);
  if (tag) memcpy(tag, s->tag, sizeof(int) * n);
  if (image) memcpy(image, s->image, sizeof(int) * n);
}
void orc_get_ghosts(orc_sys *s, double *x, int *tag)
{
  memcpy(x, &s->x[3 * s->nlocal], sizeof(double) * 3 * s->nghost);
  memcpy(tag, &s->tag[s->nlocal], sizeof(int) * s->nghost);
}
/* stored pairs as (tag_i, tag_j, code) triples, code as in oracle/ref_extract.cpp */
void orc_get_pairs(orc_sys *s, int *out)
{
  int maxtag = 0;
  for (int i = 0; i < s->nlocal; i++) maxtag = MAX(maxtag, s->tag[i]);
  int *owner = (int *) xmalloc(sizeof(int) * (maxtag + 1));
  for (int i = 0; i < s->nlocal; i++) owner[s->tag[i]] = i;
  int64_t k = 0;
  for (int i = 0; i < s->nlocal; i++)
    for (int jj = 0; jj < s->numneigh[i]; jj++) {
      int raw = s->neighbors[s->firstneigh[i] + jj];
      int j = raw & NEIGHMASK, special = (raw >> SBBITS) & 3;
      int o = owner[s->tag[j]];
      int code = special << 6;
      for (int d = 0; d < 3; d++) {
        int sh = (int) lround((s->x[3 * j + d] - s->x[3 * o + d]) / s->prd[d]);
        code |= (sh + 1) << (2 * d);
      }
      out[3 * k] = s->tag[i]; out[3 * k + 1] = s->tag[j]; out[3 * k + 2] = code;
      k++;
    }
  free(owner);
}
