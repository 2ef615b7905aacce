#!/bin/sh
# TEST INFRASTRUCTURE (oracle build) -- writes the include lists the reference's
# sources expect (style_*.h, packages_*.h, lmpinstalledpkgs.h, lmpgitversion.h) into
# oracle/_ref/gen/.  Restates what /root/reference/src/Make.sh:15-40,84-106 and
# src/Makefile:332-364 do with grep/echo; nothing is written into the reference tree.
#
# usage: gen_style_headers.sh <ref-src-dir> <out-dir> PKG/file PKG/file ...
set -e
LC_ALL=C; export LC_ALL
SRC=$1; GEN=$2; shift 2
PKGHDRS=""
for f in "$@"; do
  [ -f "$SRC/$f.h" ] && PKGHDRS="$PKGHDRS $SRC/$f.h"
done
# OPENMP headers that carry no style macro but no .cpp either are found via -I

style() {  # $1 = macro keyword, $2 = file prefix, $3 = style name
  out="$GEN/style_$3.h"
  : > "$out.tmp"
  for h in "$SRC"/$2*.h; do
    [ -f "$h" ] || continue
    if grep -qs "$1" "$h"; then echo "#include \"$(basename "$h")\"" >> "$out.tmp"; fi
  done
  for h in $PKGHDRS; do
    b=$(basename "$h")
    case "$b" in
      $2*) if grep -qs "$1" "$h"; then echo "#include \"$b\"" >> "$out.tmp"; fi ;;
    esac
  done
  if [ -f "$out" ] && cmp -s "$out" "$out.tmp"; then rm -f "$out.tmp"; else mv "$out.tmp" "$out"; fi
  [ -f "$GEN/packages_$3.h" ] || : > "$GEN/packages_$3.h"
}

style ANGLE_CLASS        angle_         angle
style ATOM_CLASS         atom_vec_      atom
style BODY_CLASS         body_          body
style BOND_CLASS         bond_          bond
style COMMAND_CLASS      ""             command
style COMPUTE_CLASS      compute_       compute
style DIHEDRAL_CLASS     dihedral_      dihedral
style DUMP_CLASS         dump_          dump
style FIX_CLASS          fix_           fix
style GRAN_SUB_MOD_CLASS gran_sub_mod_  gran_sub_mod
style IMPROPER_CLASS     improper_      improper
style INTEGRATE_CLASS    ""             integrate
style KSPACE_CLASS       ""             kspace
style MINIMIZE_CLASS     min_           minimize
style NBIN_CLASS         nbin_          nbin
style NPAIR_CLASS        npair_         npair
style NSTENCIL_CLASS     nstencil_      nstencil
style NTOPO_CLASS        ntopo_         ntopo
style PAIR_CLASS         pair_          pair
style READER_CLASS       reader_        reader
style REGION_CLASS       region_        region

cat > "$GEN/lmpinstalledpkgs.h" <<EOF
#ifndef LMP_INSTALLED_PKGS_H
#define LMP_INSTALLED_PKGS_H
const char * LAMMPS_NS::LAMMPS::installed_packages[] = {
"MANYBODY", "MOLECULE", "OPENMP", "OPT",
 NULL };
#endif
EOF
cat > "$GEN/lmpgitversion.h" <<EOF
#ifndef LMP_GIT_VERSION_H
#define LMP_GIT_VERSION_H
bool LAMMPS_NS::LAMMPS::has_git_info() { return false; }
const char *LAMMPS_NS::LAMMPS::git_commit() { return "(unknown)"; }
const char *LAMMPS_NS::LAMMPS::git_branch() { return "(unknown)"; }
const char *LAMMPS_NS::LAMMPS::git_descriptor() { return "(unknown)"; }
#endif
EOF
