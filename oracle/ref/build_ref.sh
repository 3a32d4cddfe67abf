#!/bin/bash
# build_ref.sh -- builds the UNMODIFIED reference LOWESS path (Fortran 2008) with whatever
# Fortran compiler the machine has, from the sources where they lie under $REF (default
# /root/reference), plus oracle/ref/ref_driver.f90; outputs only into oracle/_ref/.
# Then runs the driver on the inputs tests/golden/make_ref_inputs.py writes and leaves the
# dumps in oracle/_ref/ for tests/test_oracle.py::test_oracle_matches_reference_dumps.
#
# STATUS: never executed -- neither the development container nor the B200 boxes have a
# Fortran compiler (`which gfortran flang flang-new nvfortran ifx ifort f2c lfortran` prints
# nothing on both; probed in rounds 1 and 2).  The script exits 3
# in that case and the dump test skips: parity stays "unpinned".
# Flags = the reference's Release flags (CMakeLists.txt:114).
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/../_ref
FC=${FC:-}
for c in gfortran flang-new flang nvfortran ifx ifort lfortran; do
  if [ -z "$FC" ] && command -v $c >/dev/null 2>&1; then FC=$c; fi
done
if [ -z "$FC" ]; then echo "build_ref.sh: no Fortran compiler on this machine" >&2; exit 3; fi
if [ ! -d "$REF/src" ]; then echo "build_ref.sh: $REF/src not found" >&2; exit 4; fi
mkdir -p "$OUT/obj" && cd "$OUT/obj"
FFLAGS="-O3 -funroll-all-loops -fno-f2c -Wl,--allow-multiple-definition"
NL=$REF/src/external/nonlin/src
LA=$NL/external/linalg/src
FE=$LA/external/ferror/src
LP=$LA/external/lapack
# modules first, their submodules after them (dependency order read off the `use` statements)
SRC="$FE/ferror.f90 $LA/linalg_constants.f90 $LA/linalg_core.f90 $LA/linalg_basic.f90 $LA/linalg_factor.f90
     $LA/linalg_solve.f90 $LA/linalg_eigen.f90 $LA/linalg_sorting.f90 $LA/linalg_immutable.f90
     $NL/nonlin_constants.f90 $NL/nonlin_core.f90 $NL/nonlin_vecfcn_helper.f90 $NL/nonlin_fcn1var_helper.f90
     $NL/nonlin_fcnnvar_helper.f90 $NL/nonlin_equation_solver.f90 $NL/nonlin_equation_solver_1var.f90
     $NL/nonlin_equation_optimizer.f90 $NL/nonlin_least_squares.f90 $NL/nonlin_linesearch.f90
     $NL/nonlin_polynomials.f90 $NL/nonlin_solve.f90 $NL/nonlin_solve_brent.f90 $NL/nonlin_solve_line_search.f90
     $NL/nonlin_solve_newton.f90 $NL/nonlin_solve_newton1var.f90 $NL/nonlin_solve_quasi_newton.f90
     $NL/nonlin_optimize.f90 $NL/nonlin_optimize_bfgs.f90 $NL/nonlin_optimize_line_search.f90
     $NL/nonlin_optimize_nelder_mead.f90
     $REF/src/curvefit_core.f90 $REF/src/curvefit_statistics.f90 $REF/src/curvefit_regression.f90"
for f in $SRC; do $FC $FFLAGS -c "$f"; done
# LAPACK / BLAS: the system's if present, else the vendored Reference-LAPACK sources (double
# precision routines and their auxiliaries; DLASRT is the one the LOWESS path itself calls)
if echo 'end' | $FC -x f95 - -llapack -lblas -o /dev/null 2>/dev/null; then
  LIBS="-llapack -lblas"
else
  for f in $LP/SRC/d*.f $LP/SRC/i*.f $LP/SRC/xerbla.f $LP/SRC/lsame.f $LP/BLAS/SRC/d*.f $LP/BLAS/SRC/i*.f \
           $LP/INSTALL/dlamch.f $LP/INSTALL/lsame.f; do
    [ -f "$f" ] && $FC -O2 -c "$f" -o "lp_$(basename "$f" .f).o" || true
  done
  ar rcs liblapack_ref.a lp_*.o
  LIBS="liblapack_ref.a"
fi
$FC $FFLAGS -c "$HERE/ref_driver.f90"
OBJ=$(ls *.o | grep -v '^lp_')
$FC $FFLAGS -o "$OUT/ref_driver" $OBJ $LIBS
echo "built $OUT/ref_driver with $FC"
# dumps for the committed inputs
python "$HERE/../../tests/golden/make_ref_inputs.py" "$OUT"
for in in "$OUT"/in_*.bin; do
  "$OUT/ref_driver" "$in" "${in/in_/out_}"
done
ls -la "$OUT"
