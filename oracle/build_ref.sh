#!/bin/bash
# TEST INFRASTRUCTURE ONLY -- builds the reference's own CPU tile kernels into oracle/_ref/.
#
# Compiles, from the sources where they lie under /root/reference (nothing is copied
# into the repo), the reference's CPU tile kernels and matrix generators that sit on the
# tiled-Cholesky path:
#   src/cores/core_z{potrf,trsm,herk,syrk,gemm,plgsy,plghe,plrnt,trtri,lauum,trmm}.c  (d and z precisions,
#   expanded by the reference's own tools/PrecisionGenerator exactly like
#   cmake_modules/PrecisionGenerator.cmake does: one file + one precision per call)
# and links them with the image's bundled OpenBLAS 0.3.15 (CBLAS + LAPACKE, LP64).
# The reference's full path (PaRSEC JDF runtime) cannot be built here: the parsec/
# submodule is absent.  Output: oracle/_ref/libdplasma_ref.so = those
# reference kernels + oracle/ref_driver.c (an OpenMP-task restatement of the DAG of
# src/zpotrf_L.jdf that calls them).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
R="${DPLASMA_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
PY="${PYTHON:-python3}"
if [ ! -d "$R/src/cores" ]; then
  echo "build_ref.sh: $R not present; keeping prebuilt oracle/_ref (if any)"; exit 0
fi
OB_DIR="$($PY - <<'PYEOF'
import glob, os, sys, sysconfig
sp = sysconfig.get_paths()["purelib"]
c = glob.glob(os.path.join(sp, "opencv_python_headless.libs", "libopenblasp-*.so"))
print(os.path.dirname(c[0]) if c else "")
PYEOF
)"
if [ -z "$OB_DIR" ]; then echo "build_ref.sh: bundled OpenBLAS not found"; exit 0; fi
OB_LIB="$(ls "$OB_DIR"/libopenblasp-*.so | head -1)"
mkdir -p "$OUT/gen" "$OUT/inc/dplasma" "$OUT/inc/cores"
# the two headers CMake would have configured (src/include/dplasma/config.h.in, CMakeLists.txt:77-82)
cat > "$OUT/inc/dplasma/config.h" <<'H'
#ifndef DPLASMA_CONFIG_H_HAS_BEEN_INCLUDED
#define DPLASMA_CONFIG_H_HAS_BEEN_INCLUDED
#define DPLASMA_HAVE_COMPLEX_H
#endif
H
cat > "$OUT/inc/dplasma_f2c.h" <<'H'
#ifndef DPLASMA_F2C_HEADER_INCLUDED
#define DPLASMA_F2C_HEADER_INCLUDED
#define DPLASMA_F2C_GLOBAL(name,NAME) name##_
#endif
H
gen() { # file precision outdir
  (cd "$OUT" && $PY "$R/tools/PrecisionGenerator/PrecisionGenerator.py" -f "$1" -p "$2" -P "$3/" >/dev/null)
}
for p in s d c z; do gen "$R/src/cores/core_zblas.h" $p "$OUT/inc/cores"; done
gen "$R/src/cores/core_zcblas.h" ds "$OUT/inc/cores" || true
gen "$R/src/cores/core_zcblas.h" zc "$OUT/inc/cores" || true
OBJS=()
CF="-O2 -fPIC -w -I$OUT/inc -I$OUT/inc/cores -I$R/src -I$R/src/include -I$R/src/cores"
for f in potrf trsm herk syrk gemm plgsy plghe plrnt trtri lauum trmm; do
  for p in d z; do
    gen "$R/src/cores/core_z$f.c" $p "$OUT/gen"
  done
done
for c in "$OUT"/gen/core_*.c; do
  gcc $CF -c "$c" -o "${c%.c}.o"; OBJS+=("${c%.c}.o")
done
gcc $CF -c "$R/src/cores/global.c" -o "$OUT/gen/global.o"; OBJS+=("$OUT/gen/global.o")
gcc -O2 -fopenmp $CF -c "$HERE/ref_driver.c" -o "$OUT/gen/ref_driver.o"; OBJS+=("$OUT/gen/ref_driver.o")
gcc -shared -fopenmp -o "$OUT/libdplasma_ref.so" "${OBJS[@]}" "$OB_LIB" -Wl,--disable-new-dtags -Wl,-rpath,"$OB_DIR" -lm
echo "build_ref.sh: built $OUT/libdplasma_ref.so (OpenBLAS: $OB_LIB)"
