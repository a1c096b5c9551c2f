#!/bin/bash
# Builds the UNMODIFIED reference into oracle/_ref/ (git-ignored, travels to the GPU box).
# TEST INFRASTRUCTURE ONLY. Recipe = SURVEY.md App. C (the reference's own cmake flow does not
# work in this image: cmake 4 policy floor, curses.h missing, vendored gtest configure fails).
# Sources are compiled where they lie under $REF; nothing is copied into the repository.
#   oracle/_ref/libref_harness.so   function-level oracle (ref_harness.c + reference TUs + libbam.a)
#   oracle/_ref/bam-somaticsniper   the reference CLI (-O3 -DNDEBUG -Wall, release flags)
#   oracle/_ref/samtools            samtools-0.1.6 (view/index) for synthetic BAM construction
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${SSN_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/src/lib/sniper" ]; then
    echo "build_ref: $REF not present; keeping prebuilt $OUT" >&2
    exit 0
fi
mkdir -p "$OUT/build"
ST="$OUT/build/samtools-0.1.6"
if [ ! -f "$ST/libbam.a" ]; then
    tar xzf "$REF/vendor/samtools-0.1.6.tar.gz" -C "$OUT/build"
    # -fPIC so libbam.a can be linked into the harness .so; otherwise the tarball's own flags (-g -Wall -O2)
    (cd "$ST" && make lib CFLAGS="-g -Wall -O2 -fPIC" >"$OUT/build/samtools_lib.log" 2>&1)
fi
if [ ! -x "$OUT/samtools" ]; then
    (cd "$ST" && make samtools CFLAGS="-g -Wall -O2 -fPIC" DFLAGS="-D_FILE_OFFSET_BITS=64 -D_USE_KNETFILE -D_CURSES_LIB=0" LIBCURSES= \
        >"$OUT/build/samtools_bin.log" 2>&1 || true
     gcc -O2 -o "$OUT/samtools" bam_tview.o bam_maqcns.o bam_plcmd.o sam_view.o bam_rmdup.o bam_rmdupse.o \
        bam_mate.o bam_stat.o bam_color.o bamtk.o -L. -lbam -lm -lz)
fi
mkdir -p "$OUT/build/version"
sed -e 's/@FULL_VERSION@/unstable/' -e 's/@COMMIT_HASH@/nogit/' -e 's/@CMAKE_BUILD_TYPE@/release/' \
    "$REF/version/version.h.in" > "$OUT/build/version/version.h"
INC="-I$REF/src/lib -I$REF/src/lib/sniper -I$ST -I$OUT/build/version"
# reference CLI, release flags (build-common/cmake/ProjectHelper.cmake:11-15, CMakeLists.txt:38-40)
gcc -Wall -O3 -DNDEBUG $INC "$REF"/src/lib/sniper/*.c "$REF/src/exe/bam-somaticsniper/main.c" \
    -o "$OUT/bam-somaticsniper" "$ST/libbam.a" -lm -lz 2>"$OUT/build/ref_cli.log"
# function-level harness: somatic_sniper.c is #included by ref_harness.c, so leave it out of the list
SRCS=""
for f in "$REF"/src/lib/sniper/*.c; do
    case "$f" in */somatic_sniper.c) ;; *) SRCS="$SRCS $f" ;; esac
done
gcc -Wall -O3 -DNDEBUG -fPIC -shared $INC "$HERE/ref_harness.c" $SRCS \
    -o "$OUT/libref_harness.so" "$ST/libbam.a" -lm -lz 2>"$OUT/build/ref_harness.log"
# INTEGRATION.md's binding, compiled: the reference's own main()/walker/formatters with the per-column callback replaced
# by tools/shim/ssn_shim.c (-> libssn.so).  Needs the built library; skipped (with a note) when it is not there yet.
ROOT="$(cd "$HERE/.." && pwd)"
LIBDIR="$ROOT/somatic_sniper_b200/lib"
if [ -f "$LIBDIR/libssn.so" ]; then
    OTHERS=""
    for f in "$REF"/src/lib/sniper/*.c; do OTHERS="$OTHERS $f"; done
    gcc -Wall -O2 -DNDEBUG $INC -I"$ROOT/include" -I"$REF/src/lib" \
        -Dglf_somatic=ssn_shim_column -DmakeSoloPrior=ssn_shim_solo_prior -Dbam_header_destroy=ssn_shim_finish \
        -c "$REF/src/exe/bam-somaticsniper/main.c" -o "$OUT/build/main_shim.o" 2>"$OUT/build/ref_shim.log"
    gcc -Wall -O2 -DNDEBUG $INC -I"$ROOT/include" -I"$REF/src/lib" "$OUT/build/main_shim.o" "$ROOT/tools/shim/ssn_shim.c" $OTHERS \
        -o "$OUT/bam-somaticsniper-shim" "$ST/libbam.a" -L"$LIBDIR" -lssn -Wl,-rpath,"$LIBDIR" -lm -lz 2>>"$OUT/build/ref_shim.log"
else
    echo "build_ref: libssn.so not built yet; skipping bam-somaticsniper-shim" >&2
fi
echo "build_ref: ok -> $OUT"
