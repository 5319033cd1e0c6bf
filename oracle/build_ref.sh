#!/bin/bash
# TEST INFRASTRUCTURE ONLY -- builds the *unmodified* reference (sfiligoi/omp-bowtie2-prime)
# from the sources where they lie under /root/reference into oracle/_ref/.
# No reference source is copied into this repo; outputs (binaries) go only to oracle/_ref/
# (git-ignored, travels to the GPU box with the gpurun snapshot).
#
# Flags = the reference Makefile's (Makefile:39,75-84,208-233,294-302) minus the nvc++-only
# -DASM_PREFETCH (bt2_idx.h:63-68), compiled with g++ directly (the reference's build system
# is NOT run).  bowtie2-build-l is linked by hand (its Makefile rule has an empty SHARED_CPPS).
set -e
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
if [ ! -d "$REF" ]; then
  echo "build_ref.sh: $REF absent (GPU box?) -- using prebuilt files in $OUT" >&2
  exit 0
fi
mkdir -p "$OUT"
GCCLIB=$(dirname "$(g++ -print-libgcc-file-name)")
COMMON="-std=c++17 -DFORCE_ALL_OMP -fopenmp -B$GCCLIB -L$GCCLIB -DPOPCNT_CAPABILITY -DNO_SPINLOCK \
 -DWITH_QUEUELOCK=1 -w -fno-strict-aliasing -DBOWTIE2_VERSION=\"2.5.4\" -DBUILD_HOST=\"graft\" \
 -DBUILD_TIME=\"na\" -DCOMPILER_VERSION=\"gcc\" -DCOMPILER_OPTIONS=\"-O3\" -D_LARGEFILE_SOURCE \
 -D_FILE_OFFSET_BITS=64 -D_GNU_SOURCE -DBOWTIE_MM -DBOWTIE2 -DBOWTIE_64BIT_INDEX -DNDEBUG \
 -I$REF -I$REF/third_party"

build_align() { # name extra-flags
  local name=$1; shift
  if [ "$OUT/$name" -nt "$REF/bt2_search.cpp" ] && [ -z "$FORCE" ]; then return; fi
  echo "[build_ref] $name" >&2
  g++ -O3 -msse2 $COMMON "$@" -o "$OUT/$name" "$REF/bowtie2-align.cpp" -lpthread -lz
}

build_build() {
  if [ "$OUT/bowtie2-build-l" -nt "$REF/bt2_build.cpp" ] && [ -z "$FORCE" ]; then return; fi
  echo "[build_ref] bowtie2-build-l" >&2
  printf '#include <utility>\n#include <string>\n#include <iostream>\nusing namespace std;\n' > "$OUT/pre.h"
  local S=""
  for f in bt2_build bowtie_build_main diff_sample ccnt_lut ref_read alphabet shmem edit bt2_idx bt2_io \
           bt2_locks bt2_util reference ds multikey_qsort limit random_source; do S="$S $REF/$f.cpp"; done
  g++ -include "$OUT/pre.h" -O2 -msse2 $COMMON -o "$OUT/bowtie2-build-l" $S -lpthread -lz
}

# The reference with its phases P1-P6 redirected to libbt2gpu.so (the drop-in, built for real): a scratch copy of the
# reference + integration/bt2_search_gpu.patch (3 friend declarations, one if() around P1-P6) + integration/gpu_batch.cpp (ours),
# compiled with the flags above and linked against the in-tree library.  Runs the GPU path only with BT2GPU=1 in the environment.
build_gpu() {
  local LIB="$HERE/../omp-bowtie2-prime_b200/libbt2gpu.so" INT="$HERE/../integration"
  if [ ! -f "$LIB" ]; then echo "[build_ref] libbt2gpu.so not built yet: skipping bowtie2-align-l-gpu" >&2; return; fi
  if [ "$OUT/bowtie2-align-l-gpu" -nt "$INT/gpu_batch.cpp" ] && [ "$OUT/bowtie2-align-l-gpu" -nt "$INT/bt2_search_gpu.patch" ] && \
     [ "$OUT/bowtie2-align-l-gpu" -nt "$HERE/../include/bt2gpu.h" ] && [ -z "$FORCE" ]; then return; fi
  echo "[build_ref] bowtie2-align-l-gpu" >&2
  local T; T=$(mktemp -d /tmp/bt2ref_gpu.XXXXXX)
  cp -r "$REF"/. "$T"/ && chmod -R u+w "$T"
  (cd "$T" && patch -p1 -s < "$INT/bt2_search_gpu.patch") || { echo "[build_ref] patch does not apply" >&2; rm -rf "$T"; return 1; }
  cp "$INT/gpu_batch.cpp" "$HERE/../include/bt2gpu.h" "$T"/
  local C2="${COMMON//$REF/$T}"
  g++ -O3 -msse2 $C2 -DBT2GPU -o "$OUT/bowtie2-align-l-gpu" "$T/bowtie2-align.cpp" -L"$(dirname "$LIB")" -lbt2gpu \
      -Wl,-rpath,'$ORIGIN/../../omp-bowtie2-prime_b200' -lpthread -lz || { rm -rf "$T"; return 1; }
  rm -rf "$T"
}

# Oracle D: the -DENABLE_I16 reference with its two size constants raised in a scratch copy (aligner_result.h:42-43
# ALN_MAX_ROWS 160 -> 1000, ALN_MAX_COLS 200 -> 1072) -- NOT the reference as shipped: it pins what "the same algorithm without the
# 160-row limit" prints, for reads the shipped binary leaves unaligned.  Run it with -p 1: its DP buffers are sized per read slot.
build_long() {
  if [ "$OUT/bowtie2-align-l-long" -nt "$REF/bt2_search.cpp" ] && [ -z "$FORCE" ]; then return; fi
  echo "[build_ref] bowtie2-align-l-long (constants raised: not the reference as shipped)" >&2
  local T; T=$(mktemp -d /tmp/bt2ref_long.XXXXXX)
  cp -r "$REF"/. "$T"/ && chmod -R u+w "$T"
  sed -i 's/^#define ALN_MAX_ROWS 160/#define ALN_MAX_ROWS 1000/; s/^#define ALN_MAX_COLS 200/#define ALN_MAX_COLS 1072/' "$T/aligner_result.h"
  grep -q "ALN_MAX_ROWS 1000" "$T/aligner_result.h" && grep -q "ALN_MAX_COLS 1072" "$T/aligner_result.h" || { echo "[build_ref] constants not found" >&2; rm -rf "$T"; return 1; }
  local C2="${COMMON//$REF/$T}"
  g++ -O3 -msse2 $C2 -DENABLE_I16 -o "$OUT/bowtie2-align-l-long" "$T/bowtie2-align.cpp" -lpthread -lz || { rm -rf "$T"; return 1; }
  rm -rf "$T"
}

build_align bowtie2-align-l &
build_long &
build_build &
build_gpu &
# the opt-in 16-bit build (SURVEY 8c oracle mode B): pins the i16 protocol (tests/golden/*_i16.sam)
if [ -z "$NO_I16" ]; then build_align bowtie2-align-l-i16 -DENABLE_I16 & fi
wait
ls -la "$OUT" >&2
