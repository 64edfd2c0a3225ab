#!/usr/bin/env bash
# Build the UNMODIFIED reference (gol-main.c + gol-with-cuda.cu) into baseline/_ref/.
#   baseline/_ref/src/            verbatim copies of the two reference sources (sha256 recorded)
#   baseline/_ref/bin/gol_ref     the reference as shipped (official bar)
#   baseline/_ref/bin/gol_ref_noprintf   steelman: the device printf at gol-with-cuda.cu:223
#                                 removed in a build-time copy (the ONLY change)
#   baseline/_ref/bin/gol_ref_rect[_noprintf]   strong-scaling variant: the reference's main with ONE change made
#                                 at build time - gol_initMaster (gol-main.c:76) gets (W, H) =
#                                 (worldSize, $GOL_REF_ROWS) instead of (worldSize, worldSize); the .cu is
#                                 untouched (its API already takes W and H, gol-with-cuda.cu:286)
#   baseline/_ref/bin/mpirun      minimpi launcher (this image has no MPI)
# The reference ships no build file; flags follow SURVEY.md section 6.3.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${1:-/root/reference}"
OUT="$HERE/_ref"
mkdir -p "$OUT/src" "$OUT/bin" "$OUT/obj"
if [ "$(cd "$REF" && pwd)" != "$(cd "$OUT/src" && pwd)" ]; then   # rebuild in place: sources already there
  cp "$REF/gol-main.c" "$REF/gol-with-cuda.cu" "$OUT/src/"
fi
( cd "$OUT/src" && sha256sum gol-main.c gol-with-cuda.cu > SHA256SUMS )
ARCH="-gencode arch=compute_100,code=sm_100"
gcc -O2 -c "$HERE/minimpi/minimpi.c" -o "$OUT/obj/minimpi.o"
gcc -O2 -o "$OUT/bin/mpirun" "$HERE/minimpi/minimpirun.c" "$OUT/obj/minimpi.o" -lrt
gcc -O2 -w -I"$HERE/minimpi" -c "$OUT/src/gol-main.c" -o "$OUT/obj/gol-main.o"
nvcc -O3 $ARCH -w -c "$OUT/src/gol-with-cuda.cu" -o "$OUT/obj/gol-with-cuda.o"
nvcc $ARCH -o "$OUT/bin/gol_ref" "$OUT/obj/gol-main.o" "$OUT/obj/gol-with-cuda.o" "$OUT/obj/minimpi.o" -lrt
# steelman variant: drop the one debugging printf line
sed '223{/printf("BEFORE AND");/d}' "$OUT/src/gol-with-cuda.cu" > "$OUT/obj/gol-with-cuda-noprintf.cu"
if cmp -s "$OUT/src/gol-with-cuda.cu" "$OUT/obj/gol-with-cuda-noprintf.cu"; then
  echo "build_ref: printf line not found at gol-with-cuda.cu:223" >&2; exit 1
fi
nvcc -O3 $ARCH -w -c "$OUT/obj/gol-with-cuda-noprintf.cu" -o "$OUT/obj/gol-with-cuda-noprintf.o"
nvcc $ARCH -o "$OUT/bin/gol_ref_noprintf" "$OUT/obj/gol-main.o" "$OUT/obj/gol-with-cuda-noprintf.o" "$OUT/obj/minimpi.o" -lrt
# strong-scaling variant: rectangular slabs through the unchanged .cu API
sed '76s/gol_initMaster(pattern, worldSize, worldSize, myRank, numRank);/gol_initMaster(pattern, worldSize, getenv("GOL_REF_ROWS") ? (unsigned int)atoi(getenv("GOL_REF_ROWS")) : worldSize, myRank, numRank);/' \
  "$OUT/src/gol-main.c" > "$OUT/obj/gol-main-rect.c"
if cmp -s "$OUT/src/gol-main.c" "$OUT/obj/gol-main-rect.c"; then
  echo "build_ref: gol_initMaster call not found at gol-main.c:76" >&2; exit 1
fi
if [ "$(diff "$OUT/src/gol-main.c" "$OUT/obj/gol-main-rect.c" | grep -c '^[<>]')" != "2" ]; then
  echo "build_ref: the rectangular variant differs from gol-main.c in more than one line" >&2; exit 1
fi
gcc -O2 -w -I"$HERE/minimpi" -c "$OUT/obj/gol-main-rect.c" -o "$OUT/obj/gol-main-rect.o"
nvcc $ARCH -o "$OUT/bin/gol_ref_rect" "$OUT/obj/gol-main-rect.o" "$OUT/obj/gol-with-cuda.o" "$OUT/obj/minimpi.o" -lrt
nvcc $ARCH -o "$OUT/bin/gol_ref_rect_noprintf" "$OUT/obj/gol-main-rect.o" "$OUT/obj/gol-with-cuda-noprintf.o" "$OUT/obj/minimpi.o" -lrt
echo "built: $(ls "$OUT/bin")"
