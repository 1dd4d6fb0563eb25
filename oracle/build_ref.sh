#!/usr/bin/env bash
# build_ref.sh <reference_root> <out_dir>      TEST INFRASTRUCTURE ONLY.
#
# Compiles the REFERENCE's own sources (from a headless-patched scratch copy under /tmp, see patch_ref.py;
# the reference's CMake build is not used and no reference source enters the repo) into
#   <out_dir>/libtitan_ref_host.so   host-callable reference code + oracle/ref_host_shim.cu   (CPU, runs here)
#   <out_dir>/titan_ref_cuda         the reference's CUDA engine + tests/cpp/titan_api_driver.cu (needs a GPU)
#   <out_dir>/titan_ref_cuda_verlet  the same with the reference's VERLET option (CMakeLists.txt:13,90-93)
#   <out_dir>/titan_ref_cuda_rk2     the same with its RK2 option (CMakeLists.txt:14,95-100), call sites repaired by patch_ref.py
# Flags follow the reference's CMakeLists.txt: C++14, separable compilation, -DCONSTRAINTS, no GRAPHICS.
set -euo pipefail
REF="${1:-/root/reference}"
OUT="${2:-$(dirname "$0")/_ref}"
HERE="$(cd "$(dirname "$0")" && pwd)"
if [ ! -d "$REF/src" ]; then echo "build_ref: no reference at $REF (skipped)"; exit 0; fi
mkdir -p "$OUT"
SCRATCH="$(mktemp -d /tmp/titan_ref_build.XXXXXX)"
trap 'rm -rf "$SCRATCH"' EXIT
python3 "$HERE/patch_ref.py" "$REF" "$SCRATCH/ref"
S="$SCRATCH/ref"
# the distro g++ as host compiler (the exported CXX=/opt/gcc/bin/g++ lacks pieces, see oracle/Makefile)
NVFLAGS="-std=c++14 -O3 -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a -rdc=true -DCONSTRAINTS \
  -Xcompiler -fPIC -Xcompiler -ffp-contract=off -I$S/include -I$S/include/Titan -w"
for f in vec mass spring object sim; do
  nvcc $NVFLAGS -c "$S/src/$f.cu" -o "$SCRATCH/$f.o" &
done
nvcc $NVFLAGS -DVERLET -c "$S/src/sim.cu" -o "$SCRATCH/sim_verlet.o" &
# the RK2 option needs C++17 (if constexpr) and is a PUBLIC define (CUDA_MASS grows two backup vectors): everything is rebuilt
NVFLAGS_RK2="${NVFLAGS/-std=c++14/-std=c++17} -DRK2"
for f in vec mass spring object sim; do
  nvcc $NVFLAGS_RK2 -c "$S/src/$f.cu" -o "$SCRATCH/${f}_rk2.o" &
done
nvcc $NVFLAGS_RK2 -c "$HERE/../tests/cpp/titan_api_driver.cu" -o "$SCRATCH/dump_rk2.o" &
nvcc $NVFLAGS -c "$HERE/ref_host_shim.cu" -o "$SCRATCH/shim.o" &
nvcc $NVFLAGS -c "$HERE/../tests/cpp/titan_api_driver.cu" -o "$SCRATCH/dump.o" &
wait
nvcc -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a -shared -Xcompiler -fPIC \
  "$SCRATCH/vec.o" "$SCRATCH/mass.o" "$SCRATCH/spring.o" "$SCRATCH/object.o" "$SCRATCH/shim.o" \
  -o "$OUT/libtitan_ref_host.so" -lcudart
nvcc -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a \
  "$SCRATCH/vec.o" "$SCRATCH/mass.o" "$SCRATCH/spring.o" "$SCRATCH/object.o" "$SCRATCH/sim.o" "$SCRATCH/dump.o" \
  -o "$OUT/titan_ref_cuda" -lcudart
nvcc -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a \
  "$SCRATCH/vec.o" "$SCRATCH/mass.o" "$SCRATCH/spring.o" "$SCRATCH/object.o" "$SCRATCH/sim_verlet.o" "$SCRATCH/dump.o" \
  -o "$OUT/titan_ref_cuda_verlet" -lcudart
nvcc -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a \
  "$SCRATCH/vec_rk2.o" "$SCRATCH/mass_rk2.o" "$SCRATCH/spring_rk2.o" "$SCRATCH/object_rk2.o" "$SCRATCH/sim_rk2.o" "$SCRATCH/dump_rk2.o" \
  -o "$OUT/titan_ref_cuda_rk2" -lcudart
echo "build_ref: built $OUT/libtitan_ref_host.so, $OUT/titan_ref_cuda and $OUT/titan_ref_cuda_verlet"
