#!/usr/bin/env bash
# Build the UNMODIFIED reference programs from /root/reference into baseline/_ref (git-ignored).
# The image has no MPI, so the MPI programs are compiled against the stand-in baseline/mpi_shim/mpi.h
# (shared-memory mailboxes, single node) — disclosed next to every reference number.
#   baseline/_ref/src/      verbatim copy of the reference sources
#   baseline/_ref/bin/      cuda_sol (sm_100a), openmp_sol, mpi_sol, mpi_new, hybrid_sol, hybrid_new
#   baseline/_ref/p17/      cuda_sol_p17: same sources with `out.precision(17)` injected by sed into a COPY,
#                           used only by the parity test to read full-precision maxima
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF_SRC="${REF_SRC:-/root/reference}"
OUT="$HERE/_ref"
ARCH="${ARCH:-sm_100a}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
mkdir -p "$OUT/src" "$OUT/bin" "$OUT/p17" "$OUT/obj"
if [ -d "$REF_SRC" ]; then
  cp -f "$REF_SRC"/*.cpp "$REF_SRC"/*.cu "$REF_SRC"/Makefile "$REF_SRC"/README.txt "$OUT/src/"
fi
[ -f "$OUT/src/cuda_sol.cpp" ] || { echo "reference sources not found"; exit 1; }
cd "$OUT"
SHIM="-I$HERE/mpi_shim"
# --- MPI+CUDA program (the one the new solver is measured against); flags as in the reference Makefile
$NVCC -arch=$ARCH -O3 -Xptxas -v -c src/cuda_sol_kernels.cu -o obj/cuda_sol_kernels.o 2> obj/cuda_sol_kernels.ptxas.log
g++ -std=c++11 -O3 -I/usr/local/cuda/include $SHIM -c src/cuda_sol.cpp -o obj/cuda_sol.o
$NVCC -arch=$ARCH -o bin/cuda_sol obj/cuda_sol.o obj/cuda_sol_kernels.o -lrt
# --- full-precision copy for the parity test only
sed 's/out.open(/out.precision(17); out.open(/' src/cuda_sol.cpp > p17/cuda_sol_p17.cpp
g++ -std=c++11 -O3 -I/usr/local/cuda/include $SHIM -c p17/cuda_sol_p17.cpp -o obj/cuda_sol_p17.o
$NVCC -arch=$ARCH -o p17/cuda_sol_p17 obj/cuda_sol_p17.o obj/cuda_sol_kernels.o -lrt
# --- CPU programs (compile lines from README.txt, g++ instead of xlc++)
g++ -std=c++11 -O3 -fopenmp src/openmp_sol.cpp -o bin/openmp_sol
for p in mpi_sol mpi_new; do g++ -std=c++11 -O3 $SHIM src/$p.cpp -o bin/$p -lrt; done
for p in hybrid_sol hybrid_new; do g++ -std=c++11 -O3 -fopenmp $SHIM src/$p.cpp -o bin/$p -lrt; done
ls -la bin p17/cuda_sol_p17
