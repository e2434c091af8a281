#!/bin/bash
# build a variant of the library next to the in-tree one: tools/build_variant.sh <suffix> <nvcc defines...>
# (run here; the .so travels to the GPU box; tools/gpu_ab.sh compares it with the in-tree build)
sfx=$1; shift
QUADE_LIB_PATH=$PWD/quade_b200/libquade_b200_${sfx}.so QUADE_BUILD_DEFS="$*" python -m quade_b200.build --force 2>&1 | grep -v RuntimeWarning | tail -1
