#!/bin/bash
# A/B builds of one kernel file: scripts/diag/build_variants.sh <file.cu> <name>:<-Dflags...> [<name>:<flags> ...]
# -> pde_control_b200/variants/libpdec_<name>.so (other objects taken from pde_control_b200/build); select with PDEC_LIB_PATH.
set -e
cd "$(dirname "$0")/../.."
python -m pde_control_b200.build >/dev/null
SRC=$1; shift
FMAD=""; case $SRC in burgers.cu|burgers_fast2.cu|fluid.cu) FMAD="-fmad=false";; esac
mkdir -p pde_control_b200/variants
for spec in "$@"; do
  name=${spec%%:*}; flags=${spec#*:}
  obj=/tmp/variant_${name}.o
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $FMAD $flags -c pde_control_b200/csrc/$SRC -o $obj
  others=$(ls pde_control_b200/build/*.o | grep -v "/${SRC}.o")
  nvcc -shared -o pde_control_b200/variants/libpdec_${name}.so $others $obj -lcudart
  echo "built $name ($flags)"
done
