#!/bin/bash
# tools/variant.sh NAME SOURCE.cu [-D...]: builds experiments/variants/libNAME.so = the current objects with
# SOURCE.cu recompiled under the extra flags (A/B runs on one GPU box: CLIMA_RT_LIBRARY=experiments/variants/libNAME.so)
set -e
cd "$(dirname "$0")/.."
name=$1; src=$2; shift 2
mkdir -p experiments/variants
objs=""
for s in capi prep optics_rorr optics_rorr_fast optics_rorr_exact optics_rorr_cut jacobian_delta adiabat_profile twostream fp64_peak; do
  if [ "$s.cu" == "$src" ]; then
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr --extended-lambda "$@" -c clima_b200/csrc/$s.cu -o experiments/variants/$name.$s.o
    objs="$objs experiments/variants/$name.$s.o"
  else
    objs="$objs clima_b200/csrc/$s.o"
  fi
done
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o experiments/variants/lib$name.so $objs -lcudart
rm -f experiments/variants/$name.*.o
echo experiments/variants/lib$name.so
