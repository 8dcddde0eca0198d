#!/bin/bash
# Tuning variants of the robot kernel: tools/build_variants.sh "name:EXTRA flags" ...
# -> multibody.jl_b200/lib/variants/lib_<name>.so (select with MBB200_LIB=...)
set -e
cd "$(dirname "$0")/../multibody.jl_b200/csrc"
mkdir -p ../lib/variants
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  ( make -s MODELS=${MODELS:-robot6dof} BUILD=build_$name LIB=../lib/variants/lib_$name.so EXTRA="$flags" 2>&1 | grep -v "deprecated-gpu-targets" || true ) &
done
wait
ls -la ../lib/variants
