#!/bin/bash
# Build kernel-experiment variants (here, no GPU): scripts/exp_variants.sh build "name:-DFOO -DBAR" ...
# Time them (on the box):                          scripts/exp_variants.sh run name ...
mode=$1; shift
root=$(cd $(dirname $0)/.. && pwd)
if [ "$mode" = build ]; then
  for spec in "$@"; do
    name=${spec%%:*}; flags=${spec#*:}; [ "$flags" = "$spec" ] && flags=""
    out=$root/transplat_b200/lib_exp/$name
    mkdir -p $out
    make -C $root/transplat_b200/csrc -j8 OUT=$out EXTRA="$flags" > $out/build.log 2>&1 || { echo "build $name failed"; tail -20 $out/build.log; exit 1; }
    echo "built $name [$flags]"
  done
else
  mkdir -p $root/gpurun_out
  for name in "$@"; do
    lib=$root/transplat_b200/lib_exp/$name/libtsplat.so
    [ "$name" = base ] && lib=$root/transplat_b200/lib/libtsplat.so
    TAG=$name TSPLAT_LIB_PATH=$lib timeout 120 python $root/scripts/stage_times.py 2>&1 | tail -1 | tee -a $root/gpurun_out/exp_variants.log
  done
fi
