#!/bin/bash
# Build an experiment variant of the CUDA library into variants/<name>/ (git-ignored; travels with gpurun): tools/variant.sh <name> <extra nvcc flags>
# Run it with BNN_B200_LIB_DIR=variants/<name> python bench.py ...
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
w=/tmp/variant_$name
rm -rf $w && mkdir -p $w/pkg && cp -r $root/include $w/ && cp -r $root/bnn-pynq_b200/Makefile $root/bnn-pynq_b200/csrc $w/pkg/
make -C $w/pkg -j8 NVEXTRA="$*" lib/libbnn_b200.so 2>&1 | grep -i "error" && exit 1
mkdir -p $root/variants/$name && cp $w/pkg/lib/libbnn_b200.so $root/variants/$name/
grep -A3 "k_cnv01_tcILi0ELb1" $w/pkg/build/cnv01_tc.ptxas.log | grep Used
