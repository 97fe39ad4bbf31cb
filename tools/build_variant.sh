#!/bin/bash
# build the library of another git revision next to the working tree's: tools/build_variant.sh <rev> <out.so>
set -e
rev=$1; out=$2
tmp=$(mktemp -d)
git -C /root/repo archive $rev hevc_b200/csrc include | tar -x -C $tmp
cd $tmp/hevc_b200/csrc
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared -I$tmp/include -o $out $(ls *.cu)
rm -rf $tmp
