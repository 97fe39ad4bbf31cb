#!/bin/bash
# the working tree's library with extra nvcc flags, for A/B runs: tools/build_flags_variant.sh <name> [-DFLAG ...]
# -> tools/variants/<name>/libohcu.so (+ a copy of the C driver beside it); run with OHCU_LIB=tools/variants/<name>/libohcu.so
set -e
name=$1; shift
d=/root/repo/tools/variants/$name
mkdir -p $d
cd /root/repo/hevc_b200/csrc
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared "$@" -o $d/libohcu.so api.cu recon.cu loopfilter.cu picture.cu bs.cu prepare.cu intra.cu
cp /root/repo/hevc_b200/libohcu_drv.so $d/
