#!/bin/bash
# builds a kernel variant of libfxgpu.so: scripts/build_variant.sh NAME "-DFXG_VAR_X=1 ..."  -> fxtract_b200/_build/var/NAME.so
# (only fxg_scan.cu -- and fxg_scan_kmer.cu with KMER=1 -- are recompiled; the other objects come from the default build)
set -e
NAME=$1; FLAGS=$2
B=fxtract_b200/_build; C=fxtract_b200/csrc
mkdir -p $B/var/obj_$NAME
NV="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC"
nvcc $NV $FLAGS -c -o $B/var/obj_$NAME/fxg_scan.o $C/fxg_scan.cu &
if [ -n "$KMER" ]; then nvcc $NV $FLAGS -c -o $B/var/obj_$NAME/fxg_scan_kmer.o $C/fxg_scan_kmer.cu & fi
wait
K=$B/fxg_scan_kmer.o; [ -n "$KMER" ] && K=$B/var/obj_$NAME/fxg_scan_kmer.o
A=$B/fxg_api.o
if [ -n "$API" ]; then nvcc $NV $FLAGS -c -o $B/var/obj_$NAME/fxg_api.o $C/fxg_api.cu; A=$B/var/obj_$NAME/fxg_api.o; fi   # API=1: host tables depend on the flags too
nvcc $NV -shared -o $B/var/$NAME.so $B/var/obj_$NAME/fxg_scan.o $K $B/fxg_fields.o $B/fxg_synth.o $A
echo built $B/var/$NAME.so
