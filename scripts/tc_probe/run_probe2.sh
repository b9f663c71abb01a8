#!/bin/bash
cd "$(dirname "$0")"
mkdir -p ../../gpurun_out
L=../../gpurun_out/tc_probe2.log
: > $L
for i in 2 3 6 8; do timeout 60 ./tc_probe 2 $i >> $L 2>&1; echo "[exit $?] T2.$i" >> $L; done
cat $L
