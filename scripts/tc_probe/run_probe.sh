#!/bin/bash
# runs every probe case in its own process under a timeout (a hung or faulting case must not take the others down)
cd "$(dirname "$0")"
mkdir -p ../../gpurun_out
L=../../gpurun_out/tc_probe.log
: > $L
timeout 60 ./tc_probe 1 >> $L 2>&1; echo "[exit $?] T1" >> $L
for i in 0 1 2 3 4 5 6 7 8 9; do timeout 60 ./tc_probe 2 $i >> $L 2>&1; echo "[exit $?] T2.$i" >> $L; done
for i in 0 1 2 3 4 5 6; do timeout 60 ./tc_probe 3 $i >> $L 2>&1; echo "[exit $?] T3.$i" >> $L; done
cat $L
