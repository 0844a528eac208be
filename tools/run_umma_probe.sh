#!/bin/bash
P=tools/bin/umma_probe
O=gpurun_out/umma_probe.log
: > $O
run() { echo "== $*" >> $O; timeout 60 $P "$@" >> $O 2>&1; echo "exit $?" >> $O; }
run 0
run 1
run 0 timing=1 reps=200
run 1 timing=1 reps=200
run 0 timing=1 reps=200 nv=128
cat $O
