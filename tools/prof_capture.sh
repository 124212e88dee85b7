#!/bin/bash
# ncu captures of one featurizer kernel for profiles/: full-size launch (traffic, pipes, stalls) and a short
# launch with source counters.  usage: tools/prof_capture.sh <tag> <kind> <kernel-regex> <clips-big> <clips-small>
set -u
tag=$1; kind=$2; kre=$3; big=$4; small=$5
out=gpurun_out/$tag; mkdir -p $out
python tools/prof_run.py $kind $small > $out/run.log 2>&1 || { echo "plain run failed"; tail -5 $out/run.log; exit 1; }
timeout 900 ncu --set full --clock-control none -k regex:$kre -s 2 -c 1 -f -o $out/big python tools/prof_run.py $kind $big > $out/ncu_big.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$kre -s 2 -c 1 -f -o $out/src python tools/prof_run.py $kind $small > $out/ncu_src.log 2>&1
ls -la $out
