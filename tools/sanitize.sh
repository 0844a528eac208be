#!/bin/bash
# usage (under gpurun, ONE tool per call): tools/sanitize.sh memcheck|racecheck|initcheck|synccheck [env assignments]
tool=$1; shift
python tools/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
env "$@" timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_case.py > gpurun_out/sanitize_${tool}${1:+_umma}.log 2>&1
echo "exit $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|keypoints|Error|hazard" gpurun_out/sanitize_${tool}${1:+_umma}.log | head -20
