#!/bin/bash
# usage: ncu_capture.sh <tag> <kernel regex> <skip> <env or -> <python script and args...>
# one --set full capture of one launch; raw and SASS-source pages exported as CSV into gpurun_out/ (the report itself stays on the box)
tag=$1; re=$2; skip=$3; envs=$4; shift 4
[ "$envs" = "-" ] && envs=""
env $envs timeout 400 ncu --set full --clock-control none --import-source on -k regex:"$re" -s $skip -c 1 -f -o /tmp/$tag "$@" > gpurun_out/${tag}_ncu.log 2>&1
ncu -i /tmp/$tag.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
ncu -i /tmp/$tag.ncu-rep --page source --csv --print-source sass > gpurun_out/${tag}_sass.csv 2>/dev/null
ls -la gpurun_out/${tag}_raw.csv gpurun_out/${tag}_sass.csv
