#!/bin/bash
# All five BASELINE.json configurations on one B200 (full sizes), then the headline bench.  Writes gpurun_out/all_configs.txt
set -u
out=gpurun_out/all_configs.txt
: > $out
python scripts/run_config.py pump 4 5000      2>&1 | tail -1 >> $out
python scripts/run_config.py pump 4096 2000   2>&1 | tail -1 >> $out
python scripts/run_config.py logistic 1024 50 2>&1 | tail -1 >> $out
python scripts/run_config.py lmm 8 500        2>&1 | tail -1 >> $out
python scripts/run_config.py occupancy 64 50  2>&1 | tail -1 >> $out
python scripts/run_config.py glmm 8192 10     2>&1 | tail -1 >> $out
GLMM_WARM=2500 GLMM_ONLY_GROUPS=1 python scripts/probe_glmm.py 2>&1 | tail -2 | sed 's/^/glmm after 2500 iterations (adapted proposals): /' >> $out
python scripts/probe_occ.py                   2>&1 | tail -2 >> $out
python scripts/probe_family.py lmm 8          2>&1 | tail -1 >> $out
python scripts/probe_family.py glmm 8192      2>&1 | tail -1 >> $out
cat $out
