run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) scripts/scale_config.py "${@:2}" 2>&1 | grep -E " on [0-9]+ GPU|Error|error" | tail -2; }
run 1 glmm 8192 3 1
run 2 glmm 8192 3 1
run 4 glmm 8192 3 1
run 4 glmm 8192 3 2
run 4 glmm 8192 3 4
run 1 lmm 8 100 1
run 2 lmm 8 100 2
run 4 lmm 8 100 4
run 4 lmm 8 100 2
