#!/bin/bash
# usage: bash run_silo.sh <silo rank 1..n> <processes in the silo>
python -m torch.distributed.run --nnodes=1 --nproc-per-node "${2:-2}" --master-addr 127.0.0.1 --master-port $((29411 + $1)) client.py --cf fedml_config.yaml --rank "$1" --run_id "${RUN_ID:-example}"
