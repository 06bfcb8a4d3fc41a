#!/bin/bash
# usage: bash run.sh <gpus>
N="${1:-8}"
python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29621 main.py --cf fedml_config.yaml
