#!/bin/bash
# usage: bash run.sh <gpus>
python -m torch.distributed.run --nnodes=1 --nproc-per-node "${1:-8}" --master-addr 127.0.0.1 --master-port 29631 main.py --cf fedml_config.yaml
