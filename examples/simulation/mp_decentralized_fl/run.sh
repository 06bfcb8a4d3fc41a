#!/bin/bash
# usage: bash run.sh <processes = workers + 1>
N="${1:-5}"
python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29611 main.py --cf fedml_config.yaml
