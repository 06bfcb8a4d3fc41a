#!/bin/bash
# usage: bash run_client.sh <rank 1..n>
python client.py --cf fedml_config.yaml --rank "$1" --run_id "${RUN_ID:-example}"
