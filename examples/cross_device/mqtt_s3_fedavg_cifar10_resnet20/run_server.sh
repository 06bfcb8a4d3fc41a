#!/bin/bash
python server.py --cf fedml_config.yaml --rank 0 --run_id "${RUN_ID:-example}"
