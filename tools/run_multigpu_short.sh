#!/bin/bash
# Short multi-GPU pass: headline bench (+ the listed configs) and the multi-GPU tests.  gpurun --gpus N -- 'tools/run_multigpu_short.sh N cfg...'
N="${1:-2}"; shift
OUT="gpurun_out/short_n${N}"
mkdir -p "$OUT"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
python -c "from fedml_b200.native.build import build; build()" > /dev/null 2>&1
timeout 200 $TR --master-port 29811 bench.py --gpus "$N" --steps 5 --warmup 3 > "$OUT/headline.log" 2>&1
echo "[headline] rc=$? $(tail -1 "$OUT/headline.log" | cut -c1-200)"
for cfg in "$@"; do
  if [ "$cfg" = "tests" ]; then
    timeout 300 python -m pytest tests/test_multigpu.py -q -x -s > "$OUT/test_multigpu.log" 2>&1
    echo "[test_multigpu] rc=$? $(tail -1 "$OUT/test_multigpu.log")"
  else
    timeout 200 $TR --master-port 29812 bench.py --gpus "$N" --steps 3 --warmup 3 --config "$cfg" --no-bf16 > "$OUT/$cfg.log" 2>&1
    echo "[$cfg] rc=$? $(tail -1 "$OUT/$cfg.log" | cut -c1-200)"
  fi
done
for f in "$OUT"/*.log; do grep -h '^{' "$f" | tail -1 > "${f%.log}.json" 2>/dev/null; done
