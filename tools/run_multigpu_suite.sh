#!/bin/bash
# Everything that needs N GPUs of one box, in one gpurun call:  gpurun --gpus N -- 'tools/run_multigpu_suite.sh N'
# Outputs: gpurun_out/suite_n<N>/*.json|log  (copy the JSON lines you want judged into profiles/).
N="${1:-8}"
OUT="gpurun_out/suite_n${N}"
mkdir -p "$OUT"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() {  # name, port, command...
  local name="$1"; local port="$2"; shift 2
  timeout 200 $TR --master-port "$port" "$@" > "$OUT/$name.log" 2>&1
  echo "[$name] rc=$? $(tail -1 "$OUT/$name.log" | cut -c1-300)"
}
python -c "from fedml_b200.native.build import build; build()" > /dev/null 2>&1
run headline 29801 bench.py --gpus "$N" --steps 5 --warmup 3
for cfg in resnet18_fedopt_dp resnet18_robust_median resnet18_robust_trimmed resnet18_robust_krum distilbert_fednlp; do
  run "$cfg" 29802 bench.py --gpus "$N" --steps 3 --warmup 3 --config "$cfg" --no-bf16
done
run aggregation 29803 tools/bench_aggregation.py
MODEL=resnet18 run cheetah_resnet18 29804 tools/bench_cheetah.py
MODEL=distilbert B=32 run cheetah_distilbert 29805 tools/bench_cheetah.py
timeout 200 python -m pytest tests/test_multigpu.py -q -x -s > "$OUT/test_multigpu.log" 2>&1
echo "[test_multigpu] rc=$? $(tail -1 "$OUT/test_multigpu.log")"
for f in "$OUT"/*.log; do grep -h '^{' "$f" | tail -1 > "${f%.log}.json" 2>/dev/null; done
