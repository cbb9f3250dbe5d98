#!/bin/bash
# Lean scaling session on N GPUs: bench.py (ResNet-50 + the secondary models, bf16 gradient wire), and at N = 8 also
# ResNet-152 and the graph-timed all-reduce+optimizer sweep; at N = 2 the multi-GPU test file.
#   scripts/run_scale.sh N   (writes gpurun_out/scale_N/)
N=${1:-8}
OUT=gpurun_out/scale_$N
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29522 bench.py --gpus $N --steps 30 --warmup 5 > $OUT/bench.json 2> $OUT/bench.err
python - "$OUT/bench.json" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d["config"]["model"], d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["clocks"],
          {k: (v["value"], v["ms_per_step"]) for k, v in d.get("secondary", {}).items()})
except Exception as e:
    print("bench FAILED", e)
PY
if [ "$N" = "8" ]; then
  $TR --master-port 29525 scripts/p2p_check.py nvls_small > $OUT/nvls_small.log 2>&1; grep "^{" $OUT/nvls_small.log | cut -c1-200
  $TR --master-port 29526 bench.py --gpus $N --steps 20 --warmup 5 --comm nvls --no-secondary > $OUT/bench_nvls.json 2> $OUT/bench_nvls.err; cut -c1-230 $OUT/bench_nvls.json
  $TR --master-port 29524 bench.py --model resnet152 --gpus $N --steps 10 --warmup 3 > $OUT/bench_resnet152.json 2> $OUT/bench_resnet152.err; cut -c1-220 $OUT/bench_resnet152.json
  rm -f gpurun_out/p2p_check.jsonl
  $TR --master-port 29521 scripts/p2p_check.py sweep2 > $OUT/sweep2.log 2>&1
  cp gpurun_out/p2p_check.jsonl $OUT/p2p_sweep2_${N}gpu.jsonl 2>/dev/null; tail -2 $OUT/sweep2.log | cut -c1-300
fi
if [ "$N" = "2" ]; then
  python -m pytest tests/test_gpu_multi.py -q > $OUT/pytest_multi.log 2>&1; tail -3 $OUT/pytest_multi.log
fi
