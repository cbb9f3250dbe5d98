#!/bin/bash
# One multi-GPU session: pytest multi-GPU file, graph-timed all-reduce+optimizer sweep, training benches.
#   scripts/run_multi.sh N   (N = 2, 4 or 8 GPUs; writes gpurun_out/multi_N/)
N=${1:-8}
OUT=gpurun_out/multi_$N
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
python -m pytest tests/test_gpu_multi.py -q > $OUT/pytest_multi.log 2>&1; tail -3 $OUT/pytest_multi.log
rm -f gpurun_out/p2p_check.jsonl
$TR --master-port 29521 scripts/p2p_check.py sweep2 > $OUT/sweep2.log 2>&1
cp gpurun_out/p2p_check.jsonl $OUT/p2p_sweep2_${N}gpu.jsonl 2>/dev/null; tail -2 $OUT/sweep2.log | cut -c1-300
for cfg in "bf16:" "fp32:" "fp32:--comm nvls"; do
  wire=${cfg%%:*}; extra=${cfg#*:}
  tag=${wire}$(echo $extra | tr -d ' -')
  TDL_GRAD_WIRE=$wire $TR --master-port 29522 bench.py --gpus $N --steps 20 --warmup 5 $extra > $OUT/bench_$tag.json 2> $OUT/bench_$tag.err
  python - "$OUT/bench_$tag.json" "$tag" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], d["config"]["model"], d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"],
          {k: (v["value"], v["ms_per_step"]) for k, v in d.get("secondary", {}).items()})
except Exception as e:
    print(sys.argv[2], "FAILED", e)
PY
done
$TR --master-port 29523 bench.py --impl proxy --gpus $N --steps 20 --warmup 5 > $OUT/bench_proxy.json 2> $OUT/bench_proxy.err; cut -c1-200 $OUT/bench_proxy.json
$TR --master-port 29524 bench.py --model resnet152 --gpus $N --steps 10 --warmup 3 > $OUT/bench_resnet152.json 2> $OUT/bench_resnet152.err; cut -c1-200 $OUT/bench_resnet152.json
