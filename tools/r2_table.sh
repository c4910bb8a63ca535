#!/bin/bash
# Per-config bench lines at N GPUs of one box: tools/gpu.sh 1500 'bash tools/r2_table.sh N'  (gpurun --gpus N for N > 1).
# C3 is the headline line (it carries the c4_tiles record); C1, C2, C4, C5 through --workload.  One JSON line per
# workload into gpurun_out/r2_bench_<workload>_n<N>.json.
N=${1:-1}
O=gpurun_out
mkdir -p $O
run() {
  if [ "$N" = "1" ]; then python bench.py --gpus 1 "$@"
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; fi
}
WL="C3 C1 C2 C4 C5"; [ "$N" != "1" ] && WL="C3 C1 C2 C5"   # C4 tile-sharded rides along with C3 as c4_tiles
for W in $WL; do
  extra="--no-c4-tiles"; [ "$W" = "C3" ] && extra=""
  run --workload $W --steps 4 --warmup 3 --no-cpu-baseline $extra 2> $O/r2_bench_${W}_n${N}.err | grep "^{" > $O/r2_bench_${W}_n${N}.json
  echo "$W rc=${PIPESTATUS[0]}"
  python - <<PY
import json
try:
    d = json.loads(open("$O/r2_bench_${W}_n${N}.json").read().strip().splitlines()[-1])
    print("$W n=$N", round(d["value"], 1), d["unit"], "ms/step", round(d["ms_per_step"], 2), d["scaling"], d["config"].get("sharding"), "| e2e", round(d["e2e"]["value"], 1),
          "| c4_tiles", round(d["c4_tiles"]["value"], 1) if d.get("c4_tiles") else None)
except Exception as e:
    print("$W: no line", e)
PY
done
