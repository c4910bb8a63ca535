#!/bin/bash
# the headline line (C3 + the c4_tiles record) at N GPUs: [GPUS=N] tools/gpu.sh 600 'bash tools/r2_c3_line.sh N'
N=${1:-1}
O=gpurun_out; mkdir -p $O
if [ "$N" = "1" ]; then python bench.py --gpus 1 --steps 4 --warmup 3 --no-cpu-baseline 2> $O/r2y_n$N.err | grep "^{" > $O/r2_bench_C3_n$N.json
else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 4 --warmup 3 --no-cpu-baseline 2> $O/r2y_n$N.err | grep "^{" > $O/r2_bench_C3_n$N.json; fi
python - <<PY
import json
d=json.loads(open("$O/r2_bench_C3_n$N.json").read().strip().splitlines()[-1])
c=d["c4_tiles"]
print("N=$N C3", round(d["value"],1), round(d["ms_per_step"],2), "| c4_tiles", round(c["value"],1), round(c["ms_per_step"],3), c["kernel"], "|", c["kernel_choice"])
PY
