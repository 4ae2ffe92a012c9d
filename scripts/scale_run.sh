#!/bin/bash
# scripts/scale_run.sh N TAG : the three multi-GPU measurements on N GPUs of one box (default image-sharded, batch, candidate-sharded field)
N=$1; TAG=$2
run() { if [ "$N" = "1" ]; then python bench.py --gpus 1 "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@"; fi; }
run --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_default_n$N.json 2> gpurun_out/${TAG}_default_n$N.err
run --workload batch --images 8 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/${TAG}_batch_n$N.json 2> gpurun_out/${TAG}_batch_n$N.err
run --workload tiled4k --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_tiled4k_n$N.json 2> gpurun_out/${TAG}_tiled4k_n$N.err
for w in default batch tiled4k; do python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/${TAG}_${w}_n$N.json").read().splitlines() if l.startswith("{")][-1])
    print("$w", "N=$N", "value", round(d["value"]), "ms", round(d["ms_per_step"], 1), "e2e", round(d["e2e"]["value"]), "img/s", round(d["stats"]["images_per_s"], 2), "e2e img/s", round(d["stats"]["e2e_images_per_s"], 2), d["scaling"])
except Exception as e:
    print("$w", "N=$N", "FAILED", e); print(open("gpurun_out/${TAG}_${w}_n$N.err").read()[-1500:])
PY
done
