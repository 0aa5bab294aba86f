#!/bin/bash
# 2-GPU check of the frame-sharded clip path: bench line at N=2 (NCCL gather inside the step)
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${1}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/${1}_pytest.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/${1}_bench_n1.json 2> $O/${1}_bench_n1.err; echo "n1 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > $O/${1}_bench_n2.json 2> $O/${1}_bench_n2.err; echo "n2 rc=$?"
tail -5 $O/${1}_bench_n2.err
python - <<PY
import json
for n in (1,2):
    try:
        d=json.loads(open("$O/${1}_bench_n%d.json"%n).read().strip().splitlines()[-1])
        print(n, round(d["value"],1), round(d["ms_per_step"],2), round(d["e2e"]["value"],1), d["clocks"]["sm_mhz"])
    except Exception as e: print(n,"ERR",e)
PY
