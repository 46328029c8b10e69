#!/bin/bash
# On an 8-GPU box: host-link ceiling (tools/pcie_probe.py) and bench.py at N = 8 and 4.  Everything lands in gpurun_out/.
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi topo -m > $O/topo_8gpu.txt 2>&1
for n in 8 4; do
  $TR --nproc-per-node $n --master-port $((29600+n)) tools/pcie_probe.py 2>/dev/null | tail -n 1 > $O/pcie_${n}gpu.json
done
python tools/pcie_probe.py 2>/dev/null | tail -n 1 > $O/pcie_1gpu.json
for n in 8 4; do
  $TR --nproc-per-node $n --master-port $((29700+n)) bench.py --gpus $n --steps 10 --warmup 3 2>$O/bench_${n}gpu.err | tail -n 1 > $O/bench_${n}gpu.json
done
timeout 300 python -m pytest tests/test_gpu_multi.py -q 2>&1 | tail -n 3 > $O/multi_tests.log
cat $O/multi_tests.log
python - <<'PY'
import json
for n in (1,4,8):
    try:
        d=json.load(open("gpurun_out/pcie_%dgpu.json"%n))["sizes"]["835_MB"]
        print(n,"GPUs per-GPU GB/s d2h %.1f h2d %.1f both %.1f"%tuple(d[k]["GBps_per_gpu_per_direction"] for k in ("d2h","h2d","both")))
    except Exception as e: print(n,"probe failed",e)
for n in (4,8):
    try:
        d=json.load(open("gpurun_out/bench_%dgpu.json"%n))
        print(n,"GPUs value %.0f e2e %.0f variants %s parity %s gather %s"%(d["value"],d["e2e"]["value"],{k:round(v["value"]) for k,v in d["e2e"]["variants"].items()},d["parity"]["per_rank"],d["gathered_cloud"]))
    except Exception as e: print(n,"bench failed",e)
PY
