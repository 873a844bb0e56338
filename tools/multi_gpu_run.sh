#!/bin/bash
# tools/multi_gpu_run.sh TAG N [N ...] — on a multi-GPU box: the N-rank PCIe probe and the bench line for every N given (torchrun),
# and the 2-rank NCCL exchange test; everything lands in gpurun_out/TAG_*
TAG=$1; shift
python -m pytest tests/test_gpu_sharded_nccl.py -q > gpurun_out/${TAG}_pytest_nccl.log 2>&1; tail -2 gpurun_out/${TAG}_pytest_nccl.log
python tools/pcie_probe.py > gpurun_out/${TAG}_pcie_n1.json 2>gpurun_out/${TAG}_pcie_n1.err; cat gpurun_out/${TAG}_pcie_n1.json
for N in "$@"; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/pcie_probe.py > gpurun_out/${TAG}_pcie_n${N}.json 2>gpurun_out/${TAG}_pcie_n${N}.err; cat gpurun_out/${TAG}_pcie_n${N}.json
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_n${N}.json 2>gpurun_out/${TAG}_bench_n${N}.err
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/${TAG}_bench_n${N}.json") if l.startswith("{")][-1]
    print("N=${N} value", round(d["value"]), "sustained", round(d["sustained"]["value"]), "e2e", round(d["e2e"]["value"]), "host bound", round(d["e2e"]["host_bound_frames_per_s"] or 0),
          "cfg5", round(d["match"]["gpairs_per_s"],1), "ms", round(d["match"]["ms"],3), d["match"].get("exchange_check"))
except Exception as e:
    print("N=${N} failed", e, open("gpurun_out/${TAG}_bench_n${N}.err").read()[-1500:])
PY
done
