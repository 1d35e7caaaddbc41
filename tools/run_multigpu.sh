#!/bin/bash
# Multi-GPU measurement of BASELINE.json configs[1..4] on one box: tools/run_multigpu.sh <N> <tag>
# (launched through gpurun --gpus N; one process per GPU, torch.distributed.run as the driver does it)
N=$1; TAG=$2; PORT=29500
run() { out=$1; shift
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" \
    > gpurun_out/${out}.json 2> gpurun_out/${out}.err; echo "$out rc=$?"; PORT=$((PORT+1)); }
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/${TAG}_topo_n${N}.txt 2>&1
if [ "$N" = "8" ]; then
  run ${TAG}_caffenet_n8 --steps 500 --warmup 150 --tau-sweep 1,10,50,200          # configs[3] + configs[4]
  run ${TAG}_cifar10_full_n8 --workload cifar10_full --steps 500 --warmup 150      # configs[2]
  run ${TAG}_cifar10_quick_n8 --workload cifar10_quick --steps 200 --warmup 30     # configs[1]
else
  run ${TAG}_cifar10_quick_n${N} --workload cifar10_quick --steps 200 --warmup 30  # configs[1]
  run ${TAG}_caffenet_n${N} --steps 200 --warmup 50
fi
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/*_n[0-9].json")):
    try:
        b = json.load(open(f))
        a = b["averaging"]
        print(f, "value %.0f e2e %.0f e2e_u8 %s avg share %.4f exch %.3f ms bit_exact %s" % (b["value"], b["e2e"]["value"],
              b["e2e_u8"]["value"] if b.get("e2e_u8") else None, a["share_of_round"], a["exchange_ms"], a.get("bit_exact")))
    except Exception as e:
        print(f, "unreadable:", e)
PY
