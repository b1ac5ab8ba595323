#!/bin/bash
# 2 GPUs: the sharded bench over the real process group — peer-memory transport (default) against the NCCL baseline,
# and the C4 sub-sample cut into 2 shards (strong scaling, parity against the unmodified reference's digest).
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
N=${1:-2}
run() {  # name, extra args...
  name=$1; shift
  ( timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $N --steps 20 --warmup 5 "$@" ) > gpurun_out/n${N}_${name}.json 2> gpurun_out/n${N}_${name}.err
  echo "$name rc=$?" >> gpurun_out/rc_n${N}.txt
}
: > gpurun_out/rc_n${N}.txt
RUNS=${2:-"peer nccl c4"}
for r in $RUNS; do
  case $r in
    peer) run peer ;;
    nccl) run nccl --exchange nccl ;;
    c4) run c4 --workload C4 ;;
    c5) run c5 --workload C5 ;;
  esac
done
cat gpurun_out/rc_n${N}.txt
python - $N <<'PY'
import json, sys
N = sys.argv[1]
for k in ("peer", "nccl", "c4", "c5"):
    try:
        d = json.loads(open("gpurun_out/n%s_%s.json" % (N, k)).read().strip().splitlines()[-1])
        print(k, "ms/step", round(d["ms_per_step"], 3), "rec/s", round(d["value"] / 1e6, 1), "M launches", d["gpu_launches"], "e2e ms", round(d["e2e"]["ms_per_step"], 1),
              "parity", d["parity_ok"], "replayed", d["device_pass"]["steps_without_intermediate_sync"], d["scaling"])
        print("   ", {a: round(b, 3) for a, b in d.get("sharded_stage_ms_per_step", {}).items()})
        print("   ", {a: round(b, 3) for a, b in d.get("fetch_exchange_ms_per_step", {}).items()})
        print("   ", {a: round(b, 3) for a, b in d["stage_ms_per_step"].items()})
        if d["parity_ok"] is not True: print("   parity:", json.dumps(d.get("parity"))[:1500])
    except FileNotFoundError:
        pass
    except Exception as e:
        print(k, "parse failed", e); print(open("gpurun_out/n%s_%s.err" % (N, k)).read()[-2500:])
PY
