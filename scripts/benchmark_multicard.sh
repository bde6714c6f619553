#!/bin/bash
# Weak-scaling sweep of the headline benchmark at 1/2/4/8 GPUs, both arms (reference: scripts/benchmark_multicard.sh).
set -uo pipefail
cd "$(dirname "$0")/.."
source scripts/torch_dist/set_env.sh
mkdir -p results
for N in 1 2 4 8; do
  for IMPL in reference ours; do
    if [ "$N" -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29400 + N))"; fi
    $L bench.py --gpus "$N" --steps "${STEPS:-20}" --warmup "${WARMUP:-5}" --impl "$IMPL" "$@" 2> "results/multicard_${IMPL}_${N}.err" | tail -1 | tee "results/multicard_${IMPL}_${N}.json"
  done
done
python - <<'PY'
import json, glob
for impl in ("reference", "ours"):
    base = None
    for n in (1, 2, 4, 8):
        try:
            d = json.load(open(f"results/multicard_{impl}_{n}.json"))
        except Exception:
            continue
        if "value" not in d:
            continue
        base = base or d["value"]
        print(f"{impl:9s} N={n}: {d['value']:>12,.0f} tok/s  {d['ms_per_step']:.2f} ms/step  efficiency {d['value'] / (n * base) * 100:.1f} %")
PY
