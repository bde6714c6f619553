#!/bin/bash
# Wrapper around benchmark_comprehensive.py (reference: scripts/benchmark_comprehensive.sh): environment defaults,
# log directory, optional reference arm.    usage: scripts/benchmark_comprehensive.sh [GPUS] [FILTER] [IMPL]
set -euo pipefail
cd "$(dirname "$0")/.."
GPUS=${1:-8}; FILTER=${2:-}; IMPL=${3:-ours}
source scripts/torch_dist/set_env.sh
STAMP=$(date +%Y%m%d_%H%M%S)
OUT=results/bench_comprehensive_${STAMP}.json
mkdir -p results
python -c "import scaletorch_b200._build as b; b.build()"        # make sure the sm_100a library is current
python scripts/benchmark_comprehensive.py --gpus "$GPUS" --filter "$FILTER" --impl "$IMPL" --out "$OUT" 2>&1 | tee "results/bench_comprehensive_${STAMP}.log"
echo "results: $OUT  ${OUT%.json}.md"
