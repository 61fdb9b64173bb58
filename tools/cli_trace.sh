#!/bin/bash
# diagnostic: mg_gcn --eager -P 2 with host-side enqueue tracing on a small generated dataset
set -e
cd "$(dirname "$0")/.."
D=$(mktemp -d)
python - "$D" <<'PY'
import sys
sys.path.insert(0, ".")
from __graft_entry__ import load_package
mg = load_package(); mg.lib()
A, X, Y, C = mg.generate("reddit", seed=3, scale=0.02)
mg.write_dataset(sys.argv[1] + "/permuted/reddit", A, X, Y)
PY
mkdir -p "$D/csvs"; cd "$D"
MGGCN_TRACE=1 MGGCN_BARRIER_TIMEOUT_MS=2000 timeout 120 "$OLDPWD/mg-gcn_b200/cpp/bin/mg_gcn" "$@" -P 2 -R 1 -E 2 train "$D/permuted/reddit/" 1 64 2>&1 | head -150
