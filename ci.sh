#!/usr/bin/env bash
# CI entry point (counterpart of the reference's `./gradlew clean check`): build the sm_100a
# kernel library (nvcc cross-compiles without a GPU), byte-compile every module, run the CPU tiers.
set -euo pipefail
cd "$(dirname "$0")"
python -m tony_b200.csrc.build --selftest
python -m compileall -q tony_b200 examples tests bench.py __graft_entry__.py
python -m pytest tests -x -q -m "not gpu"
if python -c "import torch,sys; sys.exit(0 if torch.cuda.is_available() else 1)"; then
    python -m pytest tests -x -q -m gpu
fi
