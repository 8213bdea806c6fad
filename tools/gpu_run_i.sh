#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity_big.py -m gpu -x -q -k many_small_batches 2>&1 | tail -3
echo "--- with the build that lacked the barrier"
timeout 600 python - <<'PY' 2>&1 | tail -4
import importlib, os, sys
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
abi = importlib.import_module("svt-av1-mod-by-patman_b200.abi")
abi.LIB_PATH = os.path.join(os.getcwd(), "build_alt", "libsvtme_5d40bb1.so")
import pytest
sys.exit(pytest.main(["tests/test_gpu_parity_big.py", "-m", "gpu", "-x", "-q", "-k", "many_small_batches"]))
PY
