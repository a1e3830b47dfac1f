#!/bin/bash
# Quick parity pass for new items: pytest -k "$1" (default: everything GPU), no bench.
mkdir -p gpurun_out
K="${1:-}"
if [ -n "$K" ]; then
  timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -x -k "$K" > gpurun_out/pytest_quick.log 2>&1
else
  timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/pytest_quick.log 2>&1
fi
echo "pytest exit $?" >> gpurun_out/pytest_quick.log
tail -15 gpurun_out/pytest_quick.log
