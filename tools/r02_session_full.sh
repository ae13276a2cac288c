#!/bin/bash
# full GPU suite + smoke at the current tree: tools/r02_session_full.sh <tag>
OUT=gpurun_out/${1:-r02full}; mkdir -p $OUT
timeout 2400 python -m pytest tests/ -m gpu -x -q > $OUT/pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.txt; tail -4 $OUT/pytest_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke OK')" > $OUT/smoke.txt 2>&1; tail -2 $OUT/smoke.txt
timeout 300 python tools/dev_small.py > $OUT/dev_small.jsonl 2>&1; cut -c1-160 $OUT/dev_small.jsonl
