#!/bin/bash
# usage: tools/gpu_proj.sh <tag> [pytest -k expr]  -- projection GPU tests under a hard timeout; retries while the pod is busy
TAG=$1; K=${2:-}
KARG=""; if [ -n "$K" ]; then KARG="-k '$K'"; fi
CMD="timeout -s KILL 400 python -m pytest tests/test_gpu_project.py -x -q -m gpu -s $KARG > gpurun_out/${TAG}_pytest.log 2>&1; echo \"pytest rc=\$?\"; tail -40 gpurun_out/${TAG}_pytest.log"
for try in 1 2 3 4 5 6; do
  /usr/local/graft/bin/gpurun --timeout 600 -- "$CMD" > /tmp/gpu_proj_$TAG.log 2>&1
  if grep -q "status=transient" /tmp/gpu_proj_$TAG.log; then sleep 120; continue; fi
  break
done
tail -60 /tmp/gpu_proj_$TAG.log
