#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
for sw in R3DGW_NOP=1 R3DGW_GEOMETRY_CACHE=0 R3DGW_GEOM_BWD_F32=1; do
  echo "== $sw"; env $sw timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | grep "^E  \|^FAILED\|passed\|failed\|retry" | head -6 | cut -c1-300
done
