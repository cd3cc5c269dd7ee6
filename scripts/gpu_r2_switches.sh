#!/bin/bash
# GPU suite with the documented switches flipped (each must stay green)
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
for sw in R3DGW_SPLIT_FUSED_BACKWARD=0 R3DGW_DEFERRED_BACKWARD=0 R3DGW_ALPHA_CACHE=0 R3DGW_GEOMETRY_CACHE=0 R3DGW_GEOM_BWD_F32=1 R3DGW_GRAD_PAD=0; do
  echo "== $sw"; env $sw timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -2
done
