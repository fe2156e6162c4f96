#!/bin/bash
# One-off probe (VERDICT r01 / SURVEY F-note): is a MinkowskiEngine / mmcv / mmdet install reachable on the GPU box?
mkdir -p gpurun_out
{
  for m in MinkowskiEngine mmcv mmdet mmdet3d spconv; do
    python - <<P
import importlib
try:
    m = importlib.import_module("$m"); print("$m: importable", getattr(m, "__version__", "?"), getattr(m, "__file__", "?"))
except Exception as e:
    print("$m: NOT importable (%s: %s)" % (type(e).__name__, e))
P
  done
  echo "baseline/_ref:"; ls -la baseline/_ref 2>&1 | head
  echo "pip list | grep -i minkowski:"; python -m pip list 2>/dev/null | grep -i -E "minkowski|mmcv|mmdet|spconv" || echo "(none)"
  find / -iname "*minkowski*" -not -path "*/proc/*" 2>/dev/null | grep -v scipy | head
  nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv
  nproc
} > gpurun_out/me_probe.txt 2>&1
cat gpurun_out/me_probe.txt
