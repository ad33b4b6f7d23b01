#!/bin/bash
# sweep the persistent kernel's refetch / leaf thresholds on the C4 ray sets (tools/trace_stats.py)
for rb in ${RBS:-12 28 31}; do for lt in ${LTS:-1 2 4 8 16}; do
  echo "== refetch_below=$rb leaf_threshold=$lt"
  MTSGPU_REFETCH_BELOW=$rb MTSGPU_LEAF_THRESHOLD=$lt python tools/trace_stats.py c4 2048 2>&1 | grep -E "closest|any"
done; done
