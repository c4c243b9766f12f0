#!/bin/bash
# profiling aid: strip pipeline timing for build variants of the FOM kernel
for f in tools/var_*.so; do for b in ${BATCHES:-1}; do echo "== $f batch $b"; BFD_LIB=$PWD/$f timeout 120 python tools/trace_pipeline.py --batch $b --steps 26 2>&1 | grep -E "^run|^3 |^23 |rror"; done; done
