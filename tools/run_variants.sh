#!/bin/bash
# profiling aid: strip-0 time per time step for build variants of the FOM kernel
for f in tools/var_*.so; do echo "== $f"; BFD_LIB=$PWD/$f timeout 120 python tools/trace_pipeline.py --batch 1 --steps 26 2>&1 | tail -3 | head -2; done
