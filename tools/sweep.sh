#!/bin/bash
# usage: tools/sweep.sh <workload> <spp> -- env-var sweeps of the wavefront tuning knobs (GPU box)
wl=${1:-bedroom}; spp=${2:-16}
for t in 4 8 12 16 20 24; do echo "tri_thresh=$t"; B2_TRI_THRESH=$t python tools/profile_render.py $wl 3 $spp 0 2>&1 | grep "launch 2"; done
for s in 4 6 8 12 16 32; do echo "slice=$s"; B2_TRACE_SLICE=$s python tools/profile_render.py $wl 3 $spp 0 2>&1 | grep "launch 2"; done
