#!/bin/bash
# usage: tools/sweep.sh <workload> <spp> -- env-var sweeps of the wavefront tuning knobs (GPU box)
wl=${1:-bedroom}; spp=${2:-16}
for s in 1 2 3 4 8 16; do echo "slice=$s"; B2_TRACE_SLICE=$s python tools/profile_render.py $wl 3 $spp 0 2>&1 | grep "launch 2"; done
for b in 1 2; do echo "blocks=$b"; B2_TRACE_BLOCKS=$b python tools/profile_render.py $wl 3 $spp 0 2>&1 | grep "launch 2"; done
for c in 1048576 4194304 16777216; do echo "chunk=$c"; python tools/profile_render.py $wl 3 $spp 0 $c 2>&1 | grep "launch 2"; done
