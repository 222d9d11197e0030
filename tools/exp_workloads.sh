#!/bin/bash
# A/B of the bounce-0 packet trace (B2_PRIMARY_PACKET = rays per lane, 0 = per-ray kernel) on the recorded workloads (GPU box)
for w in "$@"; do
  for pk in 2 1 0; do
    echo "== $w packet=$pk: $(B2_PRIMARY_PACKET=$pk python tools/profile_render.py $w 2>&1 | grep 'launch 2')"
  done
done
