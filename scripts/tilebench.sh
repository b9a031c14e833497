#!/bin/bash
# Runs every variant of bin/gol-tilebench in its own process: scripts/tilebench.sh [W] [H] [reps] [chunk_rows]
W=${1:-16384}; H=${2:-16384}; REPS=${3:-20}; CR=${4:-85}
bin/gol-tilebench $W $H $REPS -1 $CR
for v in $(seq 0 16); do
  timeout 60 bin/gol-tilebench $W $H $REPS $v $CR 2>&1 || echo "  variant $v: failed (exit $?)"
done
