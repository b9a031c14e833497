#!/bin/bash
# Same usage as the reference's generate.sh (generate.sh:1-13 there): prints <height> lines of
# <width> random 0/1 characters to stdout.  Backed by bin/gol-generate: seeded (third argument,
# default: random), reproducible, and gigabytes per second instead of one `echo` per cell.
#   ./generate.sh 840 840 > input.txt
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
[ -x "$here/bin/gol-generate" ] || make -C "$here" bin/gol-generate >/dev/null
exec "$here/bin/gol-generate" "$1" "$2" "${3:-$RANDOM$RANDOM}"
