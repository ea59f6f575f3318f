#!/bin/bash
# Regenerates the committed fixtures.  Needs /root/reference (not present on the GPU box).
#   tsp100.txt           one output of the reference's instance generator (test3/gen.c),
#                        compiled as-is; it seeds from the clock, so each run differs --
#                        the committed file is the pinned instance.
#   oracle_digests.json  digests of oracle trajectories (tests/golden/make_digests.py)
set -e
cd "$(dirname "$0")"
gcc -O2 -o /tmp/pga_gen /root/reference/test3/gen.c
/tmp/pga_gen > tsp100.txt
python make_digests.py
