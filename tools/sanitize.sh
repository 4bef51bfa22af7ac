#!/bin/bash
# compute-sanitizer pass over one tiny training step.  memcheck by default; TOOL=synccheck|racecheck.
# (The GPU pool of this project refuses compute-sanitizer - "closed on this pool" - so this is for other boxes.)
#   bash tools/sanitize.sh            (one B200; takes about a minute)
TOOL=${TOOL:-memcheck}
timeout ${TIMEOUT:-240} compute-sanitizer --tool $TOOL --launch-timeout 0 --error-exitcode 7 --print-limit 20 \
  python tools/sanitize_target.py 2>&1 | grep -av "^\[W" | tail -${LINES:-25}
echo "sanitizer exit code: ${PIPESTATUS[0]}"
