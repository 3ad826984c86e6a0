#!/bin/bash
# compute-sanitizer is CLOSED on this GPU pool by its operators and stays closed
# (profiles/r02_compute_sanitizer_closed.txt: "runs under it have left GPUs needing a reset").
# This script used to submit seven memcheck / racecheck runs; it now only says so, so nobody re-submits a closed command.
# What stands in for it:
#   * bounded-spin traps in every kernel that waits on a flag or barrier (a hang becomes a launch error),
#   * tools/tp_check.py --stress  (102 400 exchanges x {pull, NVLS} x {2, 4, 8} ranks, checksummed; tools/gpu_multi.sh),
#   * the bit-exact / oracle parity tests under tests/ at small shapes.
echo "tools/gpu_sanitize.sh: compute-sanitizer is closed on this pool (see profiles/r02_compute_sanitizer_closed.txt)." >&2
echo "Use tools/tp_check.py --stress and the parity tests instead; nothing was run." >&2
exit 64
