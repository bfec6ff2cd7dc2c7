#!/usr/bin/env bash
# compute-sanitizer over every native kernel (tiny shapes, bench/sanitize_probe.py).
#   scripts/sanitize.sh [memcheck|racecheck|synccheck|initcheck]      (default: memcheck)
# Needs one B200; run the plain probe first (a fault under the sanitizer can wedge the GPU):
#   python bench/sanitize_probe.py && scripts/sanitize.sh memcheck
set -euo pipefail
tool="${1:-memcheck}"
cd "$(dirname "$0")/.."
exec compute-sanitizer --tool "$tool" --error-exitcode 3 --launch-timeout 120 \
     python bench/sanitize_probe.py
