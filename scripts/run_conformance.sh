#!/usr/bin/env bash
# Run the reference's own test-suite, unchanged, against THIS package.
# The reference's pyproject would put /root/reference/src first on sys.path, so we point pytest
# at our ini file / rootdir and export PYTHONPATH for the subprocess-based CLI tests.
set -euo pipefail
REPO="$(cd "$(dirname "$0")/.." && pwd)"
REF="${1:-/root/reference}"
WORK="$(mktemp -d)"
cd "$WORK"
PYTHONPATH="$REPO/src" python -m pytest -c "$REPO/pyproject.toml" --rootdir "$REPO" \
  "$REF/tests" -p no:cacheprovider -q "${@:2}"
