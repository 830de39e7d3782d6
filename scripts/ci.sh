#!/bin/bash
# What CI runs (also .github/workflows/ci.yml): the reference gates lint + format before the build
# (.travis.yml:36-39) and then only configures; here the build and the CPU test suite run too.
set -euo pipefail
cd "$(dirname "$0")/.."
python scripts/lint.py
python scripts/format_check.py -o check
python build.py --host-only
python -m pytest tests -x -q -m "not gpu"
if [ "${MINIPS_CI_SANITIZE:-0}" = "1" ]; then
  python tools/sanitize.py host --tool thread
  python tools/sanitize.py host --tool address
fi
