#!/usr/bin/env bash
# Everything that can be checked without a GPU: clean build for sm_100a, symbol resolution, the
# CPU test suite (gloo, world size 2 and 4), SASS / ptxas tables regenerated from the build.
set -euo pipefail
cd "$(dirname "$0")/.."
rm -rf build distributed_training_dl_b200/lib
python -c "import __graft_entry__ as g; g.build()"
python -m pytest tests -x -q -m "not gpu"
python tools/sass_table.py
git status --short profiles/sass | sed 's/^/changed: /' || true
