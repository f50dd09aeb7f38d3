#!/bin/bash
# Run V-Gate's own test-suite (unmodified, from /root/reference) with engine_type "b200" registered,
# as the regression gate for the control plane around the new backend.  CPU only (dry-run backends).
set -e
REPO="$(cd "$(dirname "$0")/.." && pwd)"
REF="${1:-/root/reference}"
cd "$REF"
cat > /tmp/b200_ref_conftest_plugin.py <<'PY'
import vgate_b200.register  # noqa: F401  registers engine_type "b200" before the tests import vgate
PY
VGATE_DRY_RUN=true PYTHONPATH="$REPO/tools/shims:$REPO:/tmp:$REF" \
  python -m pytest tests/ -q -p no:cacheprovider -p b200_ref_conftest_plugin "${@:2}"
