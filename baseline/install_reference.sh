#!/usr/bin/env bash
# Install the UNMODIFIED reference (read-only at /root/reference) into baseline/_ref.
# The reference ships no setup.py/pyproject, so pip cannot build it in place: we copy the tree
# to /tmp, add a packaging shim (setup.py listing the lab modules, nothing else) and install the
# copy with the offline command the task prescribes.  Every installed .py is byte-identical to
# the reference's (checked at the end).
set -euo pipefail
REPO="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
REF="${1:-/root/reference}"
TMP="$(mktemp -d /tmp/ddl25_ref_XXXX)"
cp -r "$REF"/. "$TMP"/
cat > "$TMP/setup.py" <<'PY'
from setuptools import setup
setup(
    name="ddl25spring-reference", version="0.0.0", package_dir={"": "."},
    py_modules=[
        "lab.tutorial_1a.hfl_complete", "lab.tutorial_2a.centralized", "lab.tutorial_2a.generative-modeling",
        "lab.tutorial_2b.vfl", "lab.tutorial_2b.exercise_1", "lab.tutorial_2b.exercise_2",
        "lab.tutorial_2b.exercise_3", "lab.tutorial_1b.primer.intro",
        "lab.tutorial_1b.DP.gradient_aggr.intro_DP_GA", "lab.tutorial_1b.DP.weight_aggr.intro_DP_WA",
    ],
)
PY
rm -rf "$REPO/baseline/_ref"
python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --no-deps \
    --target "$REPO/baseline/_ref" "$TMP" >/dev/null
cmp "$REPO/baseline/_ref/lab/tutorial_1a/hfl_complete.py" "$REF/lab/tutorial_1a/hfl_complete.py"
cp "$REF/lab/tutorial_2a/heart.csv" "$REPO/baseline/_ref/lab/tutorial_2a/heart.csv" 2>/dev/null || true
rm -rf "$TMP"
echo "reference installed into $REPO/baseline/_ref"
