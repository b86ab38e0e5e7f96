#!/usr/bin/env bash
# Stage the UNMODIFIED reference under baseline/_ref (git-ignored, travels to the GPU box with gpurun).
#
#   1. the offline pip install the task prescribes — fails: the reference is 11 flat scripts with neither
#      setup.py nor pyproject.toml ("Directory '/root/reference' is not installable");
#   2. fallback: a byte-for-byte copy of its src/ tree plus a SHA-256 manifest, so baseline/reference_arm.py can
#      prove at run time that what it executes is the reference's own text.
set -u
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
REF="${1:-/root/reference}"
DST="$ROOT/baseline/_ref"
mkdir -p "$DST"
python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target "$DST" "$REF" \
    > "$DST/pip_install.log" 2>&1
echo "pip exit code: $?" >> "$DST/pip_install.log"
rm -rf "$DST/src"
cp -r "$REF/src" "$DST/src"
( cd "$DST" && find src -type f -name '*.py' | sort | xargs sha256sum > SHA256SUMS )
( cd "$REF" && find src -type f -name '*.py' | sort | xargs sha256sum ) | diff - "$DST/SHA256SUMS" \
    && echo "staged $(wc -l < "$DST/SHA256SUMS") unmodified reference scripts under $DST/src"
