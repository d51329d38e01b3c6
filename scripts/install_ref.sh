#!/bin/bash
# Installs the UNMODIFIED reference packages into baseline/_ref/ (git-ignored, NOT gpurun-ignored: it travels to the
# GPU box like the built .so).  Used only by `bench.py --impl reference`, the `reference_gpu` probe and the
# "scorer-only" adoption tests (the reference's own BatchBeamSearch driving our CTCPrefixScorer).
#
# First choice is the contract's pip install; in this image it fails while generating metadata (setup.py has
# setup_requires=["pytest-runner"], which the offline wheelhouse does not hold), so the two python packages the hot
# path lives in (espnet/, espnet2/) are copied as they are.  Nothing is edited.
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
REF="${1:-/root/reference}"
DST="$ROOT/baseline/_ref"
[ -d "$REF/espnet" ] || { echo "no reference at $REF"; exit 1; }
rm -rf "$DST"; mkdir -p "$DST"
if [ -z "$LGASR_REF_COPY_ONLY" ]; then
  TMP=$(mktemp -d); cp -r "$REF" "$TMP/src"
  if (cd "$TMP" && python -m pip install -q --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
        --target "$DST" "$TMP/src" > "$TMP/pip.log" 2>&1); then
    echo "pip install ok" > "$DST/INSTALL_OUTCOME.txt"; rm -rf "$TMP"; exit 0
  fi
  tail -3 "$TMP/pip.log" | sed 's/^/  pip: /'
  rm -rf "$TMP"; rm -rf "$DST"; mkdir -p "$DST"
fi
cp -r "$REF/espnet" "$REF/espnet2" "$DST/"
find "$DST" -name __pycache__ -type d -prune -exec rm -rf {} +
echo "pip install failed (setup_requires pytest-runner not in the wheelhouse): espnet/ and espnet2/ copied unmodified from $REF" > "$DST/INSTALL_OUTCOME.txt"
echo "installed: $(du -sh "$DST" | cut -f1) in $DST"
