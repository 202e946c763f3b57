#!/bin/bash
# Installs the UNMODIFIED reference (if001/lit-llama-ja) into baseline/_ref (git-ignored; it travels to the GPU box with
# the gpurun snapshot).  The reference tree is read-only and its setup.py lists git dependencies, so the wheel is built from
# a copy under /tmp with --no-deps.  generate.py is a top-level script outside the package: it is copied next to it as
# ref_generate.py so that bench.py --impl reference can time the reference's own generate() loop.
#   bash baseline/install_ref.sh          (run where /root/reference exists)
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
SRC=${REFERENCE_SRC:-/root/reference}
[ -d "$SRC" ] || { echo "no reference tree at $SRC"; exit 1; }
TMP=$(mktemp -d)
cp -r "$SRC" "$TMP/ref"
rm -rf "$HERE/_ref"
python -m pip install -q --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target "$HERE/_ref" "$TMP/ref"
cp "$SRC/generate.py" "$HERE/_ref/ref_generate.py"
rm -rf "$TMP"
ls "$HERE/_ref"
