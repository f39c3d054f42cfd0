#!/bin/bash
# Installs the UNMODIFIED reference (yanjunzan/GMAT) into baseline/_ref with pip, offline, from a scratch copy of
# /root/reference (its setup.py builds two cffi modules into the source tree, which is read-only).  baseline/_ref is
# git-ignored and travels to the GPU box with the repo snapshot; nothing under gmat_b200/ ever imports it.
#   CC/LDSHARED: the image's default /opt/gcc has no libgomp.spec, the reference's C needs -fopenmp -> system gcc.
#   --no-deps:   pandas_plink (install_requires) is not in the offline wheelhouse; baseline/reference_arm.py supplies
#                a stand-in that reads the .bed with the reference's OWN C reader (gmat.process_plink.read_plink_bed).
set -e
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${1:-/root/reference}"
if [ ! -d "$REF/gmat" ]; then
    echo "reference tree absent; keeping prebuilt baseline/_ref (if any)"; exit 0
fi
if [ -f "$HERE/_ref/gmat/remma/remma_epiAA/remma_epiAA.py" ] && ls "$HERE"/_ref/_cremma_epi_eff_cpu*.so > /dev/null 2>&1; then
    echo "baseline/_ref already installed"; exit 0
fi
TMP="$(mktemp -d)"
cp -r "$REF" "$TMP/src"
CC=/usr/bin/gcc CXX=/usr/bin/g++ LDSHARED="/usr/bin/gcc -shared" python -m pip install --no-index \
    --no-build-isolation --find-links /opt/wheelhouse --no-deps --upgrade --target "$HERE/_ref" "$TMP/src" \
    > "$TMP/pip.log" 2>&1 || { tail -30 "$TMP/pip.log"; exit 1; }
rm -rf "$TMP"
echo "installed the unmodified reference into baseline/_ref"
