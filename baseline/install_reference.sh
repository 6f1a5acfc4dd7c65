#!/bin/bash
# Installs the UNMODIFIED reference (Crystal-wxy/grace) into the git-ignored baseline/_ref.
# 1) the prescribed offline pip install (from a copy: the build writes egg-info/build dirs and
#    /root/reference is read-only);
# 2) the reference's setup.py lists only packages=['grace_dl'] (setup.py:9),
#    so pip installs grace_dl/__init__.py alone; the sub-packages the benchmark needs
#    (grace_dl.dist.*) are completed by copying the reference's own files, byte for byte.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${1:-/root/reference}"
rm -rf /tmp/_grace_ref_src && cp -r "$REF" /tmp/_grace_ref_src
python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target "$HERE/_ref" --upgrade /tmp/_grace_ref_src
for sub in dist torch; do
  cp -r "$REF/grace_dl/$sub" "$HERE/_ref/grace_dl/"
done
find "$HERE/_ref" -name __pycache__ -type d -exec rm -rf {} +
echo "reference installed in $HERE/_ref"
