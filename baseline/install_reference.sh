#!/bin/bash
# "Installs" the reference into baseline/_ref (git-ignored).  pip cannot do it: the reference is a script
# collection without setup.py / pyproject.toml (`pip install --no-index --no-build-isolation --find-links
# /opt/wheelhouse --target baseline/_ref /root/reference` -> "Neither 'setup.py' nor 'pyproject.toml' found").
# The tree is therefore vendored by plain copy, byte for byte, and used unmodified by `bench.py --impl reference`.
set -e
here="$(cd "$(dirname "$0")" && pwd)"
src="${1:-/root/reference}"
rm -rf "$here/_ref"
mkdir -p "$here/_ref"
cp -r "$src"/. "$here/_ref"/
echo "reference copied to $here/_ref"
