#!/bin/bash
# Environment check + native build (the reference's install.sh:1-10 installs numpy, PyTorch (cudatoolkit 11.3), pandas,
# tqdm with conda/pip; on an offline B200 image those come with the image, so this script verifies them and builds the
# sm_100a library instead of downloading anything).
set -e
cd "$(dirname "$0")"
python - <<'PY'
import importlib
missing = []
for mod in ["numpy", "torch", "pandas", "tqdm", "PIL"]:
    try:
        m = importlib.import_module(mod)
        print(f"{mod:8s} {getattr(m, '__version__', '?')}")
    except ImportError:
        missing.append(mod)
if missing:
    raise SystemExit(f"missing python packages: {missing} (install them from your package index: pip3 install numpy pandas tqdm pillow torch)")
PY
echo "Building libdpt_b200.so (sm_100a)"
make -j"$(nproc)"
python -c "import __graft_entry__ as g; g.build()"
