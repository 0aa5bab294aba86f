#!/bin/bash
# build the library; exit non-zero (and show the compiler errors) when it fails
cd "$(dirname "$0")/.." && python - <<'PY'
import importlib, sys
sys.path.insert(0, '.')
b = importlib.import_module('2handedafforder_b200._build')
try:
    print(b.build_library(verbose=False))
except Exception as e:
    msg = str(e)
    print(msg[:3000])
    sys.exit(1)
PY
