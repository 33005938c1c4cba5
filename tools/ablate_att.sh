#!/bin/bash
# Rebuilds the library with one experiment macro at a time and times the 64x64-level self-attention (needs nvcc + GPU).
for v in "" "$@"; do
  B200SD_NVCC_EXTRA="$v" python -c "import importlib;importlib.import_module('k8s-nvidia-gpus_b200.build').build()" >/dev/null
  echo "== variant '$v'"
  B200SD_NVCC_EXTRA="$v" python tools/profile_step.py att
done
