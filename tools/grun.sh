#!/bin/bash
# build the library here (nvcc cross-compiles), then run a command on the GPU box: tools/grun.sh [gpurun flags] -- 'cmd'
set -e
cd "$(dirname "$0")/.."
python -c "import __graft_entry__ as g; g.build()" >&2
exec /usr/local/graft/bin/gpurun "$@"
