#!/bin/bash
# build (if stale) and run a command on the B200 box:  tools/g.sh [timeout_s] 'command'
set -e
cd "$(dirname "$0")/.."
T=600
if [[ "$1" =~ ^[0-9]+$ ]]; then T=$1; shift; fi
python -c "import __graft_entry__ as g; g.build()"
exec /usr/local/graft/bin/gpurun --timeout $T -- "$@"
