#!/bin/bash
# builds libb200net.so in-tree (nvcc, sm_100a); usable from any directory
cd "$(dirname "$0")/.." && python -c "import __graft_entry__ as g; g.build(); print('build ok')"
