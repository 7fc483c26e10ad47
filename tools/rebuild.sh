#!/bin/sh
# developer helper: incremental in-tree build, prints register / stack usage of the kernels matching $1
cd "$(dirname "$0")/.." && python -c "
import importlib.util
spec=importlib.util.spec_from_file_location('b','point_cloud_process_b200/build.py');m=importlib.util.module_from_spec(spec);spec.loader.exec_module(m);m.build()" 2>&1 | tail -30
[ -n "$1" ] && grep -h -A3 "Compiling entry function.*$1" point_cloud_process_b200/_obj/*.log | grep "Used\|stack\|Compiling" | sed 's/.*function .\(_Z[^ ]*\). for.*/\1/' | cut -c1-150
