#!/bin/bash
# usage: ab_env.sh VAR v1 v2 ...   — tools/r02_probe.py with VAR set to each value ("-" = unset)
var=$1; shift
for v in "$@"; do echo "== $var=$v"; if [ "$v" = "-" ]; then python tools/r02_probe.py --steps 16 2>&1 | grep -i "stages (ms/frame):\|in flight 3\|in flight 1" | cut -c1-200; else env $var=$v python tools/r02_probe.py --steps 16 2>&1 | grep -i "stages (ms/frame):\|in flight 3\|in flight 1" | cut -c1-200; fi; done
