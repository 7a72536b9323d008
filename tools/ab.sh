#!/bin/bash
# usage: ab.sh v1 v2 ...  (base = default lib)
for v in "$@"; do echo "== $v"; if [ $v = base ]; then python tools/r02_probe.py --steps 16 2>&1 | grep -i "stages (ms/frame):\|in flight 3\|in flight 1\|sha1" | cut -c1-200; else B200PT_LIB=build/variants/libb200pt_$v.so python tools/r02_probe.py --steps 16 2>&1 | grep -i "stages (ms/frame):\|in flight 3\|in flight 1\|sha1" | cut -c1-200; fi; done
