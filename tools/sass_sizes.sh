#!/bin/sh
# SASS instruction count per kernel/function of the built library
cuobjdump -sass "${1:-rust_raytracer_b200/libb200rt.so}" 2>/dev/null | awk '/Function :/{name=$3} /^ +\/\*[0-9a-f]+\*\//{cnt[name]++} END{for(n in cnt) print cnt[n], n}' | sort -n
