#!/bin/bash
# tools/sass_loop.sh <rt_core.o> [mangled kernel] : instruction count of the kernel and a dump in /tmp/<basename>.s
K=${2:-_ZN3rtd17trace_rays_kernelILb0ELb0ELb0EEEvNS_11DeviceSceneENS_10BounceArgsE}
cuobjdump -sass -fun "$K" "$1" | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/^\s+//; s/\s+\/\* 0x[0-9a-f]+ \*\/\s*$//' > /tmp/$(basename $(dirname $1)).s
grep -vc "NOP" /tmp/$(basename $(dirname $1)).s
