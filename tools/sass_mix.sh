#!/bin/bash
# usage: tools/sass_mix.sh <mangled-kernel-name> [object]   -- opcode histogram of one kernel's SASS
OBJ=${2:-build/clt_spec.o}
cuobjdump -sass -fun "$1" "$OBJ" > build/$1.sass
grep -E "^\s+/\*[0-9a-f]{4}\*/" build/$1.sass | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//' | sed -E 's/^@!?U?P[0-9T]+ +//' | awk '{print $1}' | sed 's/;//' | sort | uniq -c | sort -rn | head -${3:-25}
