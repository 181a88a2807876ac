#!/bin/bash
# Dev tool: static SASS instruction count (and opcode histogram with -v) of one kernel in a built library.
#   tools/sass_count.sh <lib.so> <mangled-name-regex> [-v]
lib=$1; pat=$2
cuobjdump -sass "$lib" 2>/dev/null | awk -v pat="$pat" '
  /Function : /{f = ($0 ~ pat)} f && /^[ \t]+\/\*[0-9a-f]{4}\*\//{n++; op=$2; sub(/\..*/,"",op); sub(/;/,"",op); h[op]++}
  END{print "instructions", n; if (ENVIRON["V"]) for (k in h) print h[k], k}' | sort -rn
