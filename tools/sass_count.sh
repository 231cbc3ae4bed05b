#!/bin/sh
# static SASS statistics of one kernel in an object file (development tool): total instructions, branches, spills
# usage: tools/sass_count.sh obj.o kernel-name-substring
cuobjdump -sass "$1" | awk -v k="$2" '/Function : /{f=index($0,k)>0} f' | grep -E "^\s+/\*[0-9a-f]{4,}\*/" | awk '{op=$2; if (op ~ /^@/) op=$3; sub(/\..*/,"",op); sub(/;/,"",op); c[op]++; n++} END{printf "total %d\n", n; for (o in c) if (c[o]>=40) printf "%s %d\n", o, c[o]}' | sort -k2 -n -r | tr '\n' ' '; echo
