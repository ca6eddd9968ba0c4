#!/bin/bash
# usage: tools/sass_loop.sh <object or .so> <mangled-name substring>  -> SASS of that kernel, one instruction per line
cuobjdump -sass "$1" | awk -v pat="$2" '/Function :/{p = index($0, pat) > 0} p' | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/^\s+\/\*([0-9a-f]+)\*\/\s+//; s/\s+\/\*.*$//'
