#!/bin/bash
# dump the SASS of one kernel of libsvo_gpu.so as plain "opcode operands" lines: scripts/sass_loop.sh <name-regex> > out.txt
cuobjdump -sass svo_b200/lib/libsvo_gpu.so 2>/dev/null | awk -v pat="$1" '
/Function :/ { on = ($0 ~ pat) }
on && /^[ \t]+\/\*[0-9a-f]+\*\/[ \t]+[A-Z@]/ { sub(/^[ \t]+\/\*[0-9a-f]+\*\/[ \t]+/, ""); sub(/[ \t]*\/\*.*$/, ""); print }'
