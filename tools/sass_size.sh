#!/bin/bash
# tools/sass_size.sh lib.so -> SASS instruction count per kernel (called device functions are part of the kernel's text)
cuobjdump -sass "$1" 2>/dev/null | awk '/Function :/{name=$3} /^[ \t]+\/\*[0-9a-f]+\*\/[ \t]+[A-Z@]/{n[name]++} END{for(k in n) printf "%6d %s\n", n[k], substr(k,1,110)}' | sort -rn
