#!/bin/bash
# instructions per kernel of a built library:  tools/sass_count.sh <lib.so> [filter]
cuobjdump -sass "$1" | awk '/Function :/ {name=$3} /^[ \t]+\/\*[0-9a-f]+\*\/[ \t]+[A-Z@!]/ {n[name]++} END {for (k in n) print n[k], k}' | c++filt | sed -E 's/\(anonymous namespace\):://; s/\(.*//; s/void //' | sort -k2 | grep -E "${2:-.}"
