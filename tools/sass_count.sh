#!/bin/bash
# static SASS instruction count per kernel of one object: tools/sass_count.sh vk-renderer_b200/build/taa.o
cuobjdump -sass "$1" | awk '/Function :/ {name=$3} /^ +\/\*[0-9a-f][0-9a-f][0-9a-f][0-9a-f]\*\// {n[name]++} END {for (k in n) print n[k], k}' | sort -n
