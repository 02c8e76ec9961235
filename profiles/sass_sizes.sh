#!/bin/sh
# Static SASS size of every function in the product library: sass_sizes.sh [lib.so]
LIB=${1:-cs378hgraphics_raytracer_b200/librt_b200.so}
cuobjdump -sass "$LIB" | awk '/Function :/ {if (name) print n, name; name=$3; n=0} /^ +\/\*[0-9a-f]+\*\// {n++} END {print n, name}' | sort -k2 | awk '{print $1, $2}'
