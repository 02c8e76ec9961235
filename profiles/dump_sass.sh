#!/bin/sh
# SASS listings of the trace loop's stage kernels, one file per kernel under profiles/sass/
# (instruction text only: the encoding lines of `cuobjdump -sass` are dropped).
#   sh profiles/dump_sass.sh [lib.so]
LIB=${1:-cs378hgraphics_raytracer_b200/librt_b200.so}
OUT=profiles/sass
mkdir -p "$OUT"
cuobjdump -sass "$LIB" | awk -v out="$OUT" '
/Function :/ {
	name = $3; keep = 0; file = ""
	if (name ~ /k_wave_gen|k_wave_closestE|k_wave_closest_retry|k_wave_closest_heavy|k_wave_shade|k_wave_shadowE|k_wave_shadow_any|k_wave_shadow_list|k_shadow_keys|k_closest_keys|k_resolve_grid|k_adapt_resolve|k_adapt_samples|k_pack_tiles|k_unpack_tiles|k_hits_static/) {
		short = name; sub(/^_ZN3rtd[0-9]+/, "", short); sub(/E.*/, "", short)
		file = out "/" short ".sass"; keep = 1
		print "// " name > file
	}
	next
}
keep && /^ +\/\*[0-9a-f]+\*\// { sub(/ +\/\* 0x[0-9a-f]+ \*\/$/, ""); print >> file }
'
wc -l "$OUT"/*.sass
