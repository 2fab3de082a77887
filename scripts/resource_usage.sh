#!/bin/sh
# Per-kernel registers / stack (spills) / shared memory of the built library, from cuobjdump -- static evidence that runs without a GPU.
#   sh scripts/resource_usage.sh > profiles/rNN_resource_usage.txt
LIB="${1:-ufemism2.0_b200/csrc/libufe_b200.so}"
echo "# cuobjdump -res-usage of libufe_b200.so (sm_100a), sorted by registers.  STACK is per-thread local memory: the k_vertex_diva / k_build_group entries are their per-thread column and patch arrays (indexed dynamically), the small k_mg_program / k_mg_tail entries are register spills at the 255 / 128 caps."
printf "%-4s %-6s %-7s %s\n" REG STACK SHARED kernel
cuobjdump -res-usage "$LIB" 2>/dev/null | awk '
  /^ Function /   { name = $2; sub(/:$/, "", name) }
  /^  REG:/       { split($1, r, ":"); split($2, s, ":"); split($3, h, ":"); print r[2], s[2], h[2], name }' |
while read reg stack shared name; do
  printf "%-4s %-6s %-7s %s\n" "$reg" "$stack" "$shared" "$(echo "$name" | c++filt | sed -e "s/(anonymous namespace):://g" -e "s/^void //" -e "s/(.*//")"
done | sort -k1,1nr -k4,4
