#!/bin/bash
# SASS instruction count of every device function inside thread_kernel<2,5> of an object/library (development aid)
f=$(realpath "$1"); tmp=$(mktemp -d); cd $tmp
cuobjdump -xelf all "$f" >/dev/null 2>&1
for c in *.cubin; do nvdisasm -c "$c" 2>/dev/null; done | awk '
/^_Z[A-Za-z0-9_]+:$/ { on = ($0 ~ /thread_kernelILi2ELi5E/); cur = "KERNEL" }
/^\$_Z.*\$_ZN3dbt/ { if (on) { s=$0; sub(/^.*\$_ZN3dbt[0-9]+/, "", s); sub(/(ILi|ERNS|E[a-z]).*$/, "", s); cur=s } }
/^ +\/\*[0-9a-f]+\*\// { if (on) n[cur]++ }
END { for (k in n) { print n[k], k; t+=n[k] } print t, "TOTAL" }' | sort -rn
rm -rf $tmp
