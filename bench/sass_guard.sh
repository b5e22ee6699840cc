#!/bin/bash
# SASS evidence for the built library, no GPU needed (cuobjdump only):
#   * writes one gzip'd listing per kernel to profiles/sass/ (with --write) and a census of the mnemonics that matter;
#   * checks what each kernel family must and must not contain:
#       bulkRoundKernel<*, false>  UBLKCP.S.G + UBLKCP.G.S (cp.async.bulk both ways), SYNCS.*TRANS64 (mbarrier tx), no LDL / STL
#       bulkRoundKernel<*, true>   the same + LDGMC (multimem.ld_reduce)
#       nvlsPipelinedKernel        LDGMC
#       every kernel               no LDL / STL (no register spills anywhere in the library)
# Exit code 1 if a check fails.
cd "$(dirname "$0")/.." || exit 1
LIB=akka_allreduce_b200/libaar_cuda.so
WRITE=0; [ "${1:-}" == "--write" ] && WRITE=1
rc=0
census=profiles/r2_sass_census.txt
[ $WRITE == 1 ] && { mkdir -p profiles/sass; : > $census; }
funcs=$(cuobjdump --dump-resource-usage $LIB 2>&1 | grep -o "Function [^:]*" | awk '{print $2}')
for fn in $funcs; do
  name=$(echo "$fn" | c++filt | sed -E 's/^void //; s/\(anonymous namespace\):://; s/^<unnamed>:://; s/([A-Za-z0-9_]+)(<[^(]*>)?\(.*/\1\2/; s/\(bool\)//g; s/\(int\)//g; s/ //g; s/__nv_bfloat16/bf16/g')
  sass=$(cuobjdump -sass -fun "$fn" $LIB 2>/dev/null)
  count() { echo "$sass" | grep -c -E "^\s+/\*[0-9a-f]+\*/\s+(@!?U?P[0-9T] +)?$1"; }
  line="$name: instr=$(echo "$sass" | grep -c -E '^\s+/\*[0-9a-f]{4,6}\*/') UBLKCP.S.G=$(count 'UBLKCP\.S\.G') UBLKCP.G.S=$(count 'UBLKCP\.G\.S') SYNCS=$(count 'SYNCS') LDGMC=$(count 'LDGMC') STG=$(count 'STG') LDG=$(count 'LDG\.') MEMBAR=$(count 'MEMBAR') ATOMG=$(count 'ATOMG') LDL=$(count 'LDL') STL=$(count 'STL')"
  echo "$line"
  [ $WRITE == 1 ] && { echo "$line" >> $census; echo "$sass" | gzip -9 > "profiles/sass/$(echo "$name" | tr '<>,' '___' | sed 's/_$//').sass.gz"; }
  for bad in LDL STL; do [ "$(count "$bad")" -eq 0 ] || { echo "  SPILL ($bad) in $name"; rc=1; }; done
  case "$name" in
    bulkRoundKernel*)
      for need in 'UBLKCP\.S\.G' 'UBLKCP\.G\.S' 'SYNCS.*TRANS64'; do [ "$(count "$need")" -gt 0 ] || { echo "  MISSING $need in $name"; rc=1; }; done
      case "$name" in *true*) [ "$(count LDGMC)" -gt 0 ] || { echo "  MISSING LDGMC in $name"; rc=1; };; esac ;;
    nvlsPipelinedKernel*) [ "$(count LDGMC)" -gt 0 ] || { echo "  MISSING LDGMC in $name"; rc=1; } ;;
  esac
done
exit $rc
