#!/bin/bash
# Compares the opcode sequence of the DEFAULT kernel instantiations in the built library with the archived listings the round-1
# measurements were taken with (profiles/sass/*.sass.gz).  Used when code is added next to the measured paths (opt-in template
# variants) without GPU access: identical sequences = the measured paths compile unchanged.  No GPU needed (cuobjdump only).
cd "$(dirname "$0")/.." || exit 1
LIB=akka_allreduce_b200/libaar_cuda.so
ops() { grep -E "^\s+/\*[0-9a-f]{4,6}\*/" | sed -E 's#/\*[0-9a-f]+\*/##g' | awk '{print $1}'; }
funcs=$(cuobjdump --dump-resource-usage $LIB 2>&1 | grep -o "Function [^:]*" | awk '{print $2}')
rc=0
for pair in "nvlsPipelinedKernel_f32:19nvlsPipelinedKernelIfLb0EEEv" "nvlsPipelinedKernel_bf16:19nvlsPipelinedKernelI13__nv_bfloat16Lb0EEEv" \
            "roundKernel_f32:11roundKernelIfLb0ELb0EEEv" "roundKernel_bf16:11roundKernelI13__nv_bfloat16Lb0ELb0EEEv" \
            "oneshotKernel_f32:13oneshotKernelIfEEv" "persistentKernel_f32:16persistentKernelIfEEv" "p2pPipelinedKernel_f32:18p2pPipelinedKernelIfLb0EEEv"; do
  f=${pair%%:*}; k=${pair##*:}
  fn=$(echo "$funcs" | grep "$k" | head -1)
  if [ -z "$fn" ]; then echo "$f: no function matching $k in $LIB"; rc=1; continue; fi
  now=$(cuobjdump -sass -fun "$fn" $LIB 2>/dev/null | ops)
  arch=$(zcat profiles/sass/$f.sass.gz | ops)
  if [ "$now" == "$arch" ]; then echo "$f: identical ($(echo "$now" | wc -l) instructions)"; else echo "$f: DIFFERS"; rc=1; fi
done
exit $rc
