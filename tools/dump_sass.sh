#!/bin/bash
# Full SASS listings of the hot kernels (cuobjdump -sass, per function) -> profiles/sass/*.sass, plus a mnemonic census that shows
# the Blackwell-native instructions (UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA, UBLKCP = bulk copy,
# SYNCS = mbarrier, multimem = MULTIMEM-tagged LD/ST/RED).
cd "$(dirname "$0")/.."
SO=ml_trainer_b200/lib/libmlt_b200.so
OUT=profiles/sass
mkdir -p $OUT
cuobjdump -sass $SO > /tmp/all.sass
dump() {  # name-fragment (mangled substring), output file
  awk -v pat="$1" 'BEGIN{on=0} /Function :/{on = index($0, pat) > 0} on{print}' /tmp/all.sass > $OUT/$2.sass
  echo "$2: $(wc -l < $OUT/$2.sass) lines"
}
dump "gemm_bf16_tcgen05_kernelILi128ELb0" gemm_bf16_tcgen05_bn128
dump "gemm_bf16_tcgen05_kernelILi256ELb0" gemm_bf16_tcgen05_bn256
dump "gemm_bf16_2cta_kernel" gemm_bf16_2cta_cta_group2
dump "attention_fwd_tcgen05_kernelILb0ELb0" attention_fwd_tcgen05
dump "attention_fwd_tcgen05_kernelILb1ELb0" attention_fwd_tcgen05_keybias
dump "attention_fwd_tcgen05_kernelILb1ELb1" attention_fwd_tcgen05_keybias_dropout
dump "attention_bwd_dq_kernelILb0ELb0" attention_bwd_dq
dump "attention_bwd_dkdv_kernelILb0ELb0" attention_bwd_dkdv
dump "attention_bwd_dkdv_kernelILb1ELb1" attention_bwd_dkdv_keybias_dropout
dump "bn_stats_kernel" bn_stats
dump "bn_apply_kernelI13__nv_bfloat16Lb1ELb1" bn_apply_residual_relu
dump "lenet_step_cluster_kernelINS_7SgdRuleELb0" lenet_step_cluster_sgd_single
dump "lenet_step_cluster_kernelINS_7SgdRuleELb1" lenet_step_cluster_sgd_multi
dump "allreduce_twoshot_update_kernelINS_9AdamWRuleE13__nv_bfloat16Lb1" allreduce_nvls_update_adamw_bf16
dump "allreduce_oneshot_update_kernelINS_7SgdRuleEf" allreduce_oneshot_update_sgd_fp32
for f in $OUT/*.sass; do
  echo "== $(basename $f)"; grep -oE "^\s+/\*[0-9a-f]+\*/\s+[A-Z0-9_.]+" $f | awk '{print $2}' | sed 's/\..*//' | sort | uniq -c | sort -rn | \
    grep -E "UTC|LDTM|STTM|UTMA|UBLKCP|SYNCS|MULTIMEM|HMMA|REDG|ATOM|LDG|STG|ST |LD " | head -14
done > $OUT/MNEMONICS.txt
grep -l . $OUT/*.sass | wc -l
# the largest listings are kept compressed (zcat profiles/sass/<name>.sass.gz)
for f in $OUT/*.sass; do
  if [ $(wc -l < $f) -gt 15000 ]; then gzip -9 -f $f; fi
done
ls -la $OUT
