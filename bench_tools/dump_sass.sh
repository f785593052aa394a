#!/bin/bash
# Full SASS listings (cuobjdump -sass -fun) of the kernels the docs cite, plus a mnemonic census per kernel.
#   bash bench_tools/dump_sass.sh     -> profiles/sass/*.sass, profiles/sass/CENSUS.md
SO=pytorch_distributed_b200/lib/libpdb200.so
OUT=profiles/sass
mkdir -p $OUT
rm -f $OUT/*.txt
declare -A K=(
 [conv_fprop_bn256_bf16]='_ZN3pdb17conv_fprop_kernelILi256ELi0ELb0ELb0EEEv14CUtensorMap_stS1_S1_NS_11FpropParamsE'
 [conv_fprop_pair_bf16]='_ZN3pdb17conv_fprop_kernelILi256ELi0ELb1ELb0EEEv14CUtensorMap_stS1_S1_NS_11FpropParamsE'
 [conv_fprop_pair_fp16]='_ZN3pdb17conv_fprop_kernelILi256ELi1ELb1ELb0EEEv14CUtensorMap_stS1_S1_NS_11FpropParamsE'
 [conv_fprop_bn256_tf32]='_ZN3pdb17conv_fprop_kernelILi256ELi2ELb0ELb0EEEv14CUtensorMap_stS1_S1_NS_11FpropParamsE'
 [conv_fprop_pair_tf32]='_ZN3pdb17conv_fprop_kernelILi256ELi2ELb1ELb0EEEv14CUtensorMap_stS1_S1_NS_11FpropParamsE'
 [conv_fprop_stem_bf16]='_ZN3pdb17conv_fprop_kernelILi64ELi0ELb0ELb1EEEv14CUtensorMap_stS1_S1_NS_11FpropParamsE'
 [conv_wgrad_bn256_bf16]='_ZN3pdb17conv_wgrad_kernelILi256ELi0ELb0EEEv14CUtensorMap_stS1_NS_11WgradParamsE'
 [conv_wgrad_bn256_tf32]='_ZN3pdb17conv_wgrad_kernelILi256ELi2ELb0EEEv14CUtensorMap_stS1_NS_11WgradParamsE'
 [conv_wgrad_stem_bf16]='_ZN3pdb17conv_wgrad_kernelILi256ELi0ELb1EEEv14CUtensorMap_stS1_NS_11WgradParamsE'
)
names=$(cuobjdump -sass $SO | grep -o "Function : .*" | sed 's/Function : //')
pick() { echo "$names" | grep -E "$1" | head -1; }
K[allreduce_nvls]=$(pick 'allreduce_kernelILb1ELb0E')
K[allreduce_p2p]=$(pick 'allreduce_kernelILb0ELb0E')
K[allreduce_nvls_smallcta]=$(pick 'allreduce_kernelILb1ELb1E')
K[allreduce_oneshot]=$(pick 'allreduce_oneshot_kernel')
K[allreduce_wire16_bf16]=$(pick 'allreduce_wire16_kernelI13__nv_bfloat16')
K[small_allreduce]=$(pick 'small_allreduce_kernel')
K[broadcast]=$(pick 'broadcast_kernel')
K[peer_gather]=$(pick 'peer_gather_kernel')
K[reduce_to_root]=$(pick 'reduce_to_root_kernel')
K[bn_stats_sync_bf16]=$(pick 'bn_stats_sync_kernelI13__nv_bfloat16')
K[bn_bwd_reduce_bits_bf16]=$(pick 'bn_bwd_reduce_kernelI13__nv_bfloat16Li2E')
K[bn_apply_fast_relu_bf16]=$(pick 'bn_apply_fast_kernelI13__nv_bfloat16Lb0ELb1E')
K[bn_bwd_apply_fast_bits_bf16]=$(pick 'bn_bwd_apply_fast_kernelI13__nv_bfloat16Li2ELb0E')
K[sgd]=$(pick 'sgd_kernelI6__halfLb1E')
{
echo "# SASS census (cuobjdump -sass -fun, sm_100a) — counts of the instructions that prove the hardware path"
echo
echo "| listing | kernel | lines | UTCHMMA (f16/bf16) | UTCQMMA/UTCMMA other | 2CTA | UTMALDG | IM2COL | UTMASTG | UTMAREDG | LDTM | multimem (LDGMC/STGMC/..) | st.release / ld.acquire .SYS |"
echo "|---|---|---|---|---|---|---|---|---|---|---|---|---|"
for n in $(echo "${!K[@]}" | tr ' ' '\n' | sort); do
  f=${K[$n]}
  [ -z "$f" ] && { echo "| $n | (not found) |"; continue; }
  # (the 64-bit encodings are dropped: they double the size and nobody reads them)
  cuobjdump -sass -fun "$f" $SO 2>/dev/null | sed -E 's@[[:space:]]*/\* 0x[0-9a-f]{16} \*/[[:space:]]*$@@' | grep -v '^[[:space:]]*$' > $OUT/$n.sass
  c() { grep -c -E "$1" $OUT/$n.sass; }
  echo "| $n.sass | \`$(echo $f | cut -c1-60)\` | $(wc -l < $OUT/$n.sass) | $(c 'UTCHMMA') | $(grep -E 'UTC[A-Z]*MMA' $OUT/$n.sass | grep -vc UTCHMMA) | $(c '2CTA') | $(c 'UTMALDG') | $(c 'IM2COL') | $(c 'UTMASTG') | $(c 'UTMAREDG') | $(c 'LDTM') | $(c 'LDGMC|STGMC|MULTIMEM|REDGMC') | $(c 'STRONG\.SYS|\.SYS') |"
done
} > $OUT/CENSUS.md
cat $OUT/CENSUS.md
