#!/bin/bash
# ncu --set full captures of one launch per kernel family (1 GPU).  Each capture is preceded by the same command
# without ncu (must exit 0).  Reports land in gpurun_out/ncu/<name>.ncu-rep; summarise here with scripts/summarize_ncu.py.
#   scripts/ncu_all.sh [name ...]     (default: every family below)
mkdir -p gpurun_out/ncu
declare -A K=(
  [lovasz]="lovasz:lovasz_hinge_kernel" [softmaxce]="softmaxce:softmax_ce_kernel" [segmetrics]="segmetrics:seg_metrics_kernel"
  [augment]="augment:augment_seg_kernel" [adam]="adam:adam_kernel" [maxpool_fwd]="maxpool:maxpool_fwd_rows_kernel"
  [maxpool_bwd]="maxpool:maxpool_bwd_block_kernel" [upsample_fwd]="upsample:upsample_fwd_kernel"
  [upsample_bwd]="upsample:upsample_bwd_kernel" [smallk_fwd]="smallk:fwd_kernel" [smallk_bwd_weight]="smallk:bwd_weight_kernel"
  [depthwise_fwd]="depthwise:depthwise3x3_tma_kernel" [depthwise_dgrad]="depthwise_bwd:depthwise3x3_tma_kernel" [depthwise_wgrad]="depthwise_bwd:depthwise3x3_tma_wgrad_kernel"
  [p2p_two_shot]="p2p:two_shot_kernel" [p2p_one_shot]="p2p:one_shot_kernel" [p2p_two_shot_bf16_wire]="p2p_wire:two_shot_wire_kernel"
  [conv_pair_3x3_256_14x14]="pair:conv_fprop2_kernel" [bn_bwd_partial_802816x64]="bn64:bn_bwd_partial_kernel"
  [fprop_3x3_64_56x56_resident]="conv3x3:conv_fprop_kernel" [wgrad_3x3_64_56x56_row]="wgrad64:conv_wgrad_kernel"
  [bn_bwd_partial_50176x1024]="bnwide:bn_bwd_partial_t_kernel" [bn_apply_802816x64]="bn64:bn_apply_kernel"
)
names=("$@"); [ ${#names[@]} -eq 0 ] && names=("${!K[@]}")
for name in "${names[@]}"; do
  kind=${K[$name]%%:*}; kern=${K[$name]#*:}
  python scripts/ncu_misc.py $kind > gpurun_out/ncu/$name.plain.log 2>&1 || { echo "$name: plain run failed"; continue; }
  ncu --set full --clock-control none --import-source on -k regex:$kern -s 2 -c 1 -f -o gpurun_out/ncu/$name \
      python scripts/ncu_misc.py $kind > gpurun_out/ncu/$name.ncu.log 2>&1 && echo "$name: captured" || echo "$name: ncu failed"
done
ls -la gpurun_out/ncu/*.ncu-rep 2>/dev/null | wc -l
