#!/usr/bin/env bash
# One ncu --set full capture per named kernel (single GPU).  Each capture runs only after the same
# command has exited 0 without ncu, as the profiling recipe requires.
set -u
mkdir -p gpurun_out
KEEP="gemm2 gemm conv3x3 conv_s2_dgrad attn512_fwd attn512_bwd wgrad ngd_tc_project"   # full reports kept (gpurun_out is capped at 64 MiB)
# ONLY="name name ..." restricts the run to those captures
cap() {  # name target regex skip
  if [ -n "${ONLY:-}" ]; then case " $ONLY " in *" $1 "*) ;; *) return ;; esac; fi
  if [ ! -f gpurun_out/plain_ok_$2 ]; then
    python tools/ncu_target.py "$2" > gpurun_out/plain_$2.log 2>&1 && touch gpurun_out/plain_ok_$2
  fi
  [ -f gpurun_out/plain_ok_$2 ] || { echo "plain run of $2 failed: not profiled"; return; }
  ncu --set full --clock-control none --import-source on -k "regex:$3" -s "$4" -c 1 -f -o gpurun_out/prof_$1 \
      python tools/ncu_target.py "$2" > gpurun_out/ncu_$1.log 2>&1
  tail -1 gpurun_out/ncu_$1.log
  # the raw metric page as CSV (what tools/ncu_summary.py reads); drop the report itself unless it is on the keep list
  ncu -i gpurun_out/prof_$1.ncu-rep --page raw --csv > gpurun_out/ncuraw_$1.csv 2>/dev/null
  case " $KEEP " in *" $1 "*) ;; *) rm -f gpurun_out/prof_$1.ncu-rep ;; esac
  rm -f gpurun_out/ncu_$1.log
}
cap conv3x3 conv3x3 gemm_nt 2
cap conv1x1 conv1x1 gemm_nt 2
cap conv1x1_stats conv1x1_stats gemm_nt 2
cap gemm gemm1 gemm_nt_kernel 1
cap gemm2 gemm2 gemm2_nt 1
cap wgrad wgrad wgrad_kernel 2
cap attn_fwd attn attn_fwd 1
cap attn_bwd attn attn_bwd 1
cap chan_affine chan_affine chan_affine 2
cap chan_reduce chan_reduce chan_reduce2 2
cap ngd_eig ngd_eig ngd_eig 2
cap ngd_axis ngd_axis ngd_small_axis 2
cap ngd_tc_project ngd_tc tf32_gemm 3
cap ngd_tc_deflate ngd_tc tf32_gemm 4
cap ngd_tc_stats ngd_tc tf32_gemm 5
cap lin_prep lin_prep linear_bwd_prep 2
cap wprep wprep dgrad_weights 2
cap conv_s2_fwd conv_s2 gemm_nt_kernel 2
cap conv_s2_dgrad conv_s2 gemm_nt_kernel 5
cap attn512_fwd attn512 attn_fwd 1
cap attn512_bwd attn512 attn_bwd 1
cap ln_fwd rowwise ln_fwd 1
cap ln_bwd rowwise ln_bwd 1
cap embed_fwd rowwise embed_fwd 1
cap embed_bwd rowwise embed_bwd 1
cap mixup_ce rowwise mixup_ce 1
cap head_mix rowwise head_mix_fwd 1
cap sgd_momentum optim sgd_momentum_dev 1
cap clip_wd optim clip_wd 1
cap madgrad optim madgrad_dev 1
cap input_aug_mixup input_head input_aug_mixup_kernel 1
cap input_aug_mixup_bwd input_head input_aug_mixup_bwd 1
cap pool_fc_fwd input_head pool_fc_fwd 1
cap pool_fc_bwd input_head pool_fc_bwd 1
cap ngd_small_matrix ngd_small ngd_small_matrix 1
