#!/bin/bash
# Plain runs (the "exits 0 without ncu" gate), then the ncu launch lists and the --set full captures of the dominant
# kernels.  Everything lands in gpurun_out/; tools/ncu_summary.py + `ncu -i … --page raw --csv` turn them into profiles/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
P="python tools/profile_step.py"
$P --batch 64 --ctx 4096 > gpurun_out/plain_bs64.log 2>&1 || { echo "plain bs64 failed"; tail -5 gpurun_out/plain_bs64.log; exit 1; }
$P --batch 1 --ctx 512 > gpurun_out/plain_bs1.log 2>&1 || { echo "plain bs1 failed"; tail -5 gpurun_out/plain_bs1.log; exit 1; }
python tools/bench_gemv.py --M 1 --variants 1 > gpurun_out/plain_gemv.jsonl 2>&1 || { echo "plain gemv failed"; exit 1; }
cat gpurun_out/plain_*.log
NCU="ncu --profile-from-start off --clock-control none"
$NCU --metrics gpu__time_duration.sum --csv --log-file gpurun_out/launches_bs64.csv $P --batch 64 --ctx 4096 > gpurun_out/ncu_bs64.log 2>&1
$NCU --metrics gpu__time_duration.sum --csv --log-file gpurun_out/launches_bs1.csv $P --batch 1 --ctx 512 > gpurun_out/ncu_bs1.log 2>&1
if [ "${FULL:-1}" = "1" ]; then
  F="--set full --import-source on"
  $NCU $F -k regex:paged_decode_attn -c 2 -f -o gpurun_out/full_attn_bs64 $P --batch 64 --ctx 4096 > gpurun_out/ncu_full_attn.log 2>&1
  $NCU $F -k regex:linear_small -s 10 -c 4 -f -o gpurun_out/full_linear_small_bs1 $P --batch 1 --ctx 512 > gpurun_out/ncu_full_ls.log 2>&1
  $NCU $F -k regex:gemm_oneshot -s 8 -c 4 -f -o gpurun_out/full_oneshot_bs64 $P --batch 64 --ctx 4096 > gpurun_out/ncu_full_os.log 2>&1
  $NCU $F -k regex:glue_rows -s 8 -c 4 -f -o gpurun_out/full_glue_bs64 $P --batch 64 --ctx 4096 > gpurun_out/ncu_full_gl.log 2>&1
  $NCU $F -k regex:lm_head -c 1 -f -o gpurun_out/full_lmhead_bs64 $P --batch 64 --ctx 4096 > gpurun_out/ncu_full_lm.log 2>&1
  G="ncu --clock-control none $F -k regex:ternary_gemv_stream -s 2 -c 1 -f"
  $G -o gpurun_out/full_gemv_70b_gateup python tools/bench_gemv.py --shapes 70b_gateup --variants 1 --iters 10 > gpurun_out/ncu_full_g1.log 2>&1
  $G -o gpurun_out/full_gemv_8b_gateup python tools/bench_gemv.py --shapes 8b_gateup --variants 1 --iters 10 > gpurun_out/ncu_full_g2.log 2>&1
  $G -o gpurun_out/full_gemv_2b_gateup python tools/bench_gemv.py --shapes 2b_gateup --variants 1 --iters 10 > gpurun_out/ncu_full_g3.log 2>&1
fi
ls -la gpurun_out | tail -30
