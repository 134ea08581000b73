#!/bin/bash
# Round 2, GPU call 2: fp32 merge-path + TF32 tcgen05 kernel, staged row-split kernel after the shuffle fix, and the ncu
# captures VERDICT r1 asked for (merge on cfg3a/cfg3b, shipping BCSR TMA kernel on cfg4, sliced cfg5-large k=128).
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_f32.py -m gpu -x -q > $O/r2b_pytest_f32.log 2>&1; echo "rc=$?" >> $O/r2b_pytest_f32.log; tail -15 $O/r2b_pytest_f32.log
timeout 600 python -m pytest tests/test_gpu_round2.py tests/test_nnz_carry.py tests/test_pipeline.py tests/test_gpu_parity.py -m gpu -q -k "not full_size" > $O/r2b_pytest_rest.log 2>&1; echo "rc=$?" >> $O/r2b_pytest_rest.log; tail -15 $O/r2b_pytest_rest.log
timeout 300 python scripts/tf32_probe.py --workload cfg4 > $O/r2b_tf32_cfg4.log 2>&1; cat $O/r2b_tf32_cfg4.log
timeout 300 python scripts/knob_probe.py --workload cfg2 --set TICKET=4 --set TICKET=4,STAGE=1 --set TICKET=4,STAGE=1,TICKET_CTAS=3 --set TICKET=8,STAGE=2 > $O/r2b_stage_cfg2_k32.log 2>&1; cat $O/r2b_stage_cfg2_k32.log
timeout 300 python scripts/knob_probe.py --workload cfg2 --k 64 --set TICKET=8 --set TICKET=8,STAGE=1 > $O/r2b_stage_cfg2_k64.log 2>&1; cat $O/r2b_stage_cfg2_k64.log
timeout 300 python scripts/knob_probe.py --workload cfg2 --shard 0/8 --set TICKET=4 --set TICKET=4,STAGE=1 --set TICKET=2 > $O/r2b_shard8.log 2>&1; cat $O/r2b_shard8.log
# ncu: one --set full capture of the dominant kernel + the launch list of the same command, per workload
for W in cfg3a:spmm_merge_kernel cfg3b:spmm_merge_kernel cfg4:spmm_bcsr_dmma_tma cfg5_large:spmm_rowsplit; do
  wl=${W%%:*}; kn=${W##*:}
  timeout 300 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu --no-parity > $O/r2b_plain_$wl.log 2>&1 || { tail -3 $O/r2b_plain_$wl.log; continue; }
  skip=3; [ $wl = cfg5_large ] && skip=40
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$kn -s $skip -c 1 -o $O/r2b_ncu_$wl -f python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu --no-parity > $O/r2b_ncu_full_$wl.log 2>&1
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2b_launches_$wl.csv python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu --no-parity > $O/r2b_ncu_list_$wl.log 2>&1
  tail -c 400 $O/r2b_plain_$wl.log | head -c 300; echo
done
ls -la $O | tail -30
