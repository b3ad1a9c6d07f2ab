#!/bin/bash
# Single-GPU validation of a round: the whole -m gpu suite, smoke(), the bench line, then the two ncu
# captures (launch list of the eager step; --set full of the pack / update kernels), each only after the
# same command exited 0 without ncu.  Everything lands in gpurun_out/r2_*.
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -q > $O/r2_pytest_gpu1_final.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2_pytest_gpu1_final.log
python __graft_entry__.py smoke 2>&1 | tail -1
python bench.py --steps 20 --warmup 5 > $O/r2_bench_n1_final.json 2> $O/r2_bench_n1_final.err; echo "bench rc=$?"
python tools/sampler_bench.py > $O/r2_sampler_bench.json 2>&1; tail -c 400 $O/r2_sampler_bench.json; echo
B="python bench.py --steps 2 --warmup 1 --settle 3 --step eager --no-cpu-baseline --no-kernel-bench --no-torch-baseline"
$B > $O/r2_plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 1400 -c 700 --csv \
    --log-file $O/r2_launches_n1.csv $B > $O/r2_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
K="python tools/kernel_bench.py --model resnet18 --only pack_fp32,unpack_fp32,update_direct,update_from_bucket_write_grad --no-torch"
$K > $O/r2_plain_kb.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"update_kernel|pack_kernel" \
    -s 30 -c 6 -o $O/r2_prof_kernels $K > $O/r2_ncu_full.log 2>&1; echo "ncu full rc=$?"
