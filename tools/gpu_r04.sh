#!/usr/bin/env bash
# One gpurun call for the r04 single-GPU evidence: all GPU parity tests, smoke, default bench line, prefix-prefill
# lines for the three kernels, launch list + full ncu capture of the prefix-prefill kernel.
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT
( time python -m pytest tests -m gpu -x -q ) > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/pytest_gpu.log
python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -4 $OUT/smoke.log
python bench.py > $OUT/bench_c2_v1.json 2> $OUT/bench_c2_v1.err; echo "bench exit $?"; cut -c1-300 $OUT/bench_c2_v1.json
python tools/prefix_prefill_bench.py > $OUT/prefix_prefill.jsonl 2> $OUT/prefix_prefill.err; echo "prefix bench exit $?"
B200_CTX_ATTN=tc2 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill_tc2.jsonl 2>&1
B200_CTX_ATTN=mma python tools/prefix_prefill_bench.py > $OUT/prefix_prefill_mma.jsonl 2>&1
cut -c1-200 $OUT/prefix_prefill.jsonl
c=llama7b_chunk2k_ctx2k_f16
CMD="python tools/prefix_prefill_bench.py $c 3"
$CMD > $OUT/pp_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/prefix_launches.csv $CMD > $OUT/ncu_prefix_launches.log 2>&1
echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on -k regex:context_attn_tc -s 2 -c 1 -f -o $OUT/prof_ctx_tc $CMD > $OUT/ncu_ctx_tc.log 2>&1
echo "ncu full exit $?"
