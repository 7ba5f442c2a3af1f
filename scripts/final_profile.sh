#!/bin/bash
# Round-end evidence on ONE B200 (run under gpurun): GPU tests, both bench arms, the ncu launch list of a short bench run and one
# ncu --set full capture of the dominant kernel over a whole 2,400-unknown solve.  Numbers printed under ncu are never bench values.
set -x
python -m pytest tests -m gpu -q 2>&1 | tail -3 > gpurun_out/final_pytest.txt
python bench.py --impl reference > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
python bench.py --steps 2 --warmup 1 --no-extras --no-cpu > gpurun_out/final_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/final_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-extras --no-cpu > gpurun_out/final_ncu_list.log 2>&1
VX_LDLT_NO_GRAPH=1 python scripts/ldlt_time.py > gpurun_out/final_ldlt_plain.log 2>&1 && \
VX_LDLT_NO_GRAPH=1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum --clock-control none -k regex:k_ldlt_trailing -c 74 --csv \
    --log-file gpurun_out/final_ldlt_trailing_traffic.csv python scripts/ldlt_time.py > gpurun_out/final_ncu_traffic.log 2>&1
# the --set full capture: 4 launches from the middle of the sweep (a whole sweep at --set full is 90 MB; gpurun_out carries 64)
VX_LDLT_NO_GRAPH=1 ncu --set full --clock-control none --import-source on -k regex:k_ldlt_trailing -s 20 -c 4 -o gpurun_out/final_ldlt_trailing \
    python scripts/ldlt_time.py > gpurun_out/final_ncu_full.log 2>&1
ls -la gpurun_out/ | tail -12
