#!/bin/bash
# round 2, GPU call 14: split polytomies on the device (tests + timings), DRAM bytes of a real config-3 batch, Sankoff bench
out=gpurun_out/${1:-r2p}; mkdir -p $out
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py tests/test_gpu_comm.py -m gpu -x -q --durations=5 > $out/pytest_parity.log 2>&1; echo "rc=$?" >> $out/pytest_parity.log
bash profiles/poly_r2.sh ${1:-r2p}
timeout 400 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k 'regex:fwd_kernel|bwd_kernel' -c 2 --csv \
  --log-file $out/ncu_dram_c3_batch.csv python profiles/prof_cfg.py c3 fitch 2048 0 1 > $out/ncu_dram_c3_batch.log 2>&1
timeout 600 python bench.py --algo sankoff --steps 4 --warmup 3 --no-e2e --no-c2 --cpu-cols 16 > $out/bench_c3_sankoff.json 2> $out/bench_c3_sankoff.err
