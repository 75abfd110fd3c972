#!/bin/bash
# One-GPU measurement pass behind profiles/: tests, kernel timing tables, both bench arms, ncu captures.
# Every ncu capture follows a plain run of the same command that exited 0.
mkdir -p gpurun_out
O=gpurun_out
timeout 400 python -m pytest tests -m gpu -q --timeout 200 > /tmp/pytest.log 2>&1; echo "pytest rc=$?"; tail -2 /tmp/pytest.log | tee $O/pytest_gpu.log
timeout 200 python tools/profile_kernels.py > $O/profile_kernels.log 2>&1; echo "profile_kernels rc=$?"
timeout 200 python tools/conv_check.py all --time > $O/conv_check.log 2>&1; echo "conv_check rc=$?"
timeout 200 python tools/gemm_bench.py > $O/gemm_bench.log 2>&1; echo "gemm_bench rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 > /tmp/b.log 2>&1; echo "bench ours rc=$?"; tail -1 /tmp/b.log > $O/bench_ours_bf16.json
timeout 300 python bench.py --dtype fp32 --steps 5 --warmup 3 > /tmp/b.log 2>&1; echo "bench ours fp32 rc=$?"; tail -1 /tmp/b.log > $O/bench_ours_fp32.json
timeout 400 python bench.py --impl reference --steps 5 --warmup 3 > /tmp/b.log 2>&1; echo "bench ref rc=$?"; tail -1 /tmp/b.log > $O/bench_ref_bf16.json
timeout 400 python bench.py --impl reference --dtype fp32 --steps 3 --warmup 3 > /tmp/b.log 2>&1; echo "bench ref fp32 rc=$?"; tail -1 /tmp/b.log > $O/bench_ref_fp32.json
cap() {  # name, kernel regex, conv_one args...
  local name=$1 regex=$2; shift 2
  if timeout 60 python tools/conv_one.py "$@" > /dev/null 2>&1; then
    timeout 200 ncu --set full --clock-control none --import-source on -k "regex:$regex" -s 2 -c 1 -f -o $O/ncu_$name python tools/conv_one.py "$@" > $O/ncu_$name.log 2>&1
    echo "ncu $name rc=$?"
  else echo "plain $name failed"; fi
}
cap conv_halo_fprop conv_halo fprop 64 64 32 32 64 3 1 1 4
cap conv_splitk_fprop conv_igemm_splitk fprop 64 256 8 8 256 3 1 1 4
cap conv_igemm_s2 conv_igemm_kernel fprop 64 64 32 32 128 3 2 1 4
cap conv_wgrad conv_wgrad wgrad 64 64 32 32 64 3 1 1 4
ls -la $O | head -40
