for mr in 8192 2048; do for b in 2048 4096 8192 16384; do
POZ_TC3_X32_MIN_ROWS=$mr timeout 100 python bench.py --workload c3 --batch $b --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('minrows',$mr,'batch',$b, round(d['ms_per_step']*1000,1),'us', d['roofline']['kernels_ms'])"
done; done
