N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus $N --no-extra --no-cpu-baseline > gpurun_out/r2_bench_c4_${N}gpu.json 2> gpurun_out/mg_${N}.err
$TR bench.py --gpus $N --no-extra --no-cpu-baseline --workload c4lite > gpurun_out/r2_bench_c4lite_${N}gpu.json 2>> gpurun_out/mg_${N}.err
$TR tests/mgpu_check.py 2>&1 | tail -12
python tools/bench_brief.py gpurun_out/r2_bench_c4_${N}gpu.json | head -2
python tools/bench_brief.py gpurun_out/r2_bench_c4lite_${N}gpu.json | head -2
