#!/bin/bash
# First GPU calls of the next round, in order, each one gpurun call.  Everything here was written after round 1's GPU budget
# was spent and has NOT run on hardware yet.  Outputs go to gpurun_out/ (scratch); copy what is worth keeping into profiles/.
#
#   gpurun --timeout 900 -- 'bash bench/next_round_gpu_plan.sh one'            # 1 GPU
#   gpurun --gpus 2 --timeout 900 -- 'bash bench/next_round_gpu_plan.sh two'   # 2 GPUs
#   gpurun --gpus 8 --timeout 1200 -- 'bash bench/next_round_gpu_plan.sh eight'
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR() { n=$1; port=$2; shift 2; python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@"; }

case "${1:-one}" in
one)
  # 1. the regular suite must still be green, then the unverified paths (flow-controlled broadcast, engine.start_at)
  timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/r2_gpu_suite.txt
  AAR_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_gpu_experimental.py -q -m gpu 2>&1 | tail -15 | tee gpurun_out/r2_experimental.txt
  timeout 300 python bench.py --steps 30 --warmup 5 | tee gpurun_out/r2_bench_n1.json
  timeout 300 python -m pytest tests/test_geometry_buffer.py -q -k native_engine_example 2>&1 | tail -3 | tee gpurun_out/r2_c_example.txt
  AAR_NATIVE_ENGINE=1 timeout 300 python bench.py --steps 30 --warmup 5 --no-e2e | tee gpurun_out/r2_bench_n1_native.json
  ;;
two)
  # 2. real peers: regular multi-GPU suite, then the same suite with the strict variants switched on (AAR_STRICT=1: broadcast
  #    flow control at thresholds < 1, one eager-gather decision per chunk at thresholds = 1)
  timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/r2_multi.txt
  AAR_STRICT=1 timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/r2_multi_fc.txt
  ;;
eight)
  # 3. does the flow control cost anything in straggler mode?  then the knobs that were added blind
  TR 8 29611 bench/straggler_bench.py 2>&1 | grep "^{" | tee gpurun_out/r2_straggler_default.json
  AAR_STRICT=1 TR 8 29612 bench/straggler_bench.py 2>&1 | grep "^{" | tee gpurun_out/r2_straggler_fc.json
  AAR_ELIDE_LATE=1 TR 8 29617 bench/straggler_bench.py 2>&1 | grep "^{" | tee gpurun_out/r2_straggler_elide.json
  TR 8 29613 bench.py --gpus 8 --steps 50 --warmup 5 --no-e2e 2>&1 | grep "^{" | tee gpurun_out/r2_bench_n8.json
  AAR_STRICT=1 TR 8 29616 bench.py --gpus 8 --steps 50 --warmup 5 --no-e2e 2>&1 | grep "^{" | tee gpurun_out/r2_bench_n8_strict.json
  TR 8 29614 bench.py --gpus 8 --steps 50 --warmup 5 --no-e2e --stage-input 2>&1 | grep "^{" | tee gpurun_out/r2_bench_n8_stage_input.json
  TR 8 29618 bench.py --gpus 8 --steps 50 --warmup 5 --no-e2e --no-nccl --view 2>&1 | grep "^{" | tee gpurun_out/r2_bench_n8_view.json
  TR 8 29619 bench.py --gpus 8 --steps 50 --warmup 5 --no-e2e --no-nccl --view --stage-input 2>&1 | grep "^{" | tee gpurun_out/r2_bench_n8_view_stage.json
  for ctas in 48:4 48:6 48:8 32:8 64:8 80:8; do
    TR 8 29615 bench.py --gpus 8 --steps 50 --warmup 5 --no-e2e --no-nccl --nvls-ctas $ctas 2>&1 | grep "^{" | tee gpurun_out/r2_bench_n8_ctas_${ctas/:/_}.json
  done
  ;;
esac
# Further sweeps once the above is green (each is one call; N = 8 unless noted):
#   TR 8 P bench/sweep.py --sizes 1048576,2097152,4194304 --algos auto,pipelined,nvls --grid 32      (and 16, 64): fewer CTAs, less sync
#   AAR_ONESHOT_MAX_BYTES=4194304 TR 8 P bench/sweep.py --sizes 262144,524288,1048576,2097152 --algos oneshot,nvls   one-shot crossover
#   TR 4 P bench/gpt2_dp.py --comm-ctas 32   (and 16, 48, 64; --algo pipelined)                      SM budget for the DDP buckets
# Host-side launch cost (small and mid sizes are bound by how fast one Python thread can enqueue rounds, ~28-38 us each):
#   * residentGrid() runs cudaOccupancyMaxActiveBlocksPerMultiprocessor + cudaDeviceGetAttribute on EVERY launch: cache per
#     (kernel pointer, device) -- the template parameter K is the function-pointer TYPE, shared by kernels with one signature,
#     so key the cache by pointer value, not by a static local per instantiation.
#   * enqueue_round() makes 3-4 torch stream/event calls per round; one C entry point (wait event, launch, record event) would
#     replace them.  Measure with `python -X importtime`-style timers around enqueue_round before and after.
#   TR 2 P bench/link_bench.py --bulk --mib 64        cp.async.bulk copy role: GB/s per GPU with 32-thread CTAs, local and to a peer
