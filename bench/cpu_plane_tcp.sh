#!/bin/bash
# BASELINE.json config 2 on the CPU data plane: master + N OS-process workers over loopback TCP (bin/aar_master, bin/aar_worker).
# usage: bench/cpu_plane_tcp.sh [port] [workers=8] [dataSize=16777216] [maxChunkSize=1048576] [maxRound=4]
cd "$(dirname "$0")/.." && mkdir -p build
PORT=${1:-27311}; N=${2:-8}; DS=${3:-16777216}; CH=${4:-1048576}; R=${5:-4}
timeout 300 ./bin/aar_master $PORT $N $DS $CH --set allreduce.max-round=$R --set allreduce.th-reduce=1.0 --set allreduce.th-complete=1.0 --set akka.loglevel=WARNING --timeout 280 > build/cpu_plane_master.log 2>&1 &
MP=$!
sleep 0.3
pids=()
for i in $(seq 1 $N); do
  timeout 300 ./bin/aar_worker 0 $DS --set allreduce.transport.port=$PORT --set akka.loglevel=WARNING --timeout 280 > build/cpu_plane_w$i.log 2>&1 &
  pids+=($!)
done
wait $MP
echo "master rc=$?"; cat build/cpu_plane_master.log | tail -2
for p in "${pids[@]}"; do wait $p; done
tail -1 build/cpu_plane_w1.log
