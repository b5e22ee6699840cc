#!/bin/bash
# CPU data plane over loopback TCP with the contributor-count check on (bin/aar_worker --check): one configuration per call.
# usage: bench/cpu_plane_stress.sh port workers dataSize maxChunkSize maxRound threshold maxLag pipelineDepth   (LOGLEVEL=INFO for logs)
# Prints one line; workers_bad=1 means a worker saw a mismatch, a lost master or a non-zero exit.
cd "$(dirname "$0")/.." && mkdir -p build
PORT=$1; N=$2; DS=$3; CH=$4; R=$5; TH=$6; LAG=$7; D=$8
timeout 120 ./bin/aar_master $PORT $N $DS $CH --set allreduce.max-round=$R --set allreduce.th-allreduce=$TH --set allreduce.th-reduce=$TH --set allreduce.th-complete=$TH --set allreduce.max-lag=$LAG --set allreduce.pipeline-depth=$D --set akka.loglevel=${LOGLEVEL:-ERROR} --timeout 100 > build/st_m.log 2>&1 &
MP=$!
sleep 0.2
pids=()
for i in $(seq 1 $N); do
  timeout 120 ./bin/aar_worker 0 $DS --check --set allreduce.transport.port=$PORT --set akka.loglevel=${LOGLEVEL:-ERROR} --timeout 100 > build/st_w$i.log 2>&1 &
  pids+=($!)
done
wait $MP; mrc=$?
bad=0
for i in $(seq 1 $N); do wait ${pids[$((i-1))]}; rc=$?; if [ $rc -ne 0 ]; then bad=1; echo "worker $i rc=$rc: $(tail -1 build/st_w$i.log)"; fi; done
echo "N=$N ds=$DS ch=$CH R=$R th=$TH lag=$LAG depth=$D master_rc=$mrc $(tail -1 build/st_m.log) workers_bad=$bad"
