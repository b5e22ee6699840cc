set -u
mkdir -p gpurun_out
python bench/ncu_target.py bulk 8 > gpurun_out/AB_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:bulkRoundKernel -s 48 -c 27 -o gpurun_out/r2_ncu_bulk -f python bench/ncu_target.py bulk 8 > gpurun_out/AB_ncu.log 2>&1
tail -3 gpurun_out/AB_plain.log gpurun_out/AB_ncu.log
ls -la gpurun_out/*.ncu-rep
