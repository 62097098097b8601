# usage: tools/ncu_tc.sh TAG  (env RBX_TC_DBG etc. pass through) -> gpurun_out/tc_TAG.ncu-rep + raw / source csv of one fused Gaussian launch
TAG=${1:-x}
RBX_TC_PANEL=0 REPS=1 ncu --set full --clock-control none --import-source on -k regex:sketch_dense_tc_kernel -c 1 -o gpurun_out/tc_$TAG -f python tools/tc_time.py G > gpurun_out/ncu_tc_$TAG.log 2>&1
ncu -i gpurun_out/tc_$TAG.ncu-rep --page raw --csv > gpurun_out/tc_${TAG}_raw.csv
ncu -i gpurun_out/tc_$TAG.ncu-rep --page source --csv --print-source sass > gpurun_out/tc_${TAG}_src.csv
