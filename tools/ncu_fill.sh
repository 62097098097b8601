# tools/ncu_fill.sh TAG -> gpurun_out/fill_TAG_{U,G}_raw.csv: ncu --set full of one linear-loop fill_dense launch per family (2^30 floats)
TAG=${1:-x}
ncu --set full --clock-control none --import-source on -k regex:fill_dense_kernel -s 2 -c 1 -o gpurun_out/fill_${TAG}_U -f python tools/micro/fill_time.py > gpurun_out/ncu_fill_U.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fill_dense_kernel -s 10 -c 1 -o gpurun_out/fill_${TAG}_G -f python tools/micro/fill_time.py > gpurun_out/ncu_fill_G.log 2>&1
for F in U G; do ncu -i gpurun_out/fill_${TAG}_$F.ncu-rep --page raw --csv > gpurun_out/fill_${TAG}_${F}_raw.csv; done
