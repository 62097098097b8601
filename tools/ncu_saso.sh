# usage: tools/ncu_saso.sh TAG "R C"  -> gpurun_out/saso_TAG_{R,C}.ncu-rep + raw csv (one saso_tile_kernel launch each, ncu --set full)
TAG=${1:-x}; LAYS=${2:-"R C"}
for L in $LAYS; do
ncu --set full --clock-control none --import-source on -k regex:saso_tile_kernel -c 1 -o gpurun_out/saso_${TAG}_$L -f python bench.py --steps 1 --warmup 1 --layout $L --no-extras --no-cpu --e2e-steps 0 > gpurun_out/ncu_saso_$L.log 2>&1
ncu -i gpurun_out/saso_${TAG}_$L.ncu-rep --page raw --csv > gpurun_out/saso_${TAG}_${L}_raw.csv
ncu -i gpurun_out/saso_${TAG}_$L.ncu-rep --page source --csv --print-source sass > gpurun_out/saso_${TAG}_${L}_src.csv
done
