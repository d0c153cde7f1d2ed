# usage: bash tools/gpu_prof.sh <kernel id> <tag> [extra bench args]   -> gpurun_out/prof_<tag>.ncu-rep
mkdir -p gpurun_out
K=${1:-3}; TAG=${2:-k$K}; shift; shift
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --kernel $K "$@" > gpurun_out/plain_$TAG.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:admm_ -s 3 -c 1 -o gpurun_out/prof_$TAG -f python bench.py --steps 2 --warmup 3 --no-cpu --kernel $K "$@" > gpurun_out/ncu_$TAG.log 2>&1
echo "ncu full rc=$?"
