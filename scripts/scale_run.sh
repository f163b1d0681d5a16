#!/bin/bash
# Runs every bench configuration at N GPUs of one box (torchrun, one rank per GPU) and stores the JSON lines under
# gpurun_out/scale_N<N>_<config>.json.  Usage: scripts/scale_run.sh N [configs...]
N=${1:-2}; shift
CFGS=${@:-"ngp nerf neus ngp_nerfpp mesh"}
mkdir -p gpurun_out
run() {   # name, env, args...
    name=$1; envs=$2; shift 2
    if [ "$N" = "1" ]; then
        env $envs timeout 400 python bench.py --gpus 1 "$@" > gpurun_out/scale_N${N}_${name}.json 2> gpurun_out/scale_N${N}_${name}.err
    else
        env $envs timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) \
            bench.py --gpus $N "$@" > gpurun_out/scale_N${N}_${name}.json 2> gpurun_out/scale_N${N}_${name}.err
    fi
    echo "== $name rc=$?"; tail -c 600 gpurun_out/scale_N${N}_${name}.json | head -c 600; echo; grep -v "Warning\|warn\|^$\|OMP_NUM" gpurun_out/scale_N${N}_${name}.err | tail -4
}
for c in $CFGS; do
    case $c in
        ngp)  run ngp_peer16 ARCB200_EXCHANGE=peer16 --config ngp --steps 200 --warmup 20 --no-cpu-baseline ;;
        ngp_variants)
              run ngp_mc16 ARCB200_EXCHANGE=mc16 --config ngp --steps 200 --warmup 20 --no-cpu-baseline --no-render
              run ngp_nccl ARCB200_EXCHANGE=nccl --config ngp --steps 200 --warmup 20 --no-cpu-baseline --no-render ;;
        nerf) run nerf X=1 --config nerf --steps 20 --warmup 3 --no-cpu-baseline ;;
        neus) run neus X=1 --config neus --steps 20 --warmup 3 ;;
        ngp_nerfpp) run ngp_nerfpp X=1 --config ngp_nerfpp --steps 20 --warmup 3 ;;
        mesh) run mesh X=1 --config mesh --steps 5 --warmup 3 --no-cpu-baseline ;;
    esac
done
