#!/bin/bash
# One-GPU evidence run: the GPU test suite and every bench configuration at N = 1; JSON lines under gpurun_out/final_*.json
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu > gpurun_out/r02_pytest_gpu.log 2>&1; tail -2 gpurun_out/r02_pytest_gpu.log
timeout 400 python bench.py > gpurun_out/final_ngp.json 2> gpurun_out/final_ngp.err; echo "ngp rc=$?"
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_ngp_reference.json 2> gpurun_out/final_ngp_reference.err; echo "ngp reference rc=$?"
timeout 400 python bench.py --config nerf > gpurun_out/final_nerf.json 2> gpurun_out/final_nerf.err; echo "nerf rc=$?"
timeout 400 python bench.py --config neus > gpurun_out/final_neus.json 2> gpurun_out/final_neus.err; echo "neus rc=$?"
timeout 400 python bench.py --config ngp_nerfpp > gpurun_out/final_ngp_nerfpp.json 2> gpurun_out/final_ngp_nerfpp.err; echo "ngp_nerfpp rc=$?"
timeout 400 python bench.py --config mesh > gpurun_out/final_mesh.json 2> gpurun_out/final_mesh.err; echo "mesh rc=$?"
