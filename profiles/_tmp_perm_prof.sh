set -e
python profiles/run_kernel.py --batch 8 --iters 5 -- transpose 2:0:1:3 float32 3:3840:2160:1 | tail -3
ncu --set full --clock-control none --import-source on -s 1 -c 1 -f -o gpurun_out/prof_permvec_201 python profiles/run_kernel.py --batch 8 --iters 3 -- transpose 2:0:1:3 float32 3:3840:2160:1 > gpurun_out/ncu_permvec_201.log 2>&1
