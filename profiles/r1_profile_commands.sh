# Round-1 per-kernel ncu captures (run on the GPU box through gpurun; one kernel family each).
# Each command first runs plainly (exit 0 required), then under ncu; -s 1 skips the cold launch.
set -e
run() { # name, run_kernel.py args...
  name=$1; shift
  python profiles/run_kernel.py "$@" > gpurun_out/plain_$name.log 2>&1 && ncu --set full --clock-control none --import-source on -s 1 -c 2 -f -o gpurun_out/prof_$name python profiles/run_kernel.py "$@" > gpurun_out/ncu_$name.log 2>&1
  tail -1 gpurun_out/plain_$name.log
}
run plane_u8 --batch 16 --iters 4 -- transpose 1:2:0:3 uint8 3:3840:2160:1
run clamp --iters 4 -- clamp -33.3:77.7 float32 3:1920:1080:64
run typecast --iters 4 -- typecast float32 uint8 3:1920:1080:32
run permute_vec --batch 8 --iters 4 -- transpose 2:0:1:3 float32 3:3840:2160:1
run fchain_f32_perch --batch 16 --iters 4 -- arithmetic "per-channel:true@0,add:-123.675@0,add:-116.28@1,add:-103.53@2,div:58.395@0,div:57.12@1,div:57.375@2" float32 3:1920:1080:1
# launch list of the whole bench (r1_launches_bench.csv):
# python bench.py --steps 2 --warmup 3 --no-cpu-baseline && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline
