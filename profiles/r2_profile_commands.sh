# Round-2 per-kernel ncu captures (run on the GPU box through gpurun; one kernel family each).
# Each command first runs plainly (exit 0 required), then under ncu; -s 1 skips the cold launch.
# Summaries: python profiles/summarize.py gpurun_out/prof_<name>.ncu-rep profiles/r2_<name>
run() { # name, kernel regex, run_kernel.py args...
  name=$1; shift; rx=$1; shift
  python profiles/run_kernel.py "$@" > gpurun_out/plain_$name.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:$rx -s 1 -c 1 -f -o gpurun_out/prof_$name python profiles/run_kernel.py "$@" > gpurun_out/ncu_$name.log 2>&1
  tail -1 gpurun_out/plain_$name.log
}
mkdir -p gpurun_out
# stand: fused single-read kernel (1 and 3 channels) and the three-launch kernels (one 400 MB tensor)
[ -n "$SKIP_STAND" ] || run stand_fused stand_fused --batch 64 --iters 3 -- stand default float32 3:1920:1080:1
[ -n "$SKIP_STAND" ] || run stand_fused_perch stand_fused --batch 64 --iters 3 -- stand default,per-channel:true float32 3:1920:1080:1
[ -n "$SKIP_STAND" ] || NNSB_STAND_NO_FUSED=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:stand_ -c 3 -f -o gpurun_out/prof_stand_three_launch python profiles/run_kernel.py stand default float32 3:1920:1080:64 --iters 1 > gpurun_out/ncu_stand3.log 2>&1
# headline kernel (bench workload), planar f32, strip, merge, permutes
run lut_chain_u8_f32 lut_chain --batch 32 --iters 3 -- arithmetic "typecast:float32,per-channel:true@0,add:-123.675@0,add:-116.28@1,add:-103.53@2,div:58.395@0,div:57.12@1,div:57.375@2" uint8 3:1920:1080:1
run plane_f32 plane_kernel --batch 8 --iters 3 -- transpose 1:2:0:3 float32 3:3840:2160:1
run pixel_rgb_u8 pixel_transpose --batch 32 --iters 3 -- transpose 0:2:1:3 uint8 3:1920:1080:1
run slab_u16 permute_slab --batch 256 --iters 3 -- transpose 1:2:0:3 uint16 3:225:113:1
run permute_vec4_f32 permute_vec --batch 64 --iters 3 -- transpose 2:1:0:3 float32 3:227:227:1
NNSB_PERM_NO_VEC=1 NNSB_PERM_SLAB=0 run permute_scalar permute_kernel --batch 8 --iters 3 -- transpose 2:0:1:3 float32 3:3840:2160:1
# strip + merge have no plan: profiles/run_misc.py launches them
python profiles/run_misc.py > gpurun_out/plain_misc.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:"repitch|merge_phase" -s 2 -c 2 -f -o gpurun_out/prof_misc python profiles/run_misc.py > gpurun_out/ncu_misc.log 2>&1
# launch list of the whole bench (r2_launches_bench.csv); children (tools/stream_bench) are not followed
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 && timeout 600 ncu --target-processes application-only --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
