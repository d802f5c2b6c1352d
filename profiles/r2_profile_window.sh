# ncu of the windowed slab kernel (output larger than the L2: windows on by default)
run() { name=$1; shift; rx=$1; shift
  python profiles/run_kernel.py "$@" > gpurun_out/plain_$name.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:$rx -s 1 -c 1 -f -o gpurun_out/prof_$name python profiles/run_kernel.py "$@" > gpurun_out/ncu_$name.log 2>&1; tail -1 gpurun_out/plain_$name.log; }
mkdir -p gpurun_out
run slab2_f32_210_b512_window slab_ --batch 512 --iters 3 -- transpose 2:1:0:3 float32 3:227:227:1
run slab2_u16_021_b2048_window slab_ --batch 2048 --iters 3 -- transpose 0:2:1:3 uint16 3:225:113:1
