run() { name=$1; shift
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:slab_ -s 1 -c 1 -f -o gpurun_out/prof_$name python profiles/run_kernel.py "$@" > gpurun_out/ncu_$name.log 2>&1; }
run slab2_u16_120 --batch 256 --iters 3 -- transpose 1:2:0:3 uint16 3:225:113:1
run slab2_u16_021 --batch 256 --iters 3 -- transpose 0:2:1:3 uint16 3:225:113:1
run slab2_f32_210 --batch 64 --iters 3 -- transpose 2:1:0:3 float32 3:227:227:1
ls -la gpurun_out/prof_slab2*
# yardstick: an element-wise kernel moving the same bytes (39 MB in, 39 MB out)
python profiles/sweep_perm.py --one arithmetic add:0 uint16 3:225:113:1 256 > gpurun_out/yardstick_copy_u16.txt 2>&1
python profiles/sweep_perm.py --one arithmetic add:0 float32 3:227:227:1 64 >> gpurun_out/yardstick_copy_u16.txt 2>&1
cat gpurun_out/yardstick_copy_u16.txt
