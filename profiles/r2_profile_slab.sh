# Round-2 captures of the narrow-row permute kernels in their final form (k_perm_slab.cu lane
# kernel, k_perm_pix.cu staged pixel transpose).  Run on the GPU box through gpurun; summaries:
#   python profiles/summarize.py gpurun_out/prof_<name>.ncu-rep profiles/r2_<name>
run() { name=$1; shift; rx=$1; shift
  python profiles/run_kernel.py "$@" > gpurun_out/plain_$name.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:$rx -s 1 -c 1 -f -o gpurun_out/prof_$name python profiles/run_kernel.py "$@" > gpurun_out/ncu_$name.log 2>&1; tail -1 gpurun_out/plain_$name.log; }
mkdir -p gpurun_out
run slab2_u16_120 slab_ --batch 256 --iters 3 -- transpose 1:2:0:3 uint16 3:225:113:1
run slab2_u16_021 slab_ --batch 256 --iters 3 -- transpose 0:2:1:3 uint16 3:225:113:1
run slab2_f32_210 slab_ --batch 64 --iters 3 -- transpose 2:1:0:3 float32 3:227:227:1
run slab2_f32_210_b512 slab_ --batch 512 --iters 3 -- transpose 2:1:0:3 float32 3:227:227:1
run pixel_rgb_u8 pixel_transpose --batch 32 --iters 3 -- transpose 0:2:1:3 uint8 3:1920:1080:1
