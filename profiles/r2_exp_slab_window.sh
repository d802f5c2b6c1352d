# windowed slabs (NNSB_SLAB_WINDOW=1): correctness, a memcheck pass, timings against the default
python -m pytest tests/test_gpu_perm_slab.py -x -q 2>&1 | tail -3
NNSB_SLAB_WINDOW=1 timeout 300 compute-sanitizer --tool memcheck --error-exitcode 9 python profiles/run_kernel.py --batch 3 --iters 1 -- transpose 2:1:0:3 float32 3:227:227:1 2>&1 | tail -4
NNSB_SLAB_WINDOW=1 timeout 300 compute-sanitizer --tool memcheck --error-exitcode 9 python profiles/run_kernel.py --batch 3 --iters 1 -- transpose 0:2:1:3 uint16 3:225:113:1 2>&1 | tail -4
one() { echo "## $*"; env "$1" python profiles/sweep_perm.py --one "${@:2}" 2>&1 | tail -1; }
for w in 0 1; do
  one NNSB_SLAB_WINDOW=$w transpose 2:1:0:3 float32 3:227:227:1 64
  one NNSB_SLAB_WINDOW=$w transpose 2:1:0:3 float32 3:227:227:1 512
  one NNSB_SLAB_WINDOW=$w transpose 0:2:1:3 uint16 3:225:113:1 256
  one NNSB_SLAB_WINDOW=$w transpose 0:2:1:3 uint16 3:225:113:1 2048
  one NNSB_SLAB_WINDOW=$w transpose 1:2:0:3 uint16 3:225:113:1 256
done
