# slab kernel: correctness and the five reference timings
python -m pytest tests/test_gpu_perm_slab.py -x -q 2>&1 | tail -2
one() { echo "## $*"; python profiles/sweep_perm.py --one "$@" 2>&1 | tail -1; }
one transpose 0:2:1:3 uint16 3:225:113:1 256
one transpose 0:2:1:3 uint16 3:225:113:1 2048
one transpose 2:1:0:3 float32 3:227:227:1 64
one transpose 2:1:0:3 float32 3:227:227:1 512
one transpose 1:2:0:3 uint16 3:225:113:1 256
