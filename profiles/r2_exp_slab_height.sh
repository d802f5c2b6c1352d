one() { echo "## $*"; env "$1" "$2" python profiles/sweep_perm.py --one "${@:3}" 2>&1 | tail -1; }
one A=1 B=1 transpose 2:1:0:3 float32 3:227:227:1 64
one A=1 B=1 transpose 2:1:0:3 float32 3:227:227:1 512
one NNSB_SLAB_KB=48 B=1 transpose 2:1:0:3 float32 3:227:227:1 64
one NNSB_SLAB_KB=48 B=1 transpose 2:1:0:3 float32 3:227:227:1 512
one NNSB_SLAB_KB=24 B=1 transpose 2:1:0:3 float32 3:227:227:1 512
one NNSB_SLAB_CTAS=3 B=1 transpose 2:1:0:3 float32 3:227:227:1 512
one A=1 B=1 transpose 0:2:1:3 uint16 3:225:113:1 256
one NNSB_SLAB_FOLD=1 B=1 transpose 0:2:1:3 uint16 3:225:113:1 256
one NNSB_SLAB_FOLD=1 NNSB_SLAB_KB=48 transpose 0:2:1:3 uint16 3:225:113:1 256
one NNSB_SLAB_KB=48 B=1 transpose 0:2:1:3 uint16 3:225:113:1 256
one A=1 B=1 transpose 1:2:0:3 uint16 3:225:113:1 256
python -m pytest tests/test_gpu_perm_slab.py -x -q 2>&1 | tail -2
