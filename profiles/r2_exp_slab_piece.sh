# slab load as one bulk copy or in pieces (NNSB_SLAB_PIECE bytes per copy)
one() { echo "## $*"; env "$1" python profiles/sweep_perm.py --one "${@:2}" 2>&1 | tail -1; }
for piece in 1000000 16384 4096 2048 1024; do
  one NNSB_SLAB_PIECE=$piece transpose 2:1:0:3 float32 3:227:227:1 64
  one NNSB_SLAB_PIECE=$piece transpose 2:1:0:3 float32 3:227:227:1 512
  one NNSB_SLAB_PIECE=$piece transpose 1:2:0:3 uint16 3:225:113:1 256
  one NNSB_SLAB_PIECE=$piece transpose 0:2:1:3 uint16 3:225:113:1 256
done
python -m pytest tests/test_gpu_perm_slab.py -x -q 2>&1 | tail -2
