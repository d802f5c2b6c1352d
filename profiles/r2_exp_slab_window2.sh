# windowed slabs: rows per slab (NNSB_SLAB_WIN_ROWS) against segment length
one() { echo "## $*"; env "$1" "$2" python profiles/sweep_perm.py --one "${@:3}" 2>&1 | tail -1; }
for rows in 114 76 57 38; do
  one NNSB_SLAB_WINDOW=1 NNSB_SLAB_WIN_ROWS=$rows transpose 2:1:0:3 float32 3:227:227:1 64
  one NNSB_SLAB_WINDOW=1 NNSB_SLAB_WIN_ROWS=$rows transpose 2:1:0:3 float32 3:227:227:1 512
done
for rows in 57 38; do
  one NNSB_SLAB_WINDOW=1 NNSB_SLAB_WIN_ROWS=$rows transpose 0:2:1:3 uint16 3:225:113:1 256
  one NNSB_SLAB_WINDOW=1 NNSB_SLAB_WIN_ROWS=$rows transpose 0:2:1:3 uint16 3:225:113:1 2048
done
