/*
 * perm_table_check.cu -- host-only check of the vector-permute plan builder (k_perm_vec.cu).
 *
 * TEST INFRASTRUCTURE (built and run by tests/test_perm_table_host.py; needs no GPU): includes
 * the kernel source to reach its host-side builder, builds the box table for a permutation
 * given on the command line, then replays on the CPU exactly what the kernel does with that
 * table (tile decode, predicates, swizzled slots) and compares the result with the plain
 * definition out[..] = in[permuted ..] of gsttensor_transform.c:1417-1500.
 *
 * usage: perm_table_check ES BATCH D0 D1 D2 D3 O0 O1 O2 O3      (transpose order O, dims D)
 * prints one line: "ok ..." / "ineligible" / "MISMATCH ..."
 */
#include <cstdio>
#include <vector>

#include "../../nnstreamer_b200/csrc/k_perm_vec.cu"

using namespace nnsb;

int
main (int argc, char **argv)
{
  if (argc != 11) {
    fprintf (stderr, "usage: %s ES BATCH D0 D1 D2 D3 O0 O1 O2 O3\n", argv[0]);
    return 2;
  }
  const int es = atoi (argv[1]);
  const uint32_t batch = (uint32_t) atoi (argv[2]);
  uint32_t dim[4];
  int order[4];
  for (int i = 0; i < 4; i++) {
    dim[i] = (uint32_t) atoi (argv[3 + i]);
    order[i] = atoi (argv[7 + i]);
  }
  uint64_t in_stride[4], s = 1;
  for (int i = 0; i < 4; i++) {
    in_stride[i] = s;
    s *= dim[i];
  }
  const uint64_t frame_elems = s;
  PermDesc d;
  memset (&d, 0, sizeof (d));
  d.rank = 4;
  d.es = es;
  for (int i = 0; i < 4; i++) {
    d.size[i] = dim[order[i]];
    d.in_stride[i] = in_stride[order[i]];
  }
  PermVecPlan pl;
  void *table = nullptr;
  if (perm_vec_build (d, frame_elems, &pl, &table) != 0) {
    printf ("ineligible\n");
    return 0;
  }
  const unsigned char *tab = (const unsigned char *) table;
  const uint32_t *t_ld_off = (const uint32_t *) (tab + PV_T_LD_OFF);
  const uint32_t *t_ld_slot = (const uint32_t *) (tab + PV_T_LD_SLOT);
  const uint32_t *t_st_off = (const uint32_t *) (tab + PV_T_ST_OFF);
  const uint64_t *t_ld_coord = (const uint64_t *) (tab + PV_T_LD_COORD);
  const uint64_t *t_st_coord = (const uint64_t *) (tab + PV_T_ST_COORD);
  const uint16_t *t_st_slot = (const uint16_t *) (tab + PV_T_ST_SLOT);
  const int VB = pl.vb, VE = VB / pl.es;

  /* input: byte b of the tensor = a hash of its position; reference output by definition */
  const size_t frame_bytes = (size_t) frame_elems * es, total = frame_bytes * batch;
  std::vector<unsigned char> in (total), want (total), got (total, 0xEE);
  for (size_t i = 0; i < total; i++)
    in[i] = (unsigned char) ((i * 2654435761u) >> 13);
  for (uint32_t f = 0; f < batch; f++) {
    uint64_t o = 0;
    for (uint32_t c3 = 0; c3 < d.size[3]; c3++)
      for (uint32_t c2 = 0; c2 < d.size[2]; c2++)
        for (uint32_t c1 = 0; c1 < d.size[1]; c1++)
          for (uint32_t c0 = 0; c0 < d.size[0]; c0++, o++) {
            const uint64_t src = c0 * d.in_stride[0] + c1 * d.in_stride[1] + c2 * d.in_stride[2]
                + c3 * d.in_stride[3];
            memcpy (&want[f * frame_bytes + o * es], &in[f * frame_bytes + src * es], es);
          }
  }

  /* replay of permute_vec_kernel, one tile at a time */
  int n = pl.n;
  uint32_t tile[PV_DIMS], size[PV_DIMS], ntile[PV_DIMS];
  uint64_t isb[PV_DIMS], osb[PV_DIMS];
  int box_slot[PV_DIMS];
  for (int i = 0; i < n; i++) {
    tile[i] = pl.tile[i];
    size[i] = pl.size[i];
    ntile[i] = (size[i] + tile[i] - 1) / tile[i];
    isb[i] = pl.in_stride[i] * pl.es;
    osb[i] = pl.out_stride[i] * pl.es;
    box_slot[i] = pl.box_slot[i];
  }
  if (batch > 1) {
    tile[n] = 1;
    size[n] = ntile[n] = batch;
    isb[n] = osb[n] = pl.frame_bytes;
    box_slot[n] = -1;
    n++;
  }
  const uint64_t n_tiles = pl.tiles_per_frame * batch;
  std::vector<unsigned char> stage ((size_t) PV_NVT * VB);
  auto inside = [] (uint64_t c, uint64_t lim) {
    for (int b = 0; b < PV_BOXED; b++)
      if (((c >> (16 * b)) & 0xffff) >= ((lim >> (16 * b)) & 0xffff))
        return false;
    return true;
  };
  size_t writes = 0;
  for (uint64_t t = 0; t < n_tiles; t++) {
    uint64_t q = t, ib = 0, ob = 0, lim = pl.limits_init;
    for (int dd = 0; dd < n; dd++) {
      const uint32_t td = (uint32_t) (q % ntile[dd]);
      q /= ntile[dd];
      const uint32_t o = td * tile[dd];
      ib += (uint64_t) o * isb[dd];
      ob += (uint64_t) o * osb[dd];
      const uint32_t left = size[dd] - o;
      const uint32_t l = left < tile[dd] ? left : tile[dd];
      if (box_slot[dd] >= 0)
        lim |= (uint64_t) l << (16 * box_slot[dd]);
    }
    std::fill (stage.begin (), stage.end (), 0xCD);
    for (uint32_t v = 0; v < (uint32_t) PV_NVT; v++)
      if (t_ld_off[v] != PV_INVALID && inside (t_ld_coord[v], lim)) {
        if (ib + t_ld_off[v] + VB > total || t_ld_slot[v] + VB > (uint32_t) PV_NVT * VB) {
          printf ("MISMATCH load out of range (tile %llu vector %u)\n", (unsigned long long) t, v);
          return 1;
        }
        memcpy (&stage[t_ld_slot[v]], &in[ib + t_ld_off[v]], VB);
      }
    for (uint32_t v = 0; v < (uint32_t) PV_NVT; v++)
      if (t_st_off[v] != PV_INVALID && inside (t_st_coord[v], lim)) {
        if (ob + t_st_off[v] + VB > total) {
          printf ("MISMATCH store out of range (tile %llu vector %u)\n", (unsigned long long) t, v);
          return 1;
        }
        if (pl.micro) {
          /* entry v = thread + j * 256: vector jj = j % VE of that thread's block j / VE; element
           * e comes from chunk e of the block (slot kept in entry (block, e)), position jj */
          const uint32_t t = v % PV_THREADS, j = v / PV_THREADS, b = j / VE, jj = j % VE;
          for (int e = 0; e < VE; e++) {
            const uint32_t entry = t + (b * VE + e) * PV_THREADS;
            memcpy (&got[ob + t_st_off[v] + (size_t) e * pl.es],
                &stage[t_st_slot[(size_t) entry * VE] + (size_t) jj * pl.es], pl.es);
          }
        } else {
          for (int e = 0; e < VE; e++)
            memcpy (&got[ob + t_st_off[v] + (size_t) e * pl.es], &stage[t_st_slot[(size_t) v * VE + e]],
                pl.es);
        }
        writes += VB;
      }
  }
  if (writes != total) {
    printf ("MISMATCH wrote %zu of %zu bytes\n", writes, total);
    return 1;
  }
  for (size_t i = 0; i < total; i++)
    if (got[i] != want[i]) {
      printf ("MISMATCH at byte %zu\n", i);
      return 1;
    }
  char box[128];
  int o = 0;
  for (int i = 0; i < pl.n; i++)
    o += snprintf (box + o, sizeof (box) - o, "%s%u", i ? "x" : "", pl.tile[i]);
  printf ("ok vb=%d es=%d box=%s n_vec=%u swizzle=%d/%d micro=%d wavefronts=%llu tiles=%llu\n", pl.vb, pl.es, box,
      pl.n_vec, pl.swizzle_shift, pl.swizzle_shift2, pl.micro, (unsigned long long) pl.gather_wavefronts,
      (unsigned long long) n_tiles);
  free (table);
  return 0;
}
