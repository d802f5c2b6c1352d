/*
 * agg_bench.cpp -- tensor_aggregator on 1080p RGB uint8 frames driven by ONE native streaming
 * thread through the C-ABI, the way GStreamer drives the element (gst_tensor_aggregator_chain,
 * gsttensor_aggregator.c:835-941).  bench.py also times the same calls made from Python; at
 * these sizes that figure is the interpreter's call rate, this one is the element's.
 *
 * usage: agg_bench DEVICE FRAMES_OUT FRAMES_DIM OUTPUTS
 *   every input buffer is one frame (frames-in = 1) that the producer has written into the
 *   memory the aggregator handed out (nnsb_aggregator_acquire: the tail of its ring), so pushing
 *   moves nothing; FRAMES_DIM 3 = append (an output is a share of the ring: zero bytes moved),
 *   FRAMES_DIM 1 = concatenate along the width (one strided copy launch per output).
 * Nothing in the loop waits for the GPU; the thread waits once, for its last output.
 * Prints one JSON object.  Built by __graft_entry__.build().
 */
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "nnstreamer_b200.h"

#define CHECK(call) \
  do { \
    int r__ = (call); \
    if (r__ < 0) { \
      fprintf (stderr, "%s failed (%d): %s\n", #call, r__, nnsb_last_error ()); \
      exit (1); \
    } \
  } while (0)

static nnsb_memory last;

static int
feed (nnsb_aggregator * a, size_t frame_bytes, int frames)
{
  int produced = 0;
  for (int i = 0; i < frames; i++) {
    nnsb_memory slot, outs[4];
    uint32_t n = 0;
    CHECK (nnsb_aggregator_acquire (a, frame_bytes, &slot));
    if (nnsb_aggregator_chain (a, &slot, nullptr, outs, nullptr, 4, &n) != NNSB_FLOW_OK) {
      fprintf (stderr, "chain: %s\n", nnsb_last_error ());
      exit (1);
    }
    nnsb_memory_release (&slot);
    for (uint32_t k = 0; k < n; k++) {
      nnsb_memory_release (&last);
      last = outs[k];
      produced++;
    }
  }
  return produced;
}

int
main (int argc, char **argv)
{
  if (argc < 5) {
    fprintf (stderr, "usage: %s DEVICE FRAMES_OUT FRAMES_DIM OUTPUTS\n", argv[0]);
    return 2;
  }
  const int device = atoi (argv[1]), frames_out = atoi (argv[2]), frames_dim = atoi (argv[3]);
  const int outputs = atoi (argv[4]);
  const size_t frame_bytes = (size_t) 3 * 1920 * 1080;
  nnsb_aggregator *a = nullptr;
  char out[4096], num[32];
  CHECK (nnsb_aggregator_new (device, &a));
  snprintf (num, sizeof (num), "%d", frames_out);
  CHECK (nnsb_aggregator_set_property (a, "frames-out", num));
  snprintf (num, sizeof (num), "%d", frames_dim);
  CHECK (nnsb_aggregator_set_property (a, "frames-dim", num));
  CHECK (nnsb_aggregator_set_caps (a, "other/tensor, dimension=(string)3:1920:1080:1, "
          "type=(string)uint8, framerate=(fraction)30/1", out, sizeof (out)));
  memset (&last, 0, sizeof (last));
  feed (a, frame_bytes, 2 * frames_out);        /* warm-up: two outputs */
  CHECK (nnsb_memory_sync (&last));
  uint64_t l0 = 0, l1 = 0, cin = 0, cout = 0;
  nnsb_aggregator_kernel_launches (a, &l0);
  const int frames = outputs * frames_out;
  const auto t0 = std::chrono::steady_clock::now ();
  const int produced = feed (a, frame_bytes, frames);
  CHECK (nnsb_memory_sync (&last));
  const double dt = std::chrono::duration<double> (std::chrono::steady_clock::now () - t0).count ();
  nnsb_aggregator_kernel_launches (a, &l1);
  nnsb_aggregator_copy_counters (a, &cin, &cout);
  nnsb_memory_release (&last);
  printf ("{\"frames/s\": %.1f, \"outputs\": %d, \"kernel_launches\": %llu, "
      "\"bytes_copied_into_the_adapter\": %llu, \"bytes_moved_by_concatenation\": %llu, "
      "\"seconds\": %.5f, \"GB/s\": %.1f, "
      "\"driver\": \"one native thread through the C-ABI (tools/agg_bench.cpp), frames written in "
      "place (nnsb_aggregator_acquire), one wait at the end\"}\n",
      frames / dt, produced, (unsigned long long) (l1 - l0), (unsigned long long) cin,
      (unsigned long long) cout, dt, frames_dim == 3 ? 0.0 : 2.0 * frame_bytes * frames / dt / 1e9);
  nnsb_aggregator_free (a);
  return 0;
}
