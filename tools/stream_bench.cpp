/*
 * stream_bench.cpp -- BASELINE configs[4] on ONE GPU through the C-ABI, in native code:
 * N concurrent 1080p pipelines, each
 *   tensor_converter -> tensor_transform typecast:float32 -> arithmetic add:-127.5,div:127.5
 *   -> transpose 1:2:0:3
 * with one host thread and one set of element instances (own CUDA streams) per pipeline, the way
 * GStreamer runs one streaming thread per branch.  bench.py starts one of these per rank (GPU)
 * for the multi-GPU figure: streams are independent, so the ranks never talk to each other.
 *
 * usage: stream_bench DEVICE STREAMS FRAMES MODE [WIDTH [DUMP [START_AT]]]
 *   MODE  resident : the video frame of every stream already sits in HBM
 *         h2d      : every frame is uploaded from pinned host memory first (PCIe in the loop),
 *                    the float CHW result stays on the device for a CUDA-aware consumer
 *         h2d_d2h  : ... and the result is downloaded into pinned host memory as well
 *   WIDTH 1920 (default) or e.g. 1918: rows then carry padding (pitch = a multiple of 4 bytes)
 *         which the converter strips (gsttensor_converter.c:1087-1100)
 *   DUMP  path prefix ("-" = none): DUMP.in receives the video frame of stream 0 and DUMP.out its
 *         last output tensor (float32 CHW), for the parity check bench.py makes against the CPU
 *         reference
 *   START_AT  when the timed region starts, so that the processes of all ranks load the shared
 *         host links at the same time: a wall-clock time (seconds since the epoch, fractional),
 *         or "barrier:DIR:RANK:WORLD" -- after its warm-up the process creates DIR/ready.RANK and
 *         starts once all WORLD such files exist (ranks of one node share /tmp)
 * Nothing in the loop waits for the GPU in the resident mode: memories are stream-ordered
 * (include/nnstreamer_b200.h, nnsb_memory); each thread waits once, for its last output.
 * Prints one JSON object.  Built by __graft_entry__.build().
 */
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <unistd.h>

#include "nnstreamer_b200.h"

#define CHECK(call) \
  do { \
    int r__ = (call); \
    if (r__ < 0) { \
      fprintf (stderr, "%s failed (%d): %s\n", #call, r__, nnsb_last_error ()); \
      exit (1); \
    } \
  } while (0)

static int W = 1920;
static const int H = 1080, C = 3;
enum Mode { RESIDENT, H2D, H2D_D2H };

struct Pipe
{
  nnsb_converter *conv = nullptr;
  nnsb_transform *t[3] = { nullptr, nullptr, nullptr };
  nnsb_memory frame;            /* the video frame (device block or pinned host) */
  nnsb_memory last;             /* last output kept for the final wait / dump */
  void *host_out = nullptr;     /* pinned, h2d_d2h only */
  size_t out_bytes = 0;
};

static std::string
negotiate (nnsb_transform * t, const std::string & incaps)
{
  static thread_local char a[4096], b[4096];
  CHECK (nnsb_transform_transform_caps (t, NNSB_PAD_SINK, incaps.c_str (), nullptr, a, sizeof (a)));
  CHECK (nnsb_transform_fixate_caps (t, NNSB_PAD_SINK, incaps.c_str (), a, b, sizeof (b)));
  CHECK (nnsb_transform_set_caps (t, incaps.c_str (), b));
  return std::string (b);
}

static void
run (Pipe * p, int frames, Mode mode)
{
  for (int i = 0; i < frames; i++) {
    nnsb_memory m[4], outs[NNSB_SIZE_LIMIT];
    int produced = 0;
    uint32_t n = 0;
    if (nnsb_converter_chain (p->conv, &p->frame, &m[0], &produced) != NNSB_FLOW_OK || !produced) {
      fprintf (stderr, "converter: %s\n", nnsb_last_error ());
      exit (1);
    }
    for (int k = 0; k < 3; k++) {
      if (nnsb_transform_transform (p->t[k], &m[k], 1, outs, &n) != NNSB_FLOW_OK || n != 1) {
        fprintf (stderr, "transform %d: %s\n", k, nnsb_last_error ());
        exit (1);
      }
      m[k + 1] = outs[0];
    }
    if (mode == H2D_D2H)
      CHECK (nnsb_transform_download (p->t[2], p->host_out, &m[3]));    /* waits for the result */
    for (int k = 0; k < 3; k++)
      nnsb_memory_release (&m[k]);
    nnsb_memory_release (&p->last);
    p->last = m[3];
  }
  /* the one host wait of this streaming thread */
  CHECK (nnsb_memory_sync (&p->last));
}

static uint64_t
fnv1a (const unsigned char *p, size_t n)
{
  uint64_t h = 1469598103934665603ull;
  for (size_t i = 0; i < n; i++) {
    h ^= p[i];
    h *= 1099511628211ull;
  }
  return h;
}

/* "DIR:RANK:WORLD": announce this rank, wait for the others (at most two minutes).  Returns the
 * seconds this rank waited. */
static double
file_barrier (const char *spec)
{
  std::string s (spec);
  const size_t c2 = s.rfind (':'), c1 = s.rfind (':', c2 - 1);
  if (c2 == std::string::npos || c1 == std::string::npos) {
    fprintf (stderr, "bad barrier spec %s\n", spec);
    exit (2);
  }
  const std::string dir = s.substr (0, c1);
  const int rank = atoi (s.substr (c1 + 1, c2 - c1 - 1).c_str ()), world = atoi (s.substr (c2 + 1).c_str ());
  FILE *f = fopen ((dir + "/ready." + std::to_string (rank)).c_str (), "w");
  if (!f) {
    fprintf (stderr, "cannot write into %s\n", dir.c_str ());
    exit (1);
  }
  fclose (f);
  const auto t0 = std::chrono::steady_clock::now ();
  for (int r = 0; r < world; r++) {
    const std::string name = dir + "/ready." + std::to_string (r);
    for (;;) {
      if (access (name.c_str (), F_OK) == 0)
        break;
      if (std::chrono::duration<double> (std::chrono::steady_clock::now () - t0).count () > 120.0) {
        fprintf (stderr, "rank %d never arrived at the start barrier\n", r);
        exit (1);
      }
      std::this_thread::sleep_for (std::chrono::microseconds (200));
    }
  }
  return std::chrono::duration<double> (std::chrono::steady_clock::now () - t0).count ();
}

int
main (int argc, char **argv)
{
  if (argc < 5) {
    fprintf (stderr, "usage: %s DEVICE STREAMS FRAMES resident|h2d|h2d_d2h [WIDTH [DUMP]]\n", argv[0]);
    return 2;
  }
  const int device = atoi (argv[1]), streams = atoi (argv[2]), frames = atoi (argv[3]);
  Mode mode;
  if (strcmp (argv[4], "resident") == 0)
    mode = RESIDENT;
  else if (strcmp (argv[4], "h2d") == 0)
    mode = H2D;
  else if (strcmp (argv[4], "h2d_d2h") == 0)
    mode = H2D_D2H;
  else {
    fprintf (stderr, "unknown mode %s\n", argv[4]);
    return 2;
  }
  if (argc > 5)
    W = atoi (argv[5]);
  const char *dump = (argc > 6 && strcmp (argv[6], "-") != 0) ? argv[6] : nullptr;
  const bool use_barrier = argc > 7 && strncmp (argv[7], "barrier:", 8) == 0;
  const double start_at = (argc > 7 && !use_barrier) ? atof (argv[7]) : 0.0;
  const size_t row = (size_t) W * C, pitch = (row + 3) / 4 * 4;
  const size_t frame_bytes = pitch * H;         /* GST_VIDEO_INFO_SIZE */
  const size_t elems = row * H;

  nnsb_ctx *ctx = nullptr;
  nnsb_stream *stream = nullptr;
  CHECK (nnsb_ctx_create (device, &ctx));
  CHECK (nnsb_stream_create (ctx, &stream));

  char caps[256], out[4096];
  snprintf (caps, sizeof (caps), "video/x-raw, format=(string)RGB, width=(int)%d, height=(int)%d, "
      "framerate=(fraction)30/1", W, H);
  static const char *modes[3] = { "typecast", "arithmetic", "transpose" };
  static const char *options[3] = { "float32", "add:-127.5,div:127.5", "1:2:0:3" };
  std::vector<Pipe> pipes (streams);
  for (int s = 0; s < streams; s++) {
    Pipe & p = pipes[s];
    CHECK (nnsb_converter_new (device, &p.conv));
    CHECK (nnsb_converter_set_caps (p.conv, caps, out, sizeof (out)));
    std::string c = out;
    for (int k = 0; k < 3; k++) {
      CHECK (nnsb_transform_new (device, &p.t[k]));
      CHECK (nnsb_transform_set_property (p.t[k], "mode", modes[k]));
      CHECK (nnsb_transform_set_property (p.t[k], "option", options[k]));
      c = negotiate (p.t[k], c);
    }
    nnsb_memory host;
    CHECK (nnsb_memory_alloc (ctx, frame_bytes, 0, nullptr, &host));
    unsigned int seed = 12345u + (unsigned int) s;
    for (size_t i = 0; i < frame_bytes; i++) {
      seed = seed * 1664525u + 1013904223u;
      ((unsigned char *) host.data)[i] = (unsigned char) (seed >> 24);
    }
    memset (&p.last, 0, sizeof (p.last));
    p.out_bytes = elems * sizeof (float);
    if (mode == RESIDENT) {
      CHECK (nnsb_memory_alloc (ctx, frame_bytes, 1, stream, &p.frame));
      CHECK (nnsb_h2d_async (stream, p.frame.data, host.data, frame_bytes));
      CHECK (nnsb_stream_sync (stream));
      nnsb_memory_release (&host);
    } else {
      p.frame = host;
    }
    if (mode == H2D_D2H)
      CHECK (nnsb_pinned_alloc (ctx, p.out_bytes, &p.host_out));
  }
  for (auto & p : pipes)
    run (&p, 3, mode);          /* warm-up: contexts, plans, pools */
  double late = 0.0;
  if (use_barrier)
    late = file_barrier (argv[7] + 8);
  else if (start_at > 0.0) {
    const double now = std::chrono::duration<double> (
        std::chrono::system_clock::now ().time_since_epoch ()).count ();
    if (now < start_at)
      std::this_thread::sleep_for (std::chrono::duration<double> (start_at - now));
    else
      late = now - start_at;
  }
  std::vector<std::thread> th;
  const auto t0 = std::chrono::steady_clock::now ();
  for (auto & p : pipes)
    th.emplace_back (run, &p, frames, mode);
  for (auto & t : th)
    t.join ();
  const double dt = std::chrono::duration<double> (std::chrono::steady_clock::now () - t0).count ();
  const double fps = (double) streams * frames / dt;

  /* last output of stream 0: checksum always, raw bytes on request */
  std::vector<unsigned char> last (pipes[0].out_bytes);
  CHECK (nnsb_transform_download (pipes[0].t[2], last.data (), &pipes[0].last));
  const uint64_t sum = fnv1a (last.data (), last.size ());
  if (dump) {
    std::vector<unsigned char> in0 (frame_bytes);
    if (mode == RESIDENT) {
      CHECK (nnsb_memory_sync (&pipes[0].frame));
      CHECK (nnsb_d2h_async (stream, in0.data (), pipes[0].frame.data, frame_bytes));
      CHECK (nnsb_stream_sync (stream));
    } else {
      memcpy (in0.data (), pipes[0].frame.data, frame_bytes);
    }
    const std::string names[2] = { std::string (dump) + ".in", std::string (dump) + ".out" };
    const std::vector<unsigned char> *bufs[2] = { &in0, &last };
    for (int k = 0; k < 2; k++) {
      FILE *f = fopen (names[k].c_str (), "wb");
      if (!f || fwrite (bufs[k]->data (), 1, bufs[k]->size (), f) != bufs[k]->size ()) {
        fprintf (stderr, "cannot write %s\n", names[k].c_str ());
        return 1;
      }
      fclose (f);
    }
  }
  uint64_t launches = 0;
  for (int k = 0; k < 3; k++) {
    uint64_t c = 0;
    nnsb_transform_kernel_launches (pipes[0].t[k], &c);
    launches += c;
  }
  /* algorithmic bytes per element along the chain: typecast 5, arithmetic 8, transpose 8
   * (+ 2 for the strip when rows are padded) */
  const double bpe = 21.0 + ((pitch != row && mode == RESIDENT) ? 2.0 : 0.0);
  printf ("{\"frames/s\": %.1f, \"GB/s\": %.1f, \"streams\": %d, \"frames_per_stream\": %d, "
      "\"mode\": \"%s\", \"width\": %d, \"seconds\": %.4f, \"h2d_bytes_per_frame\": %zu, "
      "\"d2h_bytes_per_frame\": %zu, \"stream0_last_output_fnv1a\": \"%016llx\", "
      "\"stream0_kernel_launches\": %llu, \"start_skew_s\": %.3f, "
      "\"driver\": \"native threads through the C-ABI (tools/stream_bench.cpp)\"}\n",
      fps, fps * (double) elems * bpe / 1e9, streams, frames, argv[4], W, dt,
      mode == RESIDENT ? (size_t) 0 : frame_bytes, mode == H2D_D2H ? pipes[0].out_bytes : (size_t) 0,
      (unsigned long long) sum, (unsigned long long) launches, late);
  for (auto & p : pipes) {
    nnsb_memory_release (&p.last);
    nnsb_memory_release (&p.frame);
    nnsb_converter_free (p.conv);
    for (int k = 0; k < 3; k++)
      nnsb_transform_free (p.t[k]);
  }
  nnsb_stream_destroy (stream);
  nnsb_ctx_destroy (ctx);
  return 0;
}
