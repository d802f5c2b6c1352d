/*
 * stream_bench.cpp -- BASELINE configs[4] on ONE GPU through the C-ABI, in native code:
 * N concurrent 1080p pipelines, each
 *   tensor_converter -> tensor_transform typecast:float32 -> arithmetic add:-127.5,div:127.5
 *   -> transpose 1:2:0:3
 * with one host thread and one set of element instances (own CUDA streams) per pipeline, the way
 * GStreamer runs one streaming thread per branch.  This is what a pipeline built from the
 * reference-facing element mirrors costs without the Python interpreter in the loop (bench.py's
 * `chain_8_streams...` entry drives the same calls from Python threads).
 *
 * usage: stream_bench DEVICE STREAMS FRAMES MODE     MODE = resident | h2d
 * prints one JSON object.  Built by __graft_entry__.build(): g++ -O2 tools/stream_bench.cpp
 * -Iinclude -Lnnstreamer_b200 -lnnstreamer_b200 -lpthread.
 */
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "nnstreamer_b200.h"

#define CHECK(call) \
  do { \
    int r__ = (call); \
    if (r__ < 0) { \
      fprintf (stderr, "%s failed (%d): %s\n", #call, r__, nnsb_last_error ()); \
      exit (1); \
    } \
  } while (0)

static const int W = 1920, H = 1080, C = 3;
static const size_t FRAME = (size_t) W * H * C;

struct Pipe
{
  nnsb_converter *conv = nullptr;
  nnsb_transform *t[3] = { nullptr, nullptr, nullptr };
  nnsb_memory frame;
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
run (Pipe * p, int frames)
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
    for (int k = 0; k < 4; k++)
      nnsb_memory_release (&m[k]);
  }
}

int
main (int argc, char **argv)
{
  if (argc != 5) {
    fprintf (stderr, "usage: %s DEVICE STREAMS FRAMES resident|h2d\n", argv[0]);
    return 2;
  }
  const int device = atoi (argv[1]), streams = atoi (argv[2]), frames = atoi (argv[3]);
  const bool resident = strcmp (argv[4], "resident") == 0;
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
    void *host = nullptr;
    CHECK (nnsb_pinned_alloc (ctx, FRAME, &host));
    unsigned int seed = 12345u + (unsigned int) s;
    for (size_t i = 0; i < FRAME; i++) {
      seed = seed * 1664525u + 1013904223u;
      ((unsigned char *) host)[i] = (unsigned char) (seed >> 24);
    }
    memset (&p.frame, 0, sizeof (p.frame));
    p.frame.size = FRAME;
    if (resident) {
      void *d = nullptr;
      CHECK (nnsb_dev_alloc (ctx, FRAME, &d));
      CHECK (nnsb_h2d_async (stream, d, host, FRAME));
      CHECK (nnsb_stream_sync (stream));
      p.frame.data = d;
      p.frame.on_device = 1;
    } else {
      p.frame.data = host;
      p.frame.on_device = 0;
    }
  }
  for (auto & p : pipes)
    run (&p, 3);                /* warm-up: contexts, plans, pools */
  std::vector<std::thread> th;
  const auto t0 = std::chrono::steady_clock::now ();
  for (auto & p : pipes)
    th.emplace_back (run, &p, frames);
  for (auto & t : th)
    t.join ();
  const double dt = std::chrono::duration<double> (std::chrono::steady_clock::now () - t0).count ();
  const double fps = (double) streams * frames / dt;
  /* algorithmic bytes per element along the chain: typecast 5, arithmetic 8, transpose 8 */
  printf ("{\"frames/s\": %.1f, \"GB/s\": %.1f, \"streams\": %d, \"frames_per_stream\": %d, "
      "\"mode\": \"%s\", \"driver\": \"native threads through the C-ABI (tools/stream_bench.cpp)\"}\n",
      fps, fps * FRAME * 21 / 1e9, streams, frames, argv[4]);
  for (auto & p : pipes) {
    nnsb_converter_free (p.conv);
    for (int k = 0; k < 3; k++)
      nnsb_transform_free (p.t[k]);
  }
  return 0;
}
