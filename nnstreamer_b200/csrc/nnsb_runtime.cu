/*
 * nnsb_runtime.cu -- device runtime and plans behind the C-ABI (include/nnstreamer_b200.h).
 *
 *  ctx     : binds one element instance to one GPU; owns a size-bucketed device pool and a
 *            pinned-host pool so steady-state streaming never calls cudaMalloc/cudaHostAlloc.
 *            This is what replaces `gst_allocator_alloc (NULL, size, NULL)` per tensor per
 *            buffer (gsttensor_transform.c:1758) and tensor_allocator.c:101-128.
 *  stream  : one CUDA stream per element instance (transform() runs on the upstream
 *            streaming thread; every entry point re-binds the device first).
 *  plan    : parsed option + input info -> kernel selection and launch geometry; replaces
 *            the mode switch of gst_tensor_transform_transform (:1773-1802).
 *
 * No CPU fallback exists: without a usable CUDA device every call fails with NNSB_ECUDA.
 */
#include <cstdio>
#include <cuda_runtime.h>
#include <sched.h>
#include <time.h>

#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <vector>

#include "nnsb_internal.h"
#include "nnsb_kernels.h"

using namespace nnsb;

/* ------------------------------------------------------------------------------------- */
/* objects                                                                                 */
/* ------------------------------------------------------------------------------------- */
/* A recorded CUDA event shared by everything that depends on the same point of a stream: the
 * "ready" mark of the blocks a launch produced and the "last use" mark of the blocks it read.
 * Events are pooled per device for the life of the process (creating one costs microseconds). */
struct nnsb_evt
{
  cudaEvent_t e;
  cudaStream_t stream;          /* the stream it was recorded on */
  int device;
  int refs;                     /* guarded by g_evt_mu */
};

/* a cached block and the events that must have completed before another stream may touch it */
struct Cached
{
  void *ptr;
  std::vector<nnsb_evt *> pending;
};

struct LiveInfo
{
  size_t bucket;
  bool is_block;                /* wrapped by an nnsb_block (survives nnsb_ctx_destroy) */
};

struct nnsb_ctx
{
  int device;
  int sm_count;
  int blocks_per_sm;            /* 0 = per-kernel default; NNSB_BLOCKS_PER_SM overrides */
  std::mutex mu;
  int refs;                     /* creator + live blocks; guarded by mu */
  bool closing;                 /* nnsb_ctx_destroy was called: nothing is cached any more */
  uint64_t n_dev_malloc;
  std::map<size_t, std::vector<Cached> > dev_free;    /* bucket -> cached blocks (oldest first) */
  std::map<size_t, int> dev_count;                    /* bucket -> blocks obtained from cudaMalloc */
  std::map<void *, LiveInfo> dev_live;                /* block -> bucket */
  std::map<size_t, std::vector<void *> > pin_free;
  std::map<void *, LiveInfo> pin_live;
};

/* One reference-counted allocation behind any number of nnsb_memory views. */
struct nnsb_block
{
  nnsb_ctx *ctx;
  void *base;
  int on_device;
  int refs;                     /* guarded by ctx->mu */
  nnsb_evt *ready;              /* contents complete after this event (NULL = now) */
  std::vector<nnsb_evt *> uses; /* at most one per stream: work that still reads the block */
};

struct nnsb_stream
{
  nnsb_ctx *ctx;
  cudaStream_t s;
  bool owned;
};

struct nnsb_event
{
  nnsb_ctx *ctx;
  cudaEvent_t e;
};

struct nnsb_plan
{
  nnsb_ctx *ctx;
  nnsb_option opt;
  nnsb_tensor_info in_info, out_info;
  size_t in_size, out_size;
  uint64_t frame_elems;
  int acceleration;             /* effective: orc_supported() */
  /* arithmetic / typecast */
  ChainArgs chain;
  nnsb_cell *d_table;
  /* transpose / dimchg */
  PermDesc perm;
  PermVecPlan perm_vec;         /* 16-byte vector path, when the box qualifies */
  void *d_perm_table;
  /* stand */
  uint32_t stand_nch;
  double *d_scratch;
  size_t scratch_doubles;
  double *d_stats;
  size_t stats_doubles;
  int no_fused;                 /* NNSB_STAND_NO_FUSED: always the three-launch scheme */
  int stand_fused_used;         /* the last launch took the single-read kernel */
  int last_launches;
};

#define CUDA_TRY(call) \
  do { \
    cudaError_t e__ = (call); \
    if (e__ != cudaSuccess) \
      return nnsb_set_error (NNSB_ECUDA, "%s failed: %s", #call, cudaGetErrorString (e__)); \
  } while (0)

static int
bind (const nnsb_ctx * ctx)
{
  if (!ctx)
    return nnsb_set_error (NNSB_EINVAL, "null context");
  CUDA_TRY (cudaSetDevice (ctx->device));
  return NNSB_OK;
}

static size_t
bucket_of (size_t size)
{
  if (size < 256)
    size = 256;
  if (size <= (2u << 20)) {
    size_t b = 256;
    while (b < size)
      b <<= 1;
    return b;
  }
  const size_t g = 2u << 20;    /* 2 MiB granularity above 2 MiB */
  return (size + g - 1) / g * g;
}

/* ------------------------------------------------------------------------------------- */
/* pooled events                                                                           */
/* ------------------------------------------------------------------------------------- */
static std::mutex g_evt_mu;
static std::map<int, std::vector<nnsb_evt *> > g_evt_free;      /* device -> idle events */

/* a new event recorded on `stream` (device `device` must be current); one reference */
static nnsb_evt *
evt_record (int device, cudaStream_t stream)
{
  nnsb_evt *ev = nullptr;
  {
    std::lock_guard<std::mutex> lk (g_evt_mu);
    auto & v = g_evt_free[device];
    if (!v.empty ()) {
      ev = v.back ();
      v.pop_back ();
    }
  }
  if (!ev) {
    cudaEvent_t e;
    if (cudaEventCreateWithFlags (&e, cudaEventDisableTiming) != cudaSuccess) {
      cudaGetLastError ();
      return nullptr;
    }
    ev = new (std::nothrow) nnsb_evt ();
    if (!ev) {
      cudaEventDestroy (e);
      return nullptr;
    }
    ev->e = e;
    ev->device = device;
  }
  ev->stream = stream;
  ev->refs = 1;
  if (cudaEventRecord (ev->e, stream) != cudaSuccess) {
    cudaGetLastError ();
    std::lock_guard<std::mutex> lk (g_evt_mu);
    g_evt_free[device].push_back (ev);
    return nullptr;
  }
  return ev;
}

static nnsb_evt *
evt_ref (nnsb_evt * ev)
{
  if (ev) {
    std::lock_guard<std::mutex> lk (g_evt_mu);
    ev->refs++;
  }
  return ev;
}

static void
evt_unref (nnsb_evt * ev)
{
  if (!ev)
    return;
  std::lock_guard<std::mutex> lk (g_evt_mu);
  if (--ev->refs == 0)
    g_evt_free[ev->device].push_back (ev);
}

/* true when work on `stream` (or anybody, if have_stream is false) may touch what the event guards */
static bool
evt_satisfied (nnsb_evt * ev, bool have_stream, cudaStream_t stream)
{
  if (have_stream && ev->stream == stream)
    return true;                /* same stream: ordered by construction */
  cudaError_t e = cudaEventQuery (ev->e);
  if (e == cudaSuccess)
    return true;
  if (e != cudaErrorNotReady)
    cudaGetLastError ();
  return false;
}

/* ------------------------------------------------------------------------------------- */
/* context / stream / event                                                                */
/* ------------------------------------------------------------------------------------- */
/* Elements own their streams, and a pipeline of a few branches easily has more of them than the
 * 8 hardware connections CUDA opens by default; streams that share a connection wait for each
 * other's event waits.  Ask for 32 unless the user chose a number (read when the CUDA context
 * is created, so this must run before the process's first CUDA call to have any effect). */
static void
prefer_more_connections ()
{
  static std::once_flag once;
  std::call_once (once, [] { setenv ("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0); });
}

extern "C" int
nnsb_device_count (int *count)
{
  prefer_more_connections ();
  if (!count)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_device_count: null argument");
  *count = 0;
  CUDA_TRY (cudaGetDeviceCount (count));
  return NNSB_OK;
}

extern "C" int
nnsb_ctx_create (int device_id, nnsb_ctx ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_ctx_create: null argument");
  *out = nullptr;
  prefer_more_connections ();
  int count = 0;
  CUDA_TRY (cudaGetDeviceCount (&count));
  if (device_id < 0 || device_id >= count)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_ctx_create: device %d of %d does not exist",
        device_id, count);
  CUDA_TRY (cudaSetDevice (device_id));
  cudaDeviceProp prop;
  CUDA_TRY (cudaGetDeviceProperties (&prop, device_id));
  if (prop.major != 10)
    return nnsb_set_error (NNSB_ENOTSUP,
        "nnsb_ctx_create: device %d is sm_%d%d; this library carries sm_100a code only",
        device_id, prop.major, prop.minor);
  nnsb_ctx *ctx = new (std::nothrow) nnsb_ctx ();
  if (!ctx)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_ctx_create: out of memory");
  ctx->device = device_id;
  ctx->sm_count = prop.multiProcessorCount;
  ctx->blocks_per_sm = 0;
  ctx->refs = 1;
  ctx->closing = false;
  ctx->n_dev_malloc = 0;
  const char *env = getenv ("NNSB_BLOCKS_PER_SM");
  if (env && atoi (env) > 0)
    ctx->blocks_per_sm = atoi (env);
  *out = ctx;
  return NNSB_OK;
}

static void
drop_pending (Cached & c)
{
  for (nnsb_evt * ev : c.pending)
    evt_unref (ev);
  c.pending.clear ();
}

extern "C" int
nnsb_dev_trim (nnsb_ctx * ctx)
{
  int r = bind (ctx);
  if (r)
    return r;
  std::map<size_t, std::vector<Cached> > dev;
  std::map<size_t, std::vector<void *> > pin;
  {
    std::lock_guard<std::mutex> lk (ctx->mu);
    dev.swap (ctx->dev_free);
    pin.swap (ctx->pin_free);
    for (auto & kv : dev)
      ctx->dev_count[kv.first] -= (int) kv.second.size ();
  }
  /* cudaFree waits for the work still touching the block */
  for (auto & kv : dev)
    for (Cached & c : kv.second) {
      drop_pending (c);
      cudaFree (c.ptr);
    }
  for (auto & kv : pin)
    for (void *p : kv.second)
      cudaFreeHost (p);
  return NNSB_OK;
}

/* drops one reference; the context goes away with the last one (mu must NOT be held) */
static void
ctx_unref (nnsb_ctx * ctx)
{
  bool last;
  {
    std::lock_guard<std::mutex> lk (ctx->mu);
    last = --ctx->refs == 0;
  }
  if (last)
    delete ctx;
}

extern "C" int
nnsb_ctx_destroy (nnsb_ctx * ctx)
{
  if (!ctx)
    return NNSB_OK;
  if (bind (ctx) == NNSB_OK) {
    cudaDeviceSynchronize ();
    std::vector<void *> dev, pin;
    {
      std::lock_guard<std::mutex> lk (ctx->mu);
      ctx->closing = true;
      /* raw allocations belong to whoever created the context; blocks still referenced by a
       * memory somewhere downstream stay alive until their last nnsb_memory_release */
      for (auto it = ctx->dev_live.begin (); it != ctx->dev_live.end ();) {
        if (!it->second.is_block) {
          dev.push_back (it->first);
          it = ctx->dev_live.erase (it);
        } else
          ++it;
      }
      for (auto it = ctx->pin_live.begin (); it != ctx->pin_live.end ();) {
        if (!it->second.is_block) {
          pin.push_back (it->first);
          it = ctx->pin_live.erase (it);
        } else
          ++it;
      }
    }
    nnsb_dev_trim (ctx);
    for (void *p : dev)
      cudaFree (p);
    for (void *p : pin)
      cudaFreeHost (p);
  }
  ctx_unref (ctx);
  return NNSB_OK;
}

extern "C" uint64_t
nnsb_ctx_device_mallocs (const nnsb_ctx * ctx)
{
  return ctx ? ctx->n_dev_malloc : 0;
}

extern "C" int
nnsb_ctx_device (const nnsb_ctx * ctx)
{
  return ctx ? ctx->device : NNSB_EINVAL;
}

extern "C" int
nnsb_ctx_sm_count (const nnsb_ctx * ctx)
{
  return ctx ? ctx->sm_count : NNSB_EINVAL;
}

extern "C" int
nnsb_stream_create (nnsb_ctx * ctx, nnsb_stream ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_stream_create: null argument");
  *out = nullptr;
  int r = bind (ctx);
  if (r)
    return r;
  cudaStream_t s;
  CUDA_TRY (cudaStreamCreateWithFlags (&s, cudaStreamNonBlocking));
  nnsb_stream *st = new (std::nothrow) nnsb_stream ();
  if (!st) {
    cudaStreamDestroy (s);
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_stream_create: out of memory");
  }
  st->ctx = ctx;
  st->s = s;
  st->owned = true;
  *out = st;
  return NNSB_OK;
}

extern "C" int
nnsb_stream_wrap (nnsb_ctx * ctx, void *cuda_stream, nnsb_stream ** out)
{
  if (!out || !ctx)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_stream_wrap: null argument");
  nnsb_stream *st = new (std::nothrow) nnsb_stream ();
  if (!st)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_stream_wrap: out of memory");
  st->ctx = ctx;
  st->s = (cudaStream_t) cuda_stream;
  st->owned = false;
  *out = st;
  return NNSB_OK;
}

extern "C" int
nnsb_stream_sync (nnsb_stream * s)
{
  if (!s)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_stream_sync: null stream");
  int r = bind (s->ctx);
  if (r)
    return r;
  CUDA_TRY (cudaStreamSynchronize (s->s));
  return NNSB_OK;
}

extern "C" int
nnsb_stream_destroy (nnsb_stream * s)
{
  if (!s)
    return NNSB_OK;
  if (s->owned && bind (s->ctx) == NNSB_OK) {
    cudaStreamSynchronize (s->s);
    cudaStreamDestroy (s->s);
  }
  delete s;
  return NNSB_OK;
}

extern "C" int
nnsb_event_create (nnsb_ctx * ctx, nnsb_event ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_event_create: null argument");
  *out = nullptr;
  int r = bind (ctx);
  if (r)
    return r;
  cudaEvent_t e;
  CUDA_TRY (cudaEventCreate (&e));
  nnsb_event *ev = new (std::nothrow) nnsb_event ();
  if (!ev) {
    cudaEventDestroy (e);
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_event_create: out of memory");
  }
  ev->ctx = ctx;
  ev->e = e;
  *out = ev;
  return NNSB_OK;
}

extern "C" int
nnsb_event_record (nnsb_event * e, nnsb_stream * s)
{
  if (!e || !s)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_event_record: null argument");
  int r = bind (e->ctx);
  if (r)
    return r;
  CUDA_TRY (cudaEventRecord (e->e, s->s));
  return NNSB_OK;
}

extern "C" int
nnsb_event_sync (nnsb_event * e)
{
  if (!e)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_event_sync: null argument");
  int r = bind (e->ctx);
  if (r)
    return r;
  CUDA_TRY (cudaEventSynchronize (e->e));
  return NNSB_OK;
}

extern "C" int
nnsb_stream_wait_event (nnsb_stream * s, nnsb_event * e)
{
  if (!e || !s)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_stream_wait_event: null argument");
  int r = bind (s->ctx);
  if (r)
    return r;
  CUDA_TRY (cudaStreamWaitEvent (s->s, e->e, 0));
  return NNSB_OK;
}

extern "C" int
nnsb_event_elapsed_ms (nnsb_event * start, nnsb_event * stop, float *ms)
{
  if (!start || !stop || !ms)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_event_elapsed_ms: null argument");
  int r = bind (start->ctx);
  if (r)
    return r;
  CUDA_TRY (cudaEventElapsedTime (ms, start->e, stop->e));
  return NNSB_OK;
}

extern "C" int
nnsb_event_destroy (nnsb_event * e)
{
  if (!e)
    return NNSB_OK;
  if (bind (e->ctx) == NNSB_OK)
    cudaEventDestroy (e->e);
  delete e;
  return NNSB_OK;
}

/* ------------------------------------------------------------------------------------- */
/* memory                                                                                  */
/* ------------------------------------------------------------------------------------- */
/* How many blocks of one size a context may obtain from cudaMalloc before an allocation has to
 * wait for a block whose last GPU work is still in flight.  That wait is taken on the HOST
 * (cudaEventSynchronize on the oldest block's events): it is the back-pressure of the pipeline.
 * Without it a producer thread runs arbitrarily far ahead of the GPU, and deep queues on more
 * streams than the device has hardware connections serialise behind each other (measured:
 * 8 pipelines x 4 elements queued 300 frames deep ran at 13 k frames/s).  The depth also decides
 * how far apart in time a tensor's producer and consumer run, i.e. whether the consumer still
 * finds it in the 126 MB L2: 8 x 1080p converter -> typecast -> arithmetic -> transpose chains
 * gave 48.3 / 43.1 / 34.0 / 36.4 k frames/s at depth 1 / 2 / 3 / 4 (profiles/r2_pool_depth.md),
 * so the default is 1: an element starts its next buffer when its consumer is done with the
 * previous one.  NNSB_POOL_DEPTH overrides. */
/* Waits for an event without monopolising the core: cudaEventSynchronize spins, and a node
 * that runs 8 ranks x 8 streaming threads on 32 cores (BASELINE configs[4]) then loses a fifth of
 * its throughput to threads spinning against each other.  Query + yield costs nothing when the
 * cores are free; after a few hundred misses the thread sleeps in 20 us slices (the work it waits
 * for is then clearly long).  NNSB_WAIT = spin | yield | sleep (default) selects the policy. */
static void
host_wait (cudaEvent_t e)
{
  static const int policy = [] {
    const char *v = getenv ("NNSB_WAIT");
    return !v ? 2 : strcmp (v, "spin") == 0 ? 0 : strcmp (v, "yield") == 0 ? 1 : 2;
  } ();
  if (policy == 0) {
    cudaEventSynchronize (e);
    return;
  }
  for (int misses = 0;; misses++) {
    const cudaError_t r = cudaEventQuery (e);
    if (r != cudaErrorNotReady) {
      if (r != cudaSuccess)
        cudaGetLastError ();
      return;
    }
    if (policy == 2 && misses >= 256) {
      const struct timespec ts = { 0, 20000 };
      nanosleep (&ts, nullptr);
    } else {
      sched_yield ();
    }
  }
}

static int
pool_depth ()
{
  static int depth = 0;
  if (depth == 0) {
    const char *env = getenv ("NNSB_POOL_DEPTH");
    depth = (env && atoi (env) > 0) ? atoi (env) : 1;
  }
  return depth;
}

/* Takes a device block of `bucket` bytes.  have_stream: the block will first be touched by
 * work on `stream`. */
static int
dev_take (nnsb_ctx * ctx, size_t b, bool have_stream, cudaStream_t stream, void **out, bool as_block)
{
  std::vector<nnsb_evt *> must_wait;
  void *p = nullptr;
  {
    std::lock_guard<std::mutex> lk (ctx->mu);
    auto it = ctx->dev_free.find (b);
    if (it != ctx->dev_free.end () && !it->second.empty ()) {
      std::vector<Cached> & v = it->second;
      /* newest first: a block nothing in flight touches (or only this stream does) */
      for (size_t i = v.size (); i-- > 0 && !p;) {
        bool ok = true;
        for (nnsb_evt * ev : v[i].pending)
          if (!evt_satisfied (ev, have_stream, stream)) {
            ok = false;
            break;
          }
        if (ok) {
          p = v[i].ptr;
          drop_pending (v[i]);
          v.erase (v.begin () + (long) i);
        }
      }
      if (!p && ctx->dev_count[b] >= pool_depth ()) {
        /* every cached block is still in use by queued work: take the oldest and wait for it */
        p = v.front ().ptr;
        must_wait.swap (v.front ().pending);
        v.erase (v.begin ());
      }
    }
    if (p)
      ctx->dev_live[p] = LiveInfo { b, as_block };
  }
  if (p) {
    for (nnsb_evt * ev : must_wait) {
      if (!(have_stream && ev->stream == stream))
        host_wait (ev->e);                      /* back-pressure: see pool_depth () */
      evt_unref (ev);
    }
    *out = p;
    return NNSB_OK;
  }
  cudaError_t e = cudaMalloc (&p, b);
  if (e != cudaSuccess) {
    /* give cached blocks back and retry once */
    cudaGetLastError ();
    nnsb_dev_trim (ctx);
    e = cudaMalloc (&p, b);
    if (e != cudaSuccess) {
      cudaGetLastError ();
      return nnsb_set_error (NNSB_ENOMEM, "cudaMalloc (%zu) failed: %s", b, cudaGetErrorString (e));
    }
  }
  std::lock_guard<std::mutex> lk (ctx->mu);
  ctx->dev_live[p] = LiveInfo { b, as_block };
  ctx->dev_count[b] += 1;
  ctx->n_dev_malloc += 1;
  *out = p;
  return NNSB_OK;
}

/* Returns a device block to the pool (or to the driver when the context is closing).
 * `pending` is consumed. */
static int
dev_give (nnsb_ctx * ctx, void *ptr, std::vector<nnsb_evt *> & pending)
{
  bool free_now = false;
  {
    std::lock_guard<std::mutex> lk (ctx->mu);
    auto it = ctx->dev_live.find (ptr);
    if (it == ctx->dev_live.end ()) {
      for (nnsb_evt * ev : pending)
        evt_unref (ev);
      pending.clear ();
      return nnsb_set_error (NNSB_EINVAL, "%p was not allocated by this context", ptr);
    }
    const size_t b = it->second.bucket;
    ctx->dev_live.erase (it);
    if (ctx->closing) {
      free_now = true;
      ctx->dev_count[b] -= 1;
    } else {
      Cached c;
      c.ptr = ptr;
      c.pending.swap (pending);
      ctx->dev_free[b].push_back (c);
    }
  }
  if (free_now) {
    for (nnsb_evt * ev : pending)
      evt_unref (ev);
    pending.clear ();
    if (cudaSetDevice (ctx->device) == cudaSuccess)
      cudaFree (ptr);           /* waits for the work still touching it */
  }
  return NNSB_OK;
}

extern "C" int
nnsb_dev_alloc (nnsb_ctx * ctx, size_t size, void **out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_dev_alloc: null argument");
  *out = nullptr;
  int r = bind (ctx);
  if (r)
    return r;
  return dev_take (ctx, bucket_of (size), false, nullptr, out, false);
}

extern "C" int
nnsb_dev_free (nnsb_ctx * ctx, void *ptr)
{
  if (!ptr)
    return NNSB_OK;
  if (!ctx)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_dev_free: null context");
  std::vector<nnsb_evt *> none;
  return dev_give (ctx, ptr, none);
}

extern "C" int
nnsb_dev_alloc_async (nnsb_stream * s, size_t size, void **out)
{
  if (!out || !s)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_dev_alloc_async: null argument");
  *out = nullptr;
  int r = bind (s->ctx);
  if (r)
    return r;
  return dev_take (s->ctx, bucket_of (size), true, s->s, out, false);
}

extern "C" int
nnsb_dev_free_async (nnsb_stream * s, void *ptr)
{
  if (!ptr)
    return NNSB_OK;
  if (!s)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_dev_free_async: null stream");
  int r = bind (s->ctx);
  if (r)
    return r;
  std::vector<nnsb_evt *> pending;
  nnsb_evt *ev = evt_record (s->ctx->device, s->s);
  if (!ev) {
    /* no event to be had: fall back to waiting here */
    cudaStreamSynchronize (s->s);
  } else {
    pending.push_back (ev);
  }
  return dev_give (s->ctx, ptr, pending);
}

static int
pinned_take (nnsb_ctx * ctx, size_t b, void **out, bool as_block)
{
  {
    std::lock_guard<std::mutex> lk (ctx->mu);
    auto it = ctx->pin_free.find (b);
    if (it != ctx->pin_free.end () && !it->second.empty ()) {
      void *p = it->second.back ();
      it->second.pop_back ();
      ctx->pin_live[p] = LiveInfo { b, as_block };
      *out = p;
      return NNSB_OK;
    }
  }
  void *p = nullptr;
  cudaError_t e = cudaHostAlloc (&p, b, cudaHostAllocDefault);
  if (e != cudaSuccess) {
    cudaGetLastError ();
    return nnsb_set_error (NNSB_ENOMEM, "cudaHostAlloc (%zu) failed: %s", b, cudaGetErrorString (e));
  }
  std::lock_guard<std::mutex> lk (ctx->mu);
  ctx->pin_live[p] = LiveInfo { b, as_block };
  *out = p;
  return NNSB_OK;
}

static int
pinned_give (nnsb_ctx * ctx, void *ptr)
{
  bool free_now = false;
  {
    std::lock_guard<std::mutex> lk (ctx->mu);
    auto it = ctx->pin_live.find (ptr);
    if (it == ctx->pin_live.end ())
      return nnsb_set_error (NNSB_EINVAL, "nnsb_pinned_free: %p was not allocated by this context", ptr);
    if (ctx->closing)
      free_now = true;
    else
      ctx->pin_free[it->second.bucket].push_back (ptr);
    ctx->pin_live.erase (it);
  }
  if (free_now && cudaSetDevice (ctx->device) == cudaSuccess)
    cudaFreeHost (ptr);
  return NNSB_OK;
}

extern "C" int
nnsb_pinned_alloc (nnsb_ctx * ctx, size_t size, void **out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_pinned_alloc: null argument");
  *out = nullptr;
  int r = bind (ctx);
  if (r)
    return r;
  return pinned_take (ctx, bucket_of (size), out, false);
}

extern "C" int
nnsb_pinned_free (nnsb_ctx * ctx, void *ptr)
{
  if (!ptr)
    return NNSB_OK;
  if (!ctx)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_pinned_free: null context");
  return pinned_give (ctx, ptr);
}

/* ------------------------------------------------------------------------------------- */
/* nnsb_memory: reference-counted, stream-ordered views                                    */
/* ------------------------------------------------------------------------------------- */
extern "C" int
nnsb_memory_alloc (nnsb_ctx * ctx, size_t size, int on_device, nnsb_stream * s, nnsb_memory * out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_memory_alloc: null argument");
  memset (out, 0, sizeof (*out));
  int r = bind (ctx);
  if (r)
    return r;
  nnsb_block *blk = new (std::nothrow) nnsb_block ();
  if (!blk)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_memory_alloc: out of memory");
  void *p = nullptr;
  const size_t b = bucket_of (size);
  r = on_device ? dev_take (ctx, b, s != nullptr, s ? s->s : nullptr, &p, true)
      : pinned_take (ctx, b, &p, true);
  if (r) {
    delete blk;
    return r;
  }
  blk->ctx = ctx;
  blk->base = p;
  blk->on_device = on_device ? 1 : 0;
  blk->refs = 1;
  blk->ready = nullptr;
  {
    std::lock_guard<std::mutex> lk (ctx->mu);
    ctx->refs++;
  }
  out->data = p;
  out->size = size;
  out->on_device = blk->on_device;
  out->owned = 1;
  out->owner = ctx;
  out->block = blk;
  return NNSB_OK;
}

/* wraps a raw allocation of `ctx` (nnsb_dev_alloc / nnsb_pinned_alloc) into an owned memory */
extern "C" int
nnsb_memory_adopt (nnsb_ctx * ctx, void *ptr, size_t size, int on_device, nnsb_memory * out)
{
  if (!ctx || !ptr || !out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_memory_adopt: null argument");
  nnsb_block *blk = new (std::nothrow) nnsb_block ();
  if (!blk)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_memory_adopt: out of memory");
  {
    std::lock_guard<std::mutex> lk (ctx->mu);
    auto & live = on_device ? ctx->dev_live : ctx->pin_live;
    auto it = live.find (ptr);
    if (it == live.end ()) {
      delete blk;
      return nnsb_set_error (NNSB_EINVAL, "nnsb_memory_adopt: %p was not allocated by this context", ptr);
    }
    it->second.is_block = true;
    ctx->refs++;
  }
  blk->ctx = ctx;
  blk->base = ptr;
  blk->on_device = on_device ? 1 : 0;
  blk->refs = 1;
  blk->ready = nullptr;
  memset (out, 0, sizeof (*out));
  out->data = ptr;
  out->size = size;
  out->on_device = blk->on_device;
  out->owned = 1;
  out->owner = ctx;
  out->block = blk;
  return NNSB_OK;
}

extern "C" int
nnsb_memory_ref (const nnsb_memory * mem, nnsb_memory * out)
{
  if (!mem || !out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_memory_ref: null argument");
  nnsb_memory m = *mem;
  nnsb_block *blk = (nnsb_block *) mem->block;
  if (blk) {
    std::lock_guard<std::mutex> lk (blk->ctx->mu);
    blk->refs++;
    m.owned = 1;
  } else {
    m.owned = 0;                /* caller-managed memory: a plain view */
  }
  *out = m;
  return NNSB_OK;
}

/* 1 when `mem` holds the only reference on its block (nobody else can be reading it) */
extern "C" int
nnsb_memory_is_unique (const nnsb_memory * mem)
{
  nnsb_block *blk = mem ? (nnsb_block *) mem->block : nullptr;
  if (!blk)
    return 0;
  std::lock_guard<std::mutex> lk (blk->ctx->mu);
  return blk->refs == 1;
}

extern "C" int
nnsb_memory_share (const nnsb_memory * mem, size_t offset, size_t size, nnsb_memory * out)
{
  if (!mem || !out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_memory_share: null argument");
  if (offset > mem->size || size > mem->size - offset)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_memory_share: [%zu, +%zu) is outside a memory of %zu bytes",
        offset, size, mem->size);
  int r = nnsb_memory_ref (mem, out);
  if (r)
    return r;
  out->data = (char *) mem->data + offset;
  out->size = size;
  return NNSB_OK;
}

extern "C" int
nnsb_memory_release (nnsb_memory * mem)
{
  if (!mem || !mem->data || !mem->owned)
    return NNSB_OK;
  nnsb_block *blk = (nnsb_block *) mem->block;
  mem->data = nullptr;
  mem->owned = 0;
  mem->block = nullptr;
  if (!blk)
    return NNSB_OK;
  nnsb_ctx *ctx = blk->ctx;
  bool last;
  {
    std::lock_guard<std::mutex> lk (ctx->mu);
    last = --blk->refs == 0;
  }
  if (!last)
    return NNSB_OK;
  int r;
  if (blk->on_device) {
    std::vector<nnsb_evt *> pending;
    pending.swap (blk->uses);
    if (blk->ready)
      pending.push_back (blk->ready);
    r = dev_give (ctx, blk->base, pending);
  } else {
    for (nnsb_evt * ev : blk->uses)
      evt_unref (ev);
    evt_unref (blk->ready);
    r = pinned_give (ctx, blk->base);
  }
  delete blk;
  ctx_unref (ctx);
  return r;
}

extern "C" int
nnsb_memory_wait (const nnsb_memory * mem, nnsb_stream * s)
{
  if (!mem || !s)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_memory_wait: null argument");
  nnsb_block *blk = (nnsb_block *) mem->block;
  if (!blk)
    return NNSB_OK;
  nnsb_evt *ev;
  {
    std::lock_guard<std::mutex> lk (blk->ctx->mu);
    ev = evt_ref (blk->ready);
  }
  if (!ev)
    return NNSB_OK;
  int r = NNSB_OK;
  if (ev->stream != s->s) {
    r = bind (s->ctx);
    if (!r) {
      cudaError_t e = cudaStreamWaitEvent (s->s, ev->e, 0);
      if (e != cudaSuccess)
        r = nnsb_set_error (NNSB_ECUDA, "cudaStreamWaitEvent failed: %s", cudaGetErrorString (e));
    }
  }
  evt_unref (ev);
  return r;
}

extern "C" int
nnsb_memory_sync (const nnsb_memory * mem)
{
  if (!mem)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_memory_sync: null argument");
  nnsb_block *blk = (nnsb_block *) mem->block;
  if (!blk)
    return NNSB_OK;
  nnsb_evt *ev;
  {
    std::lock_guard<std::mutex> lk (blk->ctx->mu);
    ev = evt_ref (blk->ready);
  }
  if (!ev)
    return NNSB_OK;
  cudaError_t e = cudaEventSynchronize (ev->e);
  evt_unref (ev);
  if (e != cudaSuccess)
    return nnsb_set_error (NNSB_ECUDA, "cudaEventSynchronize failed: %s", cudaGetErrorString (e));
  return NNSB_OK;
}

extern "C" int
nnsb_memory_mark_io (nnsb_stream * s, nnsb_memory * outs, uint32_t n_out, const nnsb_memory * ins,
    uint32_t n_in)
{
  if (!s || (n_out && !outs) || (n_in && !ins))
    return nnsb_set_error (NNSB_EINVAL, "nnsb_memory_mark_io: null argument");
  bool any = false;
  for (uint32_t i = 0; i < n_out && !any; i++)
    any = outs[i].block != nullptr;
  for (uint32_t i = 0; i < n_in && !any; i++)
    any = ins[i].block != nullptr;
  if (!any)
    return NNSB_OK;
  int r = bind (s->ctx);
  if (r)
    return r;
  nnsb_evt *ev = evt_record (s->ctx->device, s->s);
  if (!ev) {
    /* no event to be had: make everything true the slow way */
    CUDA_TRY (cudaStreamSynchronize (s->s));
    for (uint32_t i = 0; i < n_out; i++) {
      nnsb_block *blk = (nnsb_block *) outs[i].block;
      if (!blk)
        continue;
      std::lock_guard<std::mutex> lk (blk->ctx->mu);
      evt_unref (blk->ready);
      blk->ready = nullptr;
    }
    return NNSB_OK;
  }
  for (uint32_t i = 0; i < n_out; i++) {
    nnsb_block *blk = (nnsb_block *) outs[i].block;
    if (!blk)
      continue;
    std::lock_guard<std::mutex> lk (blk->ctx->mu);
    if (blk->ready == ev)
      continue;                 /* two views of one block */
    evt_unref (blk->ready);
    blk->ready = evt_ref (ev);
  }
  for (uint32_t i = 0; i < n_in; i++) {
    nnsb_block *blk = (nnsb_block *) ins[i].block;
    if (!blk)
      continue;
    std::lock_guard<std::mutex> lk (blk->ctx->mu);
    bool placed = false;
    for (auto & u : blk->uses) {
      if (u == ev) {
        placed = true;
        break;
      }
      if (u->stream == ev->stream) {
        /* a later point of the same stream covers the earlier one */
        evt_unref (u);
        u = evt_ref (ev);
        placed = true;
        break;
      }
    }
    if (!placed)
      blk->uses.push_back (evt_ref (ev));
  }
  evt_unref (ev);
  return NNSB_OK;
}

extern "C" int
nnsb_memory_mark_ready (nnsb_memory * mem, nnsb_stream * s)
{
  return nnsb_memory_mark_io (s, mem, mem ? 1 : 0, nullptr, 0);
}

extern "C" int
nnsb_memory_mark_use (const nnsb_memory * mem, nnsb_stream * s)
{
  return nnsb_memory_mark_io (s, nullptr, 0, mem, mem ? 1 : 0);
}

extern "C" int
nnsb_memory_export (const nnsb_memory * mem, nnsb_device_view * view)
{
  if (!mem || !view)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_memory_export: null argument");
  if (!mem->on_device || !mem->data)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_memory_export: not a device memory");
  memset (view, 0, sizeof (*view));
  view->device_ptr = mem->data;
  view->size = mem->size;
  view->device_id = -1;
  nnsb_block *blk = (nnsb_block *) mem->block;
  if (blk) {
    std::lock_guard<std::mutex> lk (blk->ctx->mu);
    view->device_id = blk->ctx->device;
    /* valid for as long as the caller holds its reference on the memory */
    view->ready_event = blk->ready ? (void *) blk->ready->e : nullptr;
  } else {
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes (&attr, mem->data) == cudaSuccess)
      view->device_id = attr.device;
    else
      cudaGetLastError ();
  }
  return NNSB_OK;
}

/* 1 when `p` is page-locked (or managed) memory the copy engine can read directly, 0 when it is
 * ordinary pageable memory: cudaMemcpyAsync from such a source is staged by the driver and
 * blocks the calling thread */
extern "C" int
nnsb_host_is_pinned (const void *p)
{
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes (&attr, p) != cudaSuccess) {
    cudaGetLastError ();
    return 0;
  }
  return attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged;
}

extern "C" void *
nnsb_stream_handle (nnsb_stream * s)
{
  return s ? (void *) s->s : nullptr;
}

#define STREAM_PROLOGUE(name) \
  if (!s) \
    return nnsb_set_error (NNSB_EINVAL, name ": null stream"); \
  { int r__ = bind (s->ctx); if (r__) return r__; }

extern "C" int
nnsb_dev_memset_async (nnsb_stream * s, void *dst, int value, size_t n)
{
  STREAM_PROLOGUE ("nnsb_dev_memset_async");
  CUDA_TRY (cudaMemsetAsync (dst, value, n, s->s));
  return NNSB_OK;
}

extern "C" int
nnsb_h2d_async (nnsb_stream * s, void *dst_dev, const void *src_host, size_t n)
{
  STREAM_PROLOGUE ("nnsb_h2d_async");
  if (n == 0)
    return NNSB_OK;
  CUDA_TRY (cudaMemcpyAsync (dst_dev, src_host, n, cudaMemcpyHostToDevice, s->s));
  return NNSB_OK;
}

extern "C" int
nnsb_d2h_async (nnsb_stream * s, void *dst_host, const void *src_dev, size_t n)
{
  STREAM_PROLOGUE ("nnsb_d2h_async");
  if (n == 0)
    return NNSB_OK;
  CUDA_TRY (cudaMemcpyAsync (dst_host, src_dev, n, cudaMemcpyDeviceToHost, s->s));
  return NNSB_OK;
}

extern "C" int
nnsb_d2d_async (nnsb_stream * s, void *dst_dev, const void *src_dev, size_t n)
{
  STREAM_PROLOGUE ("nnsb_d2d_async");
  if (n == 0)
    return NNSB_OK;
  CUDA_TRY (cudaMemcpyAsync (dst_dev, src_dev, n, cudaMemcpyDeviceToDevice, s->s));
  return NNSB_OK;
}

extern "C" int
nnsb_h2d_2d_async (nnsb_stream * s, void *dst_dev, size_t dst_pitch, const void *src_host,
    size_t src_pitch, size_t row_bytes, size_t rows)
{
  STREAM_PROLOGUE ("nnsb_h2d_2d_async");
  if (rows == 0 || row_bytes == 0)
    return NNSB_OK;
  CUDA_TRY (cudaMemcpy2DAsync (dst_dev, dst_pitch, src_host, src_pitch, row_bytes, rows,
          cudaMemcpyHostToDevice, s->s));
  return NNSB_OK;
}

/* ------------------------------------------------------------------------------------- */
/* plans                                                                                   */
/* ------------------------------------------------------------------------------------- */
static bool
is_64bit_int (int t)
{
  return t == NNSB_INT64 || t == NNSB_UINT64;
}

static LaunchCfg
cfg_of (const nnsb_ctx * ctx)
{
  LaunchCfg c;
  c.sm_count = ctx->sm_count;
  c.blocks_per_sm = ctx->blocks_per_sm;
  return c;
}

extern "C" int
nnsb_plan_destroy (nnsb_plan * plan)
{
  if (!plan)
    return NNSB_OK;
  if (plan->d_table)
    nnsb_dev_free (plan->ctx, plan->d_table);
  if (plan->d_perm_table)
    nnsb_dev_free (plan->ctx, plan->d_perm_table);
  if (plan->d_scratch)
    nnsb_dev_free (plan->ctx, plan->d_scratch);
  if (plan->d_stats)
    nnsb_dev_free (plan->ctx, plan->d_stats);
  delete plan;
  return NNSB_OK;
}

static int attach_perm_vec (nnsb_ctx * ctx, nnsb_plan * plan);

extern "C" int
nnsb_plan_create (nnsb_ctx * ctx, const nnsb_option * opt, const nnsb_tensor_info * in_info,
    int acceleration, nnsb_plan ** out)
{
  if (!out || !opt || !in_info)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_create: null argument");
  *out = nullptr;
  int r = bind (ctx);
  if (r)
    return r;
  if (in_info->type < 0 || in_info->type >= NNSB_TYPE_END || !nnsb_dim_is_valid (in_info->dim))
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_create: invalid input tensor info");

  nnsb_plan *plan = new (std::nothrow) nnsb_plan ();
  if (!plan)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_plan_create: out of memory");
  memset (plan, 0, sizeof (*plan));
  plan->ctx = ctx;
  plan->opt = *opt;
  plan->in_info = *in_info;
  r = nnsb_convert_dimension (opt, NNSB_PAD_SINK, in_info, &plan->out_info);
  if (r) {
    delete plan;
    return r;
  }
  if (plan->out_info.type < 0 || plan->out_info.type >= NNSB_TYPE_END) {
    delete plan;
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_create: invalid output type");
  }
  plan->in_size = nnsb_info_size (&plan->in_info);
  plan->out_size = nnsb_info_size (&plan->out_info);
  plan->frame_elems = nnsb_element_count (in_info->dim);
  /* orc_supported (gsttensor_transform.c:462-463) */
  plan->acceleration = acceleration && !is_64bit_int (plan->in_info.type)
      && !is_64bit_int (plan->out_info.type);

  switch (opt->mode) {
    case NNSB_MODE_TYPECAST:
      memset (&plan->chain, 0, sizeof (plan->chain));
      plan->chain.desc.nch = 1;
      plan->chain.desc.ch_size = 1;
      break;
    case NNSB_MODE_ARITHMETIC:
    {
      nnsb_cell *table = nullptr;
      r = nnsb_chain_compile (opt, plan->out_info.type, plan->acceleration, in_info->dim,
          &plan->chain.desc, &table);
      if (r == NNSB_OK && table) {
        const size_t bytes = sizeof (nnsb_cell) * plan->chain.desc.n_pos * plan->chain.desc.nch;
        void *d = nullptr;
        r = nnsb_dev_alloc (ctx, bytes, &d);
        if (r == NNSB_OK) {
          plan->d_table = (nnsb_cell *) d;
          cudaError_t e = cudaMemcpy (d, table, bytes, cudaMemcpyHostToDevice);
          if (e != cudaSuccess)
            r = nnsb_set_error (NNSB_ECUDA, "cudaMemcpy (chain table) failed: %s",
                cudaGetErrorString (e));
        }
        plan->chain.table = plan->d_table;
      }
      if (r == NNSB_OK && plan->out_info.type == NNSB_FLOAT32 && !getenv ("NNSB_CHAIN_NO_FAST")) {
        /* the fast float32 kernel needs a reciprocal next to every divisor */
        const nnsb_chain_desc & dsc = plan->chain.desc;
        bool ok = true;
        for (int i = 0; i < dsc.n_pos && ok; i++) {
          if (dsc.opcode[i] != NNSB_OP_DIV)
            continue;
          if (table) {
            for (int c = 0; c < dsc.nch && ok; c++)
              ok = (table[(size_t) i * dsc.nch + c].raw >> 32) != 0;
          } else {
            ok = (dsc.scalar[i].raw >> 32) != 0;
          }
        }
        plan->chain.fast_f32 = ok ? 1 : 0;
      }
      free (table);
      if (r) {
        nnsb_plan_destroy (plan);
        return r;
      }
      break;
    }
    case NNSB_MODE_TRANSPOSE:
    {
      /* out dim i <- in dim order[i] (gsttensor_transform.c:1924-1928); dims 4..7 stay */
      uint64_t in_strides[NNSB_RANK_LIMIT];
      uint64_t s = 1;
      for (int i = 0; i < NNSB_RANK_LIMIT; i++) {
        in_strides[i] = s;
        s *= in_info->dim[i];
      }
      plan->perm.rank = NNSB_RANK_LIMIT;
      plan->perm.es = (int) nnsb_element_size (in_info->type);
      for (int i = 0; i < NNSB_RANK_LIMIT; i++) {
        const int src = i < NNSB_TRANSPOSE_RANK_LIMIT ? opt->trans_order[i] : i;
        plan->perm.size[i] = in_info->dim[src];
        plan->perm.in_stride[i] = in_strides[src];
      }
      break;
    }
    case NNSB_MODE_DIMCHG:
    {
      const int from = opt->dimchg_from, to = opt->dimchg_to;
      if (from > to) {
        /* gsttensor_transform.c:1179-1188: negotiates, then fails at run time */
        nnsb_plan_destroy (plan);
        return nnsb_set_error (NNSB_ENOTSUP,
            "tensor-transform/dimchg operation is not permitted if from >= to.");
      }
      uint64_t in_strides[NNSB_RANK_LIMIT];
      uint64_t s = 1;
      for (int i = 0; i < NNSB_RANK_LIMIT; i++) {
        in_strides[i] = s;
        s *= in_info->dim[i];
      }
      plan->perm.rank = NNSB_RANK_LIMIT;
      plan->perm.es = (int) nnsb_element_size (in_info->type);
      for (int i = 0; i < NNSB_RANK_LIMIT; i++) {
        int src;
        if (i < from || i > to || from == to)
          src = i;
        else if (i == to)
          src = from;
        else
          src = i + 1;
        plan->perm.size[i] = in_info->dim[src];
        plan->perm.in_stride[i] = in_strides[src];
      }
      break;
    }
    case NNSB_MODE_STAND:
      if (opt->stand_mode != NNSB_STAND_DEFAULT && opt->stand_mode != NNSB_STAND_DC_AVERAGE) {
        nnsb_plan_destroy (plan);
        return nnsb_set_error (NNSB_EINVAL, "stand: cannot identify mode");
      }
      plan->stand_nch = opt->stand_per_channel ? in_info->dim[0] : 1;
      plan->no_fused = getenv ("NNSB_STAND_NO_FUSED") != nullptr;
      break;
    case NNSB_MODE_CLAMP:
      break;
    default:
      nnsb_plan_destroy (plan);
      return nnsb_set_error (NNSB_ENOTSUP, "Not supported tensor transform mode");
  }
  if (opt->mode == NNSB_MODE_TRANSPOSE || opt->mode == NNSB_MODE_DIMCHG) {
    r = attach_perm_vec (ctx, plan);
    if (r) {
      nnsb_plan_destroy (plan);
      return r;
    }
  }
  *out = plan;
  return NNSB_OK;
}

static int
attach_perm_vec (nnsb_ctx * ctx, nnsb_plan * plan)
{
  /* per-box table of the vector kernel: built and verified on the host, kept in HBM */
  void *host_table = nullptr;
  int r = NNSB_OK;
  if (perm_vec_build (plan->perm, plan->frame_elems, &plan->perm_vec, &host_table) == 0) {
    void *d = nullptr;
    r = nnsb_dev_alloc (ctx, plan->perm_vec.table_bytes, &d);
    if (r == NNSB_OK) {
      plan->d_perm_table = d;
      cudaError_t e = cudaMemcpy (d, host_table, plan->perm_vec.table_bytes,
          cudaMemcpyHostToDevice);
      if (e != cudaSuccess)
        r = nnsb_set_error (NNSB_ECUDA, "cudaMemcpy (permute table) failed: %s",
            cudaGetErrorString (e));
    }
    free (host_table);
  }
  return r;
}

extern "C" int
nnsb_plan_create_permute (nnsb_ctx * ctx, int es, int rank, const uint32_t *size,
    const uint64_t *in_stride, nnsb_plan ** out)
{
  if (!out || !size || !in_stride)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_create_permute: null argument");
  *out = nullptr;
  int r = bind (ctx);
  if (r)
    return r;
  if (rank < 1 || rank > NNSB_PERM_MAX_RANK)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_create_permute: rank must be 1..%d",
        NNSB_PERM_MAX_RANK);
  int type;
  switch (es) {
    case 1: type = NNSB_UINT8; break;
    case 2: type = NNSB_UINT16; break;
    case 4: type = NNSB_UINT32; break;
    case 8: type = NNSB_UINT64; break;
    default:
      return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_create_permute: element size must be 1, 2, 4 or 8");
  }
  nnsb_plan *plan = new (std::nothrow) nnsb_plan ();
  if (!plan)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_plan_create_permute: out of memory");
  memset (plan, 0, sizeof (*plan));
  plan->ctx = ctx;
  plan->opt.mode = NNSB_MODE_TRANSPOSE;
  uint64_t elems = 1;
  plan->perm.rank = rank;
  plan->perm.es = es;
  for (int i = 0; i < NNSB_RANK_LIMIT; i++)
    plan->in_info.dim[i] = 1;
  for (int i = 0; i < rank; i++) {
    if (size[i] == 0) {
      delete plan;
      return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_create_permute: zero extent");
    }
    plan->perm.size[i] = size[i];
    plan->perm.in_stride[i] = in_stride[i];
    plan->in_info.dim[i] = size[i];
    elems *= size[i];
  }
  plan->in_info.type = type;
  plan->out_info = plan->in_info;
  plan->frame_elems = elems;
  plan->in_size = plan->out_size = (size_t) elems * es;
  r = attach_perm_vec (ctx, plan);
  if (r) {
    nnsb_plan_destroy (plan);
    return r;
  }
  *out = plan;
  return NNSB_OK;
}

extern "C" int
nnsb_plan_out_info (const nnsb_plan * plan, nnsb_tensor_info * out_info)
{
  if (!plan || !out_info)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_out_info: null argument");
  *out_info = plan->out_info;
  return NNSB_OK;
}

extern "C" size_t
nnsb_plan_in_size (const nnsb_plan * plan)
{
  return plan ? plan->in_size : 0;
}

extern "C" size_t
nnsb_plan_out_size (const nnsb_plan * plan)
{
  return plan ? plan->out_size : 0;
}

extern "C" int
nnsb_plan_last_launch_count (const nnsb_plan * plan)
{
  return plan ? plan->last_launches : NNSB_EINVAL;
}

extern "C" int
nnsb_plan_describe (const nnsb_plan * plan, char *buf, size_t buflen)
{
  if (!plan || !buf || buflen == 0)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_describe: null argument");
  int len = 0;
  switch (plan->opt.mode) {
    case NNSB_MODE_TRANSPOSE:
    case NNSB_MODE_DIMCHG:
      if (plan->perm_vec.valid) {
        char box[128];
        int o = 0;
        for (int i = 0; i < plan->perm_vec.n && o < (int) sizeof (box) - 12; i++)
          o += snprintf (box + o, sizeof (box) - o, "%s%u", i ? "x" : "", plan->perm_vec.tile[i]);
        len = snprintf (buf, buflen, "permute: vector table, %d-byte elements, box %s, %u vectors, "
            "swizzle %d/%d, %s%llu gather wavefronts", plan->perm_vec.es, box, plan->perm_vec.n_vec,
            plan->perm_vec.swizzle_shift, plan->perm_vec.swizzle_shift2,
            plan->perm_vec.micro ? "register micro-transposes, " : "",
            (unsigned long long) plan->perm_vec.gather_wavefronts);
      } else {
        len = snprintf (buf, buflen, "permute: scalar box kernel");
      }
      if (len > 0 && (size_t) len < buflen) {
        /* what a batch of these tensors takes when the planar fast path does not apply */
        const char *path = permute_path_name (plan->perm, plan->frame_elems, 2, &plan->perm_vec);
        if (strcmp (path, "slab") == 0)
          len += snprintf (buf + len, buflen - len, "; batches: contiguous slabs through shared memory");
        else if (strcmp (path, "pixel") == 0)
          len += snprintf (buf + len, buflen - len, "; aligned frames: RGB pixel transpose kernel");
      }
      break;
    case NNSB_MODE_TYPECAST:
    case NNSB_MODE_ARITHMETIC:
      len = snprintf (buf, buflen, "chain: %d position(s), %d channel(s)%s", (int) plan->chain.desc.n_pos,
          (int) plan->chain.desc.nch, plan->chain.fast_f32 ? ", float32 fast path" : "");
      break;
    case NNSB_MODE_STAND:
      len = snprintf (buf, buflen, "stand: %u channel(s)%s", plan->stand_nch,
          plan->stand_fused_used ? ", single-read fused kernel" : "");
      break;
    case NNSB_MODE_CLAMP:
      len = snprintf (buf, buflen, "clamp");
      break;
    default:
      return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_describe: unknown mode");
  }
  if (len < 0 || (size_t) len >= buflen)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_describe: buffer too small");
  return len;
}

static int
grow (nnsb_ctx * ctx, double **buf, size_t * have, size_t need)
{
  if (*have >= need && *buf)
    return NNSB_OK;
  if (*buf) {
    /* a previous launch may still be reading the old buffer */
    cudaDeviceSynchronize ();
    nnsb_dev_free (ctx, *buf);
  }
  *buf = nullptr;
  *have = 0;
  void *p = nullptr;
  int r = nnsb_dev_alloc (ctx, need * sizeof (double), &p);
  if (r)
    return r;
  *buf = (double *) p;
  *have = need;
  return NNSB_OK;
}

extern "C" int
nnsb_plan_launch (nnsb_plan * plan, const void *d_in, void *d_out, uint32_t batch,
    nnsb_stream * s)
{
  if (!plan || !s)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_launch: null argument");
  int r = bind (plan->ctx);
  if (r)
    return r;
  plan->last_launches = 0;
  if (batch == 0)
    return NNSB_OK;
  if (!d_in || !d_out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_launch: null device pointer");
  const LaunchCfg cfg = cfg_of (plan->ctx);
  const size_t n = (size_t) plan->frame_elems * batch;
  int err = 0;
  switch (plan->opt.mode) {
    case NNSB_MODE_TYPECAST:
    case NNSB_MODE_ARITHMETIC:
      err = launch_chain (plan->in_info.type, plan->out_info.type, plan->acceleration != 0, d_in,
          d_out, n, plan->chain, cfg, s->s);
      plan->last_launches = 1;
      break;
    case NNSB_MODE_CLAMP:
      err = launch_clamp (plan->in_info.type, d_in, d_out, n, plan->opt.clamp_min,
          plan->opt.clamp_max, cfg, s->s);
      plan->last_launches = 1;
      break;
    case NNSB_MODE_TRANSPOSE:
    case NNSB_MODE_DIMCHG:
      err = launch_permute (plan->perm, d_in, d_out, plan->frame_elems, batch, cfg, s->s,
          &plan->perm_vec, plan->d_perm_table);
      plan->last_launches = 1;
      break;
    case NNSB_MODE_STAND:
    {
      if (!plan->no_fused && stand_fused_supported (plan->in_info.type, plan->out_info.type, d_in,
              d_out, plan->frame_elems, plan->stand_nch, batch, cfg)) {
        /* batches of float32 frames that fit the SMs' shared memory: one read of the input */
        r = grow (plan->ctx, &plan->d_scratch, &plan->scratch_doubles,
            (stand_fused_scratch_bytes (cfg) + 7) / 8);
        if (r)
          return r;
        r = grow (plan->ctx, &plan->d_stats, &plan->stats_doubles, (size_t) batch * plan->stand_nch * 2);
        if (r)
          return r;
        int nl = 0;
        err = launch_stand_fused (plan->opt.stand_mode, d_in, d_out, plan->frame_elems,
            plan->stand_nch, batch, plan->d_scratch, plan->d_stats, cfg, s->s, &nl);
        plan->last_launches = nl;
        plan->stand_fused_used = 1;
        break;
      }
      plan->stand_fused_used = 0;
      const size_t need_scratch = stand_scratch_doubles (batch, plan->stand_nch,
          plan->frame_elems, cfg);
      const size_t need_stats = (size_t) batch * plan->stand_nch * 2;
      r = grow (plan->ctx, &plan->d_scratch, &plan->scratch_doubles, need_scratch);
      if (r)
        return r;
      r = grow (plan->ctx, &plan->d_stats, &plan->stats_doubles, need_stats);
      if (r)
        return r;
      int nl = 0;
      err = launch_stand (plan->in_info.type, plan->out_info.type, plan->opt.stand_mode, d_in,
          d_out, plan->frame_elems, plan->stand_nch, batch, plan->d_scratch, plan->d_stats, cfg,
          s->s, &nl);
      plan->last_launches = nl;
      break;
    }
    default:
      return nnsb_set_error (NNSB_ENOTSUP, "Not supported tensor transform mode");
  }
  if (err)
    return nnsb_set_error (NNSB_ECUDA, "kernel launch failed: %s",
        cudaGetErrorString ((cudaError_t) err));
  return NNSB_OK;
}

extern "C" int
nnsb_plan_is_elementwise (const nnsb_plan * plan)
{
  if (!plan)
    return 0;
  return plan->opt.mode == NNSB_MODE_TYPECAST || plan->opt.mode == NNSB_MODE_ARITHMETIC
      || plan->opt.mode == NNSB_MODE_CLAMP;
}

extern "C" uint64_t
nnsb_plan_chunk_granule (const nnsb_plan * plan)
{
  /* element counts that are a multiple of this keep every launch vector-aligned and, for
   * per-channel arithmetic, start on channel 0 of a channel run */
  if (!plan)
    return 0;
  uint64_t g = 256;             /* 256 elements: 32-byte aligned for every element size */
  if (plan->opt.mode == NNSB_MODE_ARITHMETIC && plan->chain.desc.nch > 1) {
    const uint64_t period = plan->chain.desc.ch_size * (uint64_t) plan->chain.desc.nch;
    uint64_t a = g, b = period;
    while (b) { const uint64_t t = a % b; a = b; b = t; }
    g = g / a * period;
  }
  return g;
}

extern "C" int
nnsb_plan_launch_elems (nnsb_plan * plan, const void *d_in, void *d_out, uint64_t first_elem,
    uint64_t n_elems, nnsb_stream * s)
{
  if (!plan || !s || !d_in || !d_out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_launch_elems: null argument");
  if (!nnsb_plan_is_elementwise (plan))
    return nnsb_set_error (NNSB_ENOTSUP,
        "nnsb_plan_launch_elems: only typecast, arithmetic and clamp can be split by elements");
  if (first_elem % nnsb_plan_chunk_granule (plan))
    return nnsb_set_error (NNSB_EINVAL,
        "nnsb_plan_launch_elems: first element must be a multiple of nnsb_plan_chunk_granule()");
  int r = bind (plan->ctx);
  if (r)
    return r;
  plan->last_launches = 0;
  if (n_elems == 0)
    return NNSB_OK;
  const LaunchCfg cfg = cfg_of (plan->ctx);
  const char *in = (const char *) d_in + first_elem * nnsb_element_size (plan->in_info.type);
  char *out = (char *) d_out + first_elem * nnsb_element_size (plan->out_info.type);
  int err;
  if (plan->opt.mode == NNSB_MODE_CLAMP)
    err = launch_clamp (plan->in_info.type, in, out, n_elems, plan->opt.clamp_min,
        plan->opt.clamp_max, cfg, s->s);
  else
    err = launch_chain (plan->in_info.type, plan->out_info.type, plan->acceleration != 0, in, out,
        n_elems, plan->chain, cfg, s->s);
  plan->last_launches = 1;
  if (err)
    return nnsb_set_error (NNSB_ECUDA, "kernel launch failed: %s",
        cudaGetErrorString ((cudaError_t) err));
  return NNSB_OK;
}

extern "C" int
nnsb_plan_stand_stats (nnsb_plan * plan, nnsb_stream * s, uint32_t batch, double *mean_out,
    double *std_out, size_t capacity)
{
  if (!plan || !s || plan->opt.mode != NNSB_MODE_STAND)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_stand_stats: not a stand plan");
  int r = bind (plan->ctx);
  if (r)
    return r;
  const size_t cnt = (size_t) batch * plan->stand_nch;
  if (cnt > capacity || cnt * 2 > plan->stats_doubles)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_plan_stand_stats: capacity too small");
  std::vector<double> tmp (cnt * 2);
  CUDA_TRY (cudaMemcpyAsync (tmp.data (), plan->d_stats, cnt * 2 * sizeof (double),
          cudaMemcpyDeviceToHost, s->s));
  CUDA_TRY (cudaStreamSynchronize (s->s));
  for (size_t i = 0; i < cnt; i++) {
    if (mean_out)
      mean_out[i] = tmp[2 * i];
    if (std_out)
      std_out[i] = tmp[2 * i + 1];
  }
  return NNSB_OK;
}

extern "C" int
nnsb_merge_launch (nnsb_stream * s, uint32_t n, const void *const *d_in, const uint64_t *block_bytes,
    void *d_out, uint64_t rows)
{
  STREAM_PROLOGUE ("nnsb_merge_launch");
  if (!d_in || !block_bytes || !d_out || n < 1 || n > NNSB_MERGE_MAX)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_merge_launch: invalid argument");
  for (uint32_t k = 0; k < n; k++)
    if (!d_in[k] || block_bytes[k] == 0)
      return nnsb_set_error (NNSB_EINVAL, "nnsb_merge_launch: input %u is empty", k);
  int err = launch_merge ((int) n, d_in, block_bytes, d_out, rows, cfg_of (s->ctx), s->s);
  if (err)
    return nnsb_set_error (NNSB_ECUDA, "merge launch failed: %s",
        cudaGetErrorString ((cudaError_t) err));
  return NNSB_OK;
}

extern "C" int
nnsb_pad_launch (nnsb_stream * s, const void *d_src, void *d_dst, size_t row_bytes,
    size_t dst_pitch, size_t rows)
{
  STREAM_PROLOGUE ("nnsb_pad_launch");
  if (!d_src || !d_dst)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_pad_launch: null device pointer");
  int err = launch_pad (d_src, d_dst, row_bytes, dst_pitch, rows, cfg_of (s->ctx), s->s);
  if (err)
    return nnsb_set_error (NNSB_ECUDA, "pad launch failed: %s",
        cudaGetErrorString ((cudaError_t) err));
  return NNSB_OK;
}

extern "C" int
nnsb_d2h_2d_async (nnsb_stream * s, void *dst_host, size_t dst_pitch, const void *src_dev,
    size_t src_pitch, size_t row_bytes, size_t rows)
{
  STREAM_PROLOGUE ("nnsb_d2h_2d_async");
  if (rows == 0 || row_bytes == 0)
    return NNSB_OK;
  CUDA_TRY (cudaMemcpy2DAsync (dst_host, dst_pitch, src_dev, src_pitch, row_bytes, rows,
          cudaMemcpyDeviceToHost, s->s));
  return NNSB_OK;
}

extern "C" int
nnsb_strip_launch (nnsb_stream * s, const void *d_src, void *d_dst, size_t row_bytes,
    size_t src_pitch, size_t rows)
{
  STREAM_PROLOGUE ("nnsb_strip_launch");
  if (!d_src || !d_dst)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_strip_launch: null device pointer");
  int err = launch_strip (d_src, d_dst, row_bytes, src_pitch, rows, cfg_of (s->ctx), s->s);
  if (err)
    return nnsb_set_error (NNSB_ECUDA, "strip launch failed: %s",
        cudaGetErrorString ((cudaError_t) err));
  return NNSB_OK;
}
