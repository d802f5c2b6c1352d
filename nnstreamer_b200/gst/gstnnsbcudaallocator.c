/*
 * gstnnsbcudaallocator.c -- GstAllocator whose memories live in B200 HBM.
 *
 * Replaces the role of gst_allocator_alloc (NULL, ...) in the reference's transform
 * (gsttensor_transform.c:1758) and follows the precedent of tensor_allocator.c:101-128 (an
 * allocator wrapper installed by the element).  Tensors stay resident on the device between
 * chained CUDA-aware elements: such a peer recognises the memory by its type string
 * (GST_NNSB_CUDA_MEMORY_TYPE) and takes the device view with gst_nnsb_cuda_memory_get() -- the
 * view carries the block whose "ready" event the peer's stream waits on, so nothing
 * synchronises with the host.  Any other element simply maps the memory: mem_map waits for
 * the producer, copies device -> pinned host once; mem_unmap after a WRITE map copies back.
 *
 * Life time: every GstMemory holds its own reference on the device block (nnsb_memory_ref /
 * nnsb_memory_share), the block keeps its context alive, and the allocator has a context and a
 * stream of its own -- so a memory that is still held downstream when the element that
 * produced it is finalized stays valid.
 *
 * Not compiled in this repository's image (no GStreamer): `make syntax` in this directory
 * checks it against declaration stubs; see README.md.
 */
#include <string.h>
#include "gstnnsbcudaallocator.h"

typedef struct _GstNnsbCudaAllocator
{
  GstAllocator parent;
  nnsb_ctx *ctx;                /* pools for memories allocated through gst_allocator_alloc */
  nnsb_stream *stream;          /* copy stream of mem_map / mem_unmap */
} GstNnsbCudaAllocator;

typedef struct _GstNnsbCudaAllocatorClass
{
  GstAllocatorClass parent_class;
} GstNnsbCudaAllocatorClass;

G_DEFINE_TYPE (GstNnsbCudaAllocator, gst_nnsb_cuda_allocator, GST_TYPE_ALLOCATOR);

/* takes over the reference held by `dev` */
static GstMemory *
_new_memory (GstAllocator * allocator, GstMemory * parent, const nnsb_memory * dev, gsize maxsize,
    gsize offset, gsize size)
{
  GstNnsbCudaMemory *m = g_new0 (GstNnsbCudaMemory, 1);
  gst_memory_init (GST_MEMORY_CAST (m), 0, allocator, parent, maxsize, 255, offset, size);
  m->dev = *dev;
  memset (&m->host, 0, sizeof (m->host));
  g_mutex_init (&m->lock);
  return GST_MEMORY_CAST (m);
}

static GstMemory *
_alloc (GstAllocator * allocator, gsize size, GstAllocationParams * params)
{
  GstNnsbCudaAllocator *self = (GstNnsbCudaAllocator *) allocator;
  nnsb_memory dev;
  (void) params;
  if (nnsb_memory_alloc (self->ctx, size, 1, self->stream, &dev) != NNSB_OK) {
    GST_ERROR_OBJECT (allocator, "device allocation of %" G_GSIZE_FORMAT " bytes failed: %s", size,
        nnsb_last_error ());
    return NULL;
  }
  return _new_memory (allocator, NULL, &dev, size, 0, size);
}

static void
_free (GstAllocator * allocator, GstMemory * mem)
{
  GstNnsbCudaMemory *m = (GstNnsbCudaMemory *) mem;
  (void) allocator;
  nnsb_memory_release (&m->host);
  nnsb_memory_release (&m->dev);        /* the block returns to its pool with the last reference */
  g_mutex_clear (&m->lock);
  g_free (m);
}

/* a peer that is not CUDA-aware maps the memory: give it host bytes */
static gpointer
_map (GstMemory * mem, gsize maxsize, GstMapFlags flags)
{
  GstNnsbCudaMemory *m = (GstNnsbCudaMemory *) mem;
  GstNnsbCudaAllocator *self = (GstNnsbCudaAllocator *) mem->allocator;
  gpointer ret = NULL;
  (void) maxsize;
  g_mutex_lock (&m->lock);
  if (!m->host.data)
    nnsb_memory_alloc (self->ctx, m->dev.size, 0, NULL, &m->host);
  if (m->host.data) {
    gboolean ok = TRUE;
    if (!m->host_valid && (flags & GST_MAP_READ)) {
      /* wait (on the GPU) for the producer, device -> pinned host, then wait for the copy */
      ok = nnsb_memory_wait (&m->dev, self->stream) == NNSB_OK
          && nnsb_d2h_async (self->stream, m->host.data, m->dev.data, m->dev.size) == NNSB_OK
          && nnsb_stream_sync (self->stream) == NNSB_OK;
      m->host_valid = ok;
    }
    if (ok) {
      if (flags & GST_MAP_WRITE)
        m->host_dirty = TRUE;
      ret = m->host.data;
    }
  }
  g_mutex_unlock (&m->lock);
  return ret;
}

static void
_unmap (GstMemory * mem)
{
  GstNnsbCudaMemory *m = (GstNnsbCudaMemory *) mem;
  GstNnsbCudaAllocator *self = (GstNnsbCudaAllocator *) mem->allocator;
  g_mutex_lock (&m->lock);
  if (m->host_dirty) {
    /* the upload is stream-ordered: the block is "ready" again once it has landed, and the
     * pinned source must not change until then -- wait here, a WRITE map is the slow path */
    if (nnsb_memory_wait (&m->dev, self->stream) == NNSB_OK
        && nnsb_h2d_async (self->stream, m->dev.data, m->host.data, m->dev.size) == NNSB_OK)
      nnsb_memory_mark_ready (&m->dev, self->stream);
    nnsb_stream_sync (self->stream);
    m->host_dirty = FALSE;
    m->host_valid = TRUE;
  }
  g_mutex_unlock (&m->lock);
}

/* gst_memory_share: tensor_mux/demux and gst_tensor_buffer_from_config
 * (nnstreamer_plugin_api_impl.c:534) hand out sub-ranges of one block */
static GstMemory *
_share (GstMemory * mem, gssize offset, gssize size)
{
  GstNnsbCudaMemory *m = (GstNnsbCudaMemory *) mem;
  GstMemory *parent = mem->parent ? mem->parent : mem;
  nnsb_memory dev;
  if (size == -1)
    size = (gssize) mem->size - offset;
  if (nnsb_memory_share (&m->dev, (size_t) offset, (size_t) size, &dev) != NNSB_OK)
    return NULL;
  return _new_memory (mem->allocator, parent, &dev, mem->maxsize, mem->offset + (gsize) offset,
      (gsize) size);
}

static void
_finalize (GObject * object)
{
  GstNnsbCudaAllocator *self = (GstNnsbCudaAllocator *) object;
  if (self->stream)
    nnsb_stream_destroy (self->stream);
  if (self->ctx)
    nnsb_ctx_destroy (self->ctx);
  G_OBJECT_CLASS (gst_nnsb_cuda_allocator_parent_class)->finalize (object);
}

static void
gst_nnsb_cuda_allocator_class_init (GstNnsbCudaAllocatorClass * klass)
{
  GstAllocatorClass *ac = GST_ALLOCATOR_CLASS (klass);
  ac->alloc = _alloc;
  ac->free = _free;
  G_OBJECT_CLASS (klass)->finalize = _finalize;
}

static void
gst_nnsb_cuda_allocator_init (GstNnsbCudaAllocator * self)
{
  GstAllocator *alloc = GST_ALLOCATOR_CAST (self);
  alloc->mem_type = GST_NNSB_CUDA_MEMORY_TYPE;
  alloc->mem_map = _map;
  alloc->mem_unmap = _unmap;
  alloc->mem_share = _share;
  GST_OBJECT_FLAG_SET (self, GST_ALLOCATOR_FLAG_CUSTOM_ALLOC);
}

GstAllocator *
gst_nnsb_cuda_allocator_new (int device_id)
{
  GstNnsbCudaAllocator *self = g_object_new (gst_nnsb_cuda_allocator_get_type (), NULL);
  if (nnsb_ctx_create (device_id, &self->ctx) != NNSB_OK
      || nnsb_stream_create (self->ctx, &self->stream) != NNSB_OK) {
    GST_ERROR_OBJECT (self, "no usable GPU %d: %s", device_id, nnsb_last_error ());
    gst_object_unref (self);
    return NULL;
  }
  return GST_ALLOCATOR_CAST (self);
}

gboolean
gst_is_nnsb_cuda_memory (GstMemory * mem)
{
  return mem && mem->allocator && g_strcmp0 (mem->allocator->mem_type, GST_NNSB_CUDA_MEMORY_TYPE) == 0;
}

GstMemory *
gst_nnsb_cuda_memory_wrap (GstAllocator * allocator, const nnsb_memory * dev)
{
  return _new_memory (allocator, NULL, dev, dev->size, 0, dev->size);
}

gboolean
gst_nnsb_cuda_memory_get (GstMemory * mem, nnsb_memory * out)
{
  if (!gst_is_nnsb_cuda_memory (mem))
    return FALSE;
  *out = ((GstNnsbCudaMemory *) mem)->dev;
  out->owned = 0;               /* borrowed: the GstMemory keeps the reference */
  return TRUE;
}
