/*
 * gstnnsbcudaallocator.c -- GstAllocator whose memories live in B200 HBM.
 *
 * Replaces the role of gst_allocator_alloc (NULL, ...) in the reference's transform
 * (gsttensor_transform.c:1758) and follows the precedent of tensor_allocator.c:101-128 (an
 * allocator wrapper installed by the element).  Tensors stay resident on the device between
 * chained CUDA-aware elements: such a peer recognises the memory by its type string
 * (GST_NNSB_CUDA_MEMORY_TYPE) and takes the device pointer with gst_nnsb_cuda_memory_get().
 * Any other element simply maps the memory: mem_map then copies device -> pinned host once,
 * mem_unmap after a WRITE map copies back.  NOT compiled in this repository's image (no
 * GStreamer); see README.md in this directory.
 */
#include <string.h>
#include "gstnnsbcudaallocator.h"

typedef struct _GstNnsbCudaAllocator
{
  GstAllocator parent;
  nnsb_transform *owner;        /* element whose context/pools back the memories */
} GstNnsbCudaAllocator;

typedef struct _GstNnsbCudaAllocatorClass
{
  GstAllocatorClass parent_class;
} GstNnsbCudaAllocatorClass;

G_DEFINE_TYPE (GstNnsbCudaAllocator, gst_nnsb_cuda_allocator, GST_TYPE_ALLOCATOR);

static GstMemory *
_new_memory (GstAllocator * allocator, GstMemory * parent, const nnsb_memory * dev, gsize maxsize,
    gsize offset, gsize size)
{
  GstNnsbCudaMemory *m = g_new0 (GstNnsbCudaMemory, 1);
  gst_memory_init (GST_MEMORY_CAST (m), 0, allocator, parent, maxsize, 255, offset, size);
  m->dev = *dev;
  if (parent)
    m->dev.owned = 0;           /* sub-memories borrow the root's block */
  g_mutex_init (&m->lock);
  return GST_MEMORY_CAST (m);
}

static GstMemory *
_alloc (GstAllocator * allocator, gsize size, GstAllocationParams * params)
{
  GstNnsbCudaAllocator *self = (GstNnsbCudaAllocator *) allocator;
  nnsb_memory dev;
  if (nnsb_transform_alloc (self->owner, size, 1, &dev) != NNSB_OK) {
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
  if (m->host) {
    nnsb_memory h = { m->host, mem->maxsize, 0, 1, m->dev.owner };
    nnsb_memory_release (&h);
  }
  nnsb_memory_release (&m->dev);        /* no-op for sub-memories (owned == 0) */
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
  g_mutex_lock (&m->lock);
  if (!m->host) {
    nnsb_memory h;
    if (nnsb_transform_alloc (self->owner, mem->maxsize, 0, &h) == NNSB_OK)
      m->host = h.data;
  }
  if (m->host) {
    if (!m->host_valid && (flags & GST_MAP_READ)) {
      /* device -> pinned host, stream-ordered then synchronised */
      if (nnsb_transform_download (self->owner, m->host, &m->dev) == NNSB_OK)
        m->host_valid = TRUE;
    }
    if (flags & GST_MAP_WRITE)
      m->host_dirty = TRUE;
    ret = m->host;
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
    nnsb_transform_upload (self->owner, &m->dev, m->host);
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
  nnsb_memory dev = m->dev;
  if (size == -1)
    size = mem->size - offset;
  dev.data = (char *) m->dev.data + offset;
  dev.size = size;
  return _new_memory (mem->allocator, parent, &dev, mem->maxsize, mem->offset + offset, size);
}

static void
gst_nnsb_cuda_allocator_class_init (GstNnsbCudaAllocatorClass * klass)
{
  GstAllocatorClass *ac = GST_ALLOCATOR_CLASS (klass);
  ac->alloc = _alloc;
  ac->free = _free;
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
gst_nnsb_cuda_allocator_new (nnsb_transform * owner)
{
  GstNnsbCudaAllocator *self = g_object_new (gst_nnsb_cuda_allocator_get_type (), NULL);
  self->owner = owner;
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
  out->owned = 0;
  return TRUE;
}
