/* CUDA-device GstAllocator for the B200 tensor path (see gstnnsbcudaallocator.c). */
#ifndef GST_NNSB_CUDA_ALLOCATOR_H
#define GST_NNSB_CUDA_ALLOCATOR_H

#include <gst/gst.h>
#include <nnstreamer_b200.h>

G_BEGIN_DECLS

#define GST_NNSB_CUDA_MEMORY_TYPE "nnsb.cuda"

typedef struct _GstNnsbCudaMemory
{
  GstMemory mem;
  nnsb_memory dev;              /* device block (owned by the root memory only) */
  gpointer host;                /* pinned staging, filled lazily by mem_map */
  gboolean host_valid;          /* host copy reflects the device data */
  gboolean host_dirty;          /* mapped for WRITE: upload on unmap */
  GMutex lock;
} GstNnsbCudaMemory;

GstAllocator *gst_nnsb_cuda_allocator_new (nnsb_transform * owner);
gboolean gst_is_nnsb_cuda_memory (GstMemory * mem);
/* wrap a device block produced by nnsb_transform_transform (takes ownership) */
GstMemory *gst_nnsb_cuda_memory_wrap (GstAllocator * allocator, const nnsb_memory * dev);
/* device pointer of a CUDA-aware peer's memory (no host round trip) */
gboolean gst_nnsb_cuda_memory_get (GstMemory * mem, nnsb_memory * out);

G_END_DECLS
#endif
