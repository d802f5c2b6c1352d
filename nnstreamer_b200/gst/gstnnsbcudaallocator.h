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
  nnsb_memory dev;              /* one reference on the device block (root and shares alike) */
  nnsb_memory host;             /* pinned staging, allocated lazily by mem_map (data == NULL: none) */
  gboolean host_valid;          /* host copy reflects the device data */
  gboolean host_dirty;          /* mapped for WRITE: upload on unmap */
  GMutex lock;
} GstNnsbCudaMemory;

/* an allocator on GPU `device_id` with its own context and copy stream: memories keep it (and,
 * through their blocks, the pools they came from) alive after the element is gone */
GstAllocator *gst_nnsb_cuda_allocator_new (int device_id);
gboolean gst_is_nnsb_cuda_memory (GstMemory * mem);
/* wrap a device memory produced by an nnsb element; takes over the caller's reference */
GstMemory *gst_nnsb_cuda_memory_wrap (GstAllocator * allocator, const nnsb_memory * dev);
/* for CUDA-aware peers: a BORROWED view (pointer, size, block with its ready event) of the
 * device memory -- no host round trip.  Valid while the GstMemory is referenced. */
gboolean gst_nnsb_cuda_memory_get (GstMemory * mem, nnsb_memory * out);

G_END_DECLS
#endif
