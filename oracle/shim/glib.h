/*
 * oracle/shim/glib.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Minimal stand-in for <glib.h> so that the reference's own arithmetic
 * (/root/reference/gst/nnstreamer/tensor_data.c and the six per-mode functions of
 * gsttensor_transform.c) can be compiled unmodified, by oracle/build_ref.py, into
 * oracle/_ref/libnns_ref.so.  GLib itself is not installed in this image.
 * Only the handful of typedefs/macros those reference lines use are provided.
 */
#ifndef NNS_ORACLE_GLIB_SHIM_H
#define NNS_ORACLE_GLIB_SHIM_H

#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <assert.h>

typedef int gboolean;
typedef int gint;
typedef unsigned int guint;
typedef char gchar;
typedef unsigned char guint8;
typedef long glong;
typedef unsigned long gulong;
typedef size_t gsize;
typedef double gdouble;
typedef float gfloat;
typedef void *gpointer;
typedef const void *gconstpointer;
typedef int64_t gint64;
typedef uint64_t guint64;

#ifndef TRUE
#define TRUE 1
#endif
#ifndef FALSE
#define FALSE 0
#endif

#ifdef __cplusplus
#define G_BEGIN_DECLS extern "C" {
#define G_END_DECLS }
#else
#define G_BEGIN_DECLS
#define G_END_DECLS
#endif

#define g_return_val_if_fail(expr, val) do { if (!(expr)) return (val); } while (0)
#define g_return_if_fail(expr) do { if (!(expr)) return; } while (0)
#define g_assert(expr) assert (expr)
#define g_try_malloc0(n) calloc (1, (n))
#define g_free(p) free (p)

/* logging: the oracle keeps quiet; messages are not part of the result */
#define NNS_SHIM_LOG(...) do { } while (0)
#define g_info NNS_SHIM_LOG
#define g_warning NNS_SHIM_LOG
#define g_critical NNS_SHIM_LOG
#define g_debug NNS_SHIM_LOG
#define g_error NNS_SHIM_LOG

#undef CLAMP
#define CLAMP(x, low, high) (((x) > (high)) ? (high) : (((x) < (low)) ? (low) : (x)))

typedef struct _GSList GSList;
struct _GSList { gpointer data; GSList *next; };
#define g_slist_next(l) ((l) ? (((GSList *) (l))->next) : NULL)

typedef struct _GList GList;
struct _GList { gpointer data; GList *next; GList *prev; };

#endif
