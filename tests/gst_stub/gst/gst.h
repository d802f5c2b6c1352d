/*
 * tests/gst_stub/gst/gst.h -- DECLARATION-ONLY stand-in for the GLib / GObject / GStreamer 1.x
 * names the glue under nnstreamer_b200/gst/ uses.  Test infrastructure, never shipped, never
 * linked: it exists so that `make -C nnstreamer_b200/gst syntax` can put the glue through
 * `gcc -fsyntax-only -Wall` in an image without GStreamer (typos, wrong argument counts and
 * types, missing declarations).  Prototypes follow the GStreamer 1.x documentation; where
 * pkg-config finds the real headers the Makefile builds against those instead.
 */
#ifndef NNSB_GST_STUB_H
#define NNSB_GST_STUB_H
#include <stddef.h>
#include <stdint.h>

typedef int gboolean;
typedef int gint;
typedef unsigned int guint;
typedef char gchar;
typedef size_t gsize;
typedef ptrdiff_t gssize;
typedef void *gpointer;
typedef const void *gconstpointer;
typedef unsigned long GType;
typedef uint64_t guint64;
typedef int64_t gint64;
typedef uint64_t GstClockTime;
typedef void (*GDestroyNotify) (gpointer data);
#define TRUE 1
#define FALSE 0
#define G_BEGIN_DECLS
#define G_END_DECLS
#define G_GSIZE_FORMAT "zu"
#ifndef NULL
#define NULL ((void *) 0)
#endif

/* ---- GLib ---- */
typedef struct { gpointer p; } GMutex;
void g_mutex_init (GMutex * m);
void g_mutex_clear (GMutex * m);
void g_mutex_lock (GMutex * m);
void g_mutex_unlock (GMutex * m);
gpointer g_malloc0 (gsize n);
gpointer g_malloc (gsize n);
void g_free (gpointer p);
#define g_new0(T, n) ((T *) g_malloc0 (sizeof (T) * (n)))
#define g_new(T, n) ((T *) g_malloc (sizeof (T) * (n)))
int g_strcmp0 (const char *a, const char *b);
gint g_snprintf (gchar * buf, unsigned long n, const gchar * fmt, ...) __attribute__((format (printf, 3, 4)));
void g_critical (const gchar * fmt, ...) __attribute__((format (printf, 1, 2)));

/* ---- GObject ---- */
typedef struct _GObject { gpointer g_class; } GObject;
typedef struct _GParamSpec GParamSpec;
typedef struct _GValue GValue;
typedef struct _GObjectClass
{
  void (*set_property) (GObject * object, guint prop_id, const GValue * value, GParamSpec * pspec);
  void (*get_property) (GObject * object, guint prop_id, GValue * value, GParamSpec * pspec);
  void (*finalize) (GObject * object);
} GObjectClass;
typedef struct { gint value; const gchar *value_name; const gchar *value_nick; } GEnumValue;
typedef enum { G_PARAM_READABLE = 1, G_PARAM_WRITABLE = 2, G_PARAM_READWRITE = 3 } GParamFlags;
GType g_enum_register_static (const gchar * name, const GEnumValue * values);
gpointer g_object_new (GType type, const gchar * first_property, ...);
void g_object_class_install_property (GObjectClass * klass, guint id, GParamSpec * pspec);
GParamSpec *g_param_spec_boolean (const gchar * name, const gchar * nick, const gchar * blurb, gboolean def, GParamFlags flags);
GParamSpec *g_param_spec_string (const gchar * name, const gchar * nick, const gchar * blurb, const gchar * def, GParamFlags flags);
GParamSpec *g_param_spec_uint (const gchar * name, const gchar * nick, const gchar * blurb, guint min, guint max, guint def, GParamFlags flags);
GParamSpec *g_param_spec_int (const gchar * name, const gchar * nick, const gchar * blurb, gint min, gint max, gint def, GParamFlags flags);
GParamSpec *g_param_spec_enum (const gchar * name, const gchar * nick, const gchar * blurb, GType type, gint def, GParamFlags flags);
gboolean g_value_get_boolean (const GValue * v);
void g_value_set_boolean (GValue * v, gboolean b);
const gchar *g_value_get_string (const GValue * v);
void g_value_set_string (GValue * v, const gchar * s);
gint g_value_get_enum (const GValue * v);
void g_value_set_enum (GValue * v, gint e);
guint g_value_get_uint (const GValue * v);
void g_value_set_uint (GValue * v, guint u);
gint g_value_get_int (const GValue * v);
void g_value_set_int (GValue * v, gint i);
void nnsb_stub_warn_invalid_property (GObject * object, guint id, GParamSpec * pspec);
#define G_OBJECT_WARN_INVALID_PROPERTY_ID(o, id, ps) nnsb_stub_warn_invalid_property ((GObject *) (o), (id), (ps))
#define G_OBJECT_CLASS(k) ((GObjectClass *) (k))
#define G_DEFINE_TYPE(TN, t_n, T_P) \
  static void t_n##_init (TN * self); \
  static void t_n##_class_init (TN##Class * klass); \
  static gpointer t_n##_parent_class = NULL; \
  GType t_n##_get_type (void); \
  static inline void t_n##_stub_use (void) { (void) t_n##_init; (void) t_n##_class_init; (void) t_n##_parent_class; }

/* ---- GstObject / GstMiniObject ---- */
typedef struct _GstObject { GObject object; guint flags; } GstObject;
typedef struct _GstMiniObject { GType type; gint refcount; } GstMiniObject;
#define GST_OBJECT_FLAG_SET(obj, flag) (((GstObject *) (obj))->flags |= (flag))
gpointer gst_object_ref (gpointer object);
void gst_object_unref (gpointer object);

/* ---- GstMemory / GstAllocator ---- */
typedef struct _GstMemory GstMemory;
typedef struct _GstAllocator GstAllocator;
typedef struct _GstAllocationParams GstAllocationParams;
typedef enum { GST_MAP_READ = 1, GST_MAP_WRITE = 2 } GstMapFlags;
typedef guint GstMemoryFlags;
typedef struct { GstMemory *memory; GstMapFlags flags; unsigned char *data; gsize size; gsize maxsize; gpointer user_data[4]; } GstMapInfo;
struct _GstMemory
{
  GstMiniObject mini_object;
  GstAllocator *allocator;
  GstMemory *parent;
  gsize maxsize, align, offset, size;
};
typedef gpointer (*GstMemoryMapFunction) (GstMemory * mem, gsize maxsize, GstMapFlags flags);
typedef void (*GstMemoryUnmapFunction) (GstMemory * mem);
typedef GstMemory *(*GstMemoryShareFunction) (GstMemory * mem, gssize offset, gssize size);
struct _GstAllocator
{
  GstObject object;
  const gchar *mem_type;
  GstMemoryMapFunction mem_map;
  GstMemoryUnmapFunction mem_unmap;
  GstMemoryShareFunction mem_share;
};
typedef struct _GstAllocatorClass
{
  GObjectClass object_class;
  GstMemory *(*alloc) (GstAllocator * allocator, gsize size, GstAllocationParams * params);
  void (*free) (GstAllocator * allocator, GstMemory * memory);
} GstAllocatorClass;
#define GST_TYPE_ALLOCATOR ((GType) 1)
#define GST_ALLOCATOR_FLAG_CUSTOM_ALLOC (1u << 4)
#define GST_ALLOCATOR_CLASS(k) ((GstAllocatorClass *) (k))
#define GST_ALLOCATOR_CAST(o) ((GstAllocator *) (o))
#define GST_MEMORY_CAST(m) ((GstMemory *) (m))
void gst_memory_init (GstMemory * mem, GstMemoryFlags flags, GstAllocator * allocator, GstMemory * parent,
    gsize maxsize, gsize align, gsize offset, gsize size);
GstMemory *gst_memory_ref (GstMemory * mem);
void gst_memory_unref (GstMemory * mem);
gboolean gst_memory_map (GstMemory * mem, GstMapInfo * info, GstMapFlags flags);
void gst_memory_unmap (GstMemory * mem, GstMapInfo * info);
GstMemory *gst_memory_share (GstMemory * mem, gssize offset, gssize size);
GstMemory *gst_memory_new_wrapped (GstMemoryFlags flags, gpointer data, gsize maxsize, gsize offset, gsize size,
    gpointer user_data, GDestroyNotify notify);

/* ---- GstBuffer / GstCaps / GstEvent ---- */
typedef struct _GstBuffer GstBuffer;
typedef struct _GstCaps GstCaps;
typedef struct _GstEvent GstEvent;
typedef enum { GST_BUFFER_COPY_METADATA = 7 } GstBufferCopyFlags;
guint gst_buffer_n_memory (GstBuffer * buffer);
GstMemory *gst_buffer_peek_memory (GstBuffer * buffer, guint idx);
void gst_buffer_append_memory (GstBuffer * buffer, GstMemory * mem);
GstBuffer *gst_buffer_new (void);
void gst_buffer_unref (GstBuffer * buffer);
gboolean gst_buffer_copy_into (GstBuffer * dest, GstBuffer * src, GstBufferCopyFlags flags, gsize offset, gsize size);
gchar *gst_caps_to_string (const GstCaps * caps);
GstCaps *gst_caps_from_string (const gchar * s);
GstCaps *gst_caps_new_empty (void);
guint gst_caps_get_size (const GstCaps * caps);
void gst_caps_unref (GstCaps * caps);
GstCaps *gst_caps_make_writable (GstCaps * caps);
GstCaps *gst_caps_fixate (GstCaps * caps);
typedef enum { GST_EVENT_CAPS = 1, GST_EVENT_EOS = 2 } GstEventType;
GstEventType nnsb_stub_event_type (GstEvent * event);
#define GST_EVENT_TYPE(e) nnsb_stub_event_type (e)
void gst_event_parse_caps (GstEvent * event, GstCaps ** caps);
GstEvent *gst_event_new_caps (GstCaps * caps);
void gst_event_unref (GstEvent * event);

/* ---- pads / elements / plugins ---- */
typedef enum { GST_PAD_UNKNOWN, GST_PAD_SRC, GST_PAD_SINK } GstPadDirection;
typedef enum { GST_PAD_ALWAYS } GstPadPresence;
typedef enum { GST_FLOW_OK = 0, GST_FLOW_ERROR = -5, GST_FLOW_NOT_SUPPORTED = -6 } GstFlowReturn;
typedef struct _GstPad GstPad;
typedef struct _GstPadTemplate GstPadTemplate;
typedef struct { const gchar *string; } GstStaticCaps;
typedef struct { const gchar *name_template; GstPadDirection direction; GstPadPresence presence; GstStaticCaps static_caps; } GstStaticPadTemplate;
#define GST_STATIC_CAPS(s) { (s) }
#define GST_STATIC_PAD_TEMPLATE(name, dir, pres, caps) { (name), (dir), (pres), caps }
typedef struct _GstElement { GstObject object; } GstElement;
typedef struct _GstElementClass { GObjectClass parent_class; } GstElementClass;
typedef struct _GstPlugin GstPlugin;
#define GST_TYPE_ELEMENT ((GType) 2)
#define GST_ELEMENT(o) ((GstElement *) (o))
#define GST_RANK_NONE 0
typedef GstFlowReturn (*GstPadChainFunction) (GstPad * pad, GstObject * parent, GstBuffer * buffer);
typedef gboolean (*GstPadEventFunction) (GstPad * pad, GstObject * parent, GstEvent * event);
GstPadTemplate *gst_static_pad_template_get (GstStaticPadTemplate * t);
GstPad *gst_pad_new_from_static_template (GstStaticPadTemplate * t, const gchar * name);
void gst_pad_set_chain_function (GstPad * pad, GstPadChainFunction f);
void gst_pad_set_event_function (GstPad * pad, GstPadEventFunction f);
gboolean gst_element_add_pad (GstElement * element, GstPad * pad);
GstFlowReturn gst_pad_push (GstPad * pad, GstBuffer * buffer);
gboolean gst_pad_push_event (GstPad * pad, GstEvent * event);
gboolean gst_pad_event_default (GstPad * pad, GstObject * parent, GstEvent * event);
void gst_element_class_set_details_simple (GstElementClass * klass, const gchar * longname, const gchar * classification,
    const gchar * description, const gchar * author);
void gst_element_class_add_pad_template (GstElementClass * klass, GstPadTemplate * templ);
gboolean gst_element_register (GstPlugin * plugin, const gchar * name, guint rank, GType type);
#define GST_VERSION_MAJOR 1
#define GST_VERSION_MINOR 20
#define GST_PLUGIN_DEFINE(major, minor, name, description, init, version, license, package, origin) \
  gboolean nnsb_stub_plugin_entry_##name (GstPlugin * p); \
  gboolean nnsb_stub_plugin_entry_##name (GstPlugin * p) { return init (p); }

/* ---- debug ---- */
typedef struct _GstDebugCategory GstDebugCategory;
#define GST_DEBUG_CATEGORY_STATIC(cat) static GstDebugCategory *cat = NULL
GstDebugCategory *nnsb_stub_debug_category (const gchar * name);
#define GST_DEBUG_CATEGORY_INIT(cat, name, color, description) ((cat) = nnsb_stub_debug_category (name))
void nnsb_stub_log (gpointer object, const gchar * fmt, ...) __attribute__((format (printf, 2, 3)));
#define GST_ERROR_OBJECT(obj, ...) nnsb_stub_log ((gpointer) (obj), __VA_ARGS__)
#define GST_WARNING_OBJECT(obj, ...) nnsb_stub_log ((gpointer) (obj), __VA_ARGS__)
#define GST_DEBUG_FUNCPTR(f) (f)

#endif
