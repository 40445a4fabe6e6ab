/*
 * oracle/shim/glib.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * A minimal stand-in for <glib.h> so that the reference's own C files for the
 * per-pixel path (src/cprocess/{half,gammatab,filter,video_mix,video_reconstruct,
 * video_subsample,main,workspace}.c, compiled where they lie under /root/reference)
 * build on an image that has no glib.  Only what those files use is provided:
 * basic typedefs, attribute macros, malloc-backed g_malloc/g_slice, g_once_init,
 * a pthread GMutex and a small sorted doubly-linked-list GSequence.
 *
 * GSequence tie rule: g_sequence_insert_sorted places a new element AFTER all
 * elements that compare equal (same as GLib's g_sequence_search).
 */
#ifndef ORACLE_SHIM_GLIB_H
#define ORACLE_SHIM_GLIB_H

#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <stdint.h>
#include <stddef.h>
#include <limits.h>
#include <assert.h>
#include <pthread.h>

typedef size_t gsize;
typedef int gint;
typedef unsigned int guint;
typedef int gboolean;
typedef void *gpointer;
typedef const void *gconstpointer;
typedef char gchar;
typedef uint32_t GQuark;
typedef void (*GFreeFunc)( gpointer data );
typedef void (*GDestroyNotify)( gpointer data );
typedef gint (*GCompareDataFunc)( gconstpointer a, gconstpointer b, gpointer user_data );

#ifndef TRUE
#define TRUE 1
#define FALSE 0
#endif

#define G_GNUC_CONST    __attribute__((const))
#define G_GNUC_PURE     __attribute__((pure))
#define G_GNUC_MALLOC   __attribute__((malloc))
#define G_GNUC_UNUSED   __attribute__((unused))
#define G_LIKELY(x)     __builtin_expect(!!(x), 1)
#define G_UNLIKELY(x)   __builtin_expect(!!(x), 0)
#define G_STRINGIFY(x)      G_STRINGIFY_ARG(x)
#define G_STRINGIFY_ARG(x)  #x
#define G_PI        3.1415926535897932384626433832795028841971693993751
#define G_MAXINT    INT_MAX
#define G_MININT    INT_MIN
#define G_LOG_DOMAIN "oracle"

#define g_assert(x)     assert(x)
#define g_debug(...)    ((void) 0)
#define g_warning(...)  do { fprintf( stderr, "warning: " __VA_ARGS__ ); fputc( '\n', stderr ); } while( 0 )
#define g_print(...)    printf( __VA_ARGS__ )

static inline gpointer g_malloc( gsize n ) { return n ? malloc( n ) : NULL; }
static inline gpointer g_malloc0( gsize n ) { return n ? calloc( 1, n ) : NULL; }
static inline void g_free( gpointer p ) { free( p ); }
#define g_new0(type, n)         ((type *) g_malloc0( sizeof(type) * (n) ))
#define g_new(type, n)          ((type *) g_malloc( sizeof(type) * (n) ))
#define g_slice_alloc(n)        g_malloc( n )
#define g_slice_alloc0(n)       g_malloc0( n )
#define g_slice_free1(n, p)     g_free( p )
#define g_slice_new(type)       ((type *) g_malloc( sizeof(type) ))
#define g_slice_new0(type)      ((type *) g_malloc0( sizeof(type) ))
#define g_slice_free(type, p)   g_free( p )

/* g_once_init: the reference only uses it for lazily-built tables. */
static inline gboolean g_once_init_enter( volatile gsize *loc ) {
    extern pthread_mutex_t oracle_shim_once_mutex;
    if( __atomic_load_n( loc, __ATOMIC_ACQUIRE ) )
        return FALSE;
    pthread_mutex_lock( &oracle_shim_once_mutex );
    if( __atomic_load_n( loc, __ATOMIC_ACQUIRE ) ) {
        pthread_mutex_unlock( &oracle_shim_once_mutex );
        return FALSE;
    }
    return TRUE;    /* mutex stays held until g_once_init_leave */
}
static inline void g_once_init_leave( volatile gsize *loc, gsize value ) {
    extern pthread_mutex_t oracle_shim_once_mutex;
    __atomic_store_n( loc, value, __ATOMIC_RELEASE );
    pthread_mutex_unlock( &oracle_shim_once_mutex );
}

/* GMutex */
typedef struct { pthread_mutex_t m; } GMutex;
static inline void g_mutex_init( GMutex *m ) { pthread_mutex_init( &m->m, NULL ); }
static inline void g_mutex_clear( GMutex *m ) { pthread_mutex_destroy( &m->m ); }
static inline void g_mutex_lock( GMutex *m ) { pthread_mutex_lock( &m->m ); }
static inline void g_mutex_unlock( GMutex *m ) { pthread_mutex_unlock( &m->m ); }

/* Dataset / quark: only the GL halves use them; never reached on the CPU path. */
GQuark g_quark_from_static_string( const gchar *s );
gpointer g_dataset_id_get_data( gconstpointer loc, GQuark key );
void g_dataset_id_set_data_full( gconstpointer loc, GQuark key, gpointer data, GDestroyNotify destroy );

/* GSequence: sorted doubly-linked list with an end sentinel. */
typedef struct _GSequence GSequence;
typedef struct _GSequenceIter GSequenceIter;

GSequence *g_sequence_new( GDestroyNotify data_destroy );
void g_sequence_free( GSequence *seq );
gint g_sequence_get_length( GSequence *seq );
GSequenceIter *g_sequence_get_begin_iter( GSequence *seq );
GSequenceIter *g_sequence_get_end_iter( GSequence *seq );
GSequenceIter *g_sequence_get_iter_at_pos( GSequence *seq, gint pos );
GSequenceIter *g_sequence_insert_sorted( GSequence *seq, gpointer data, GCompareDataFunc cmp, gpointer user );
void g_sequence_sort_changed( GSequenceIter *iter, GCompareDataFunc cmp, gpointer user );
void g_sequence_remove( GSequenceIter *iter );
gpointer g_sequence_get( GSequenceIter *iter );
gboolean g_sequence_iter_is_begin( GSequenceIter *iter );
gboolean g_sequence_iter_is_end( GSequenceIter *iter );
GSequenceIter *g_sequence_iter_next( GSequenceIter *iter );
GSequenceIter *g_sequence_iter_prev( GSequenceIter *iter );
gint g_sequence_iter_compare( GSequenceIter *a, GSequenceIter *b );

#endif
