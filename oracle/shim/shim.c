/*
 * oracle/shim/shim.c -- TEST INFRASTRUCTURE ONLY (part of oracle/_ref/libcanvasref.so).
 *
 * Link-time companions for shim/glib.h and shim/GL/glew.h: the GSequence list,
 * the g_once mutex, and abort() stubs for every GL entry point and for the
 * reference functions that live in its GL-only files (src/cprocess/gl.c),
 * none of which the CPU path ever reaches.
 */
#include <glib.h>
#include <GL/glew.h>

pthread_mutex_t oracle_shim_once_mutex = PTHREAD_MUTEX_INITIALIZER;

/* ---- GSequence ---------------------------------------------------------- */
struct _GSequenceIter {
    GSequence *seq;
    GSequenceIter *prev, *next;
    gpointer data;
};

struct _GSequence {
    GSequenceIter end;      /* sentinel: end.next = first, end.prev = last */
    gint length;
    GDestroyNotify destroy;
};

GSequence *g_sequence_new( GDestroyNotify data_destroy ) {
    GSequence *seq = calloc( 1, sizeof(GSequence) );
    seq->end.seq = seq;
    seq->end.prev = seq->end.next = &seq->end;
    seq->destroy = data_destroy;
    return seq;
}

void g_sequence_free( GSequence *seq ) {
    GSequenceIter *it = seq->end.next;
    while( it != &seq->end ) {
        GSequenceIter *next = it->next;
        if( seq->destroy )
            seq->destroy( it->data );
        free( it );
        it = next;
    }
    free( seq );
}

gint g_sequence_get_length( GSequence *seq ) { return seq->length; }
GSequenceIter *g_sequence_get_begin_iter( GSequence *seq ) { return seq->end.next; }
GSequenceIter *g_sequence_get_end_iter( GSequence *seq ) { return &seq->end; }

GSequenceIter *g_sequence_get_iter_at_pos( GSequence *seq, gint pos ) {
    if( pos < 0 || pos >= seq->length )
        return &seq->end;
    GSequenceIter *it = seq->end.next;
    while( pos-- )
        it = it->next;
    return it;
}

static void link_before( GSequenceIter *node, GSequenceIter *where ) {
    node->prev = where->prev;
    node->next = where;
    where->prev->next = node;
    where->prev = node;
}

static void unlink_node( GSequenceIter *node ) {
    node->prev->next = node->next;
    node->next->prev = node->prev;
}

static GSequenceIter *find_slot( GSequence *seq, gpointer data, GCompareDataFunc cmp, gpointer user, GSequenceIter *skip ) {
    GSequenceIter *it = seq->end.next;
    while( it != &seq->end ) {
        if( it != skip && cmp( it->data, data, user ) > 0 )
            break;
        it = it->next;
    }
    return it;
}

GSequenceIter *g_sequence_insert_sorted( GSequence *seq, gpointer data, GCompareDataFunc cmp, gpointer user ) {
    GSequenceIter *node = calloc( 1, sizeof(GSequenceIter) );
    node->seq = seq;
    node->data = data;
    link_before( node, find_slot( seq, data, cmp, user, NULL ) );
    seq->length++;
    return node;
}

void g_sequence_sort_changed( GSequenceIter *iter, GCompareDataFunc cmp, gpointer user ) {
    GSequence *seq = iter->seq;
    unlink_node( iter );
    link_before( iter, find_slot( seq, iter->data, cmp, user, iter ) );
}

void g_sequence_remove( GSequenceIter *iter ) {
    GSequence *seq = iter->seq;
    unlink_node( iter );
    seq->length--;
    if( seq->destroy )
        seq->destroy( iter->data );
    free( iter );
}

gpointer g_sequence_get( GSequenceIter *iter ) { return iter->data; }
gboolean g_sequence_iter_is_begin( GSequenceIter *iter ) { return iter == iter->seq->end.next; }
gboolean g_sequence_iter_is_end( GSequenceIter *iter ) { return iter == &iter->seq->end; }
GSequenceIter *g_sequence_iter_next( GSequenceIter *iter ) { return (iter == &iter->seq->end) ? iter : iter->next; }
GSequenceIter *g_sequence_iter_prev( GSequenceIter *iter ) { return (iter == iter->seq->end.next) ? iter : iter->prev; }

gint g_sequence_iter_compare( GSequenceIter *a, GSequenceIter *b ) {
    if( a == b )
        return 0;
    for( GSequenceIter *it = a; it != &a->seq->end; it = it->next ) {
        if( it == b )
            return -1;
    }
    return (b == &a->seq->end) ? -1 : 1;
}

/* ---- never-reached stubs ------------------------------------------------ */
#define DEAD(name) int name() { fprintf( stderr, "oracle shim: " #name " reached (no graphics stack on this box)\n" ); abort(); }

GQuark g_quark_from_static_string( const gchar *s ) { (void) s; return 1; }
gpointer g_dataset_id_get_data( gconstpointer loc, GQuark key ) { (void) loc; (void) key; abort(); }
void g_dataset_id_set_data_full( gconstpointer loc, GQuark key, gpointer data, GDestroyNotify destroy ) {
    (void) loc; (void) key; (void) data; (void) destroy; abort();
}

DEAD(glActiveTexture) DEAD(glBindBuffer) DEAD(glBindFramebufferEXT) DEAD(glBindTexture)
DEAD(glBufferData) DEAD(glDeleteBuffers) DEAD(glDeleteFramebuffersEXT) DEAD(glDeleteProgram)
DEAD(glDeleteShader) DEAD(glDeleteTextures) DEAD(glDisable) DEAD(glDisableVertexAttribArray)
DEAD(glDrawArrays) DEAD(glDrawBuffers) DEAD(glEnable) DEAD(glEnableVertexAttribArray)
DEAD(glFramebufferTexture2DEXT) DEAD(glGenBuffers) DEAD(glGenFramebuffersEXT) DEAD(glGenTextures)
DEAD(glGetAttribLocation) DEAD(glGetIntegerv) DEAD(glGetTexImage) DEAD(glGetUniformLocation)
DEAD(glPixelStorei) DEAD(glTexImage2D) DEAD(glTexParameteri) DEAD(glUniform1f) DEAD(glUniform1i)
DEAD(glUniform2f) DEAD(glUniform2i) DEAD(glUniform2iv) DEAD(glUniformMatrix3fv) DEAD(glUseProgram)
DEAD(glVertexAttribPointer) DEAD(glViewport)

/* reference functions defined in src/cprocess/gl.c (GL-only, out of scope) */
DEAD(getCurrentGLContext) DEAD(gl_ensure_context) DEAD(gl_compile_shader) DEAD(gl_link_program)
DEAD(video_get_frame_gl) DEAD(video_make_gl_texture) DEAD(video_create_filter_program)
DEAD(video_render_gl_frame) DEAD(video_render_gl_frame_filter1) DEAD(video_render_gl_frame_filter2)
DEAD(__gl_checkError) DEAD(video_delete_filter_program)
