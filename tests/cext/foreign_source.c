/*
 * tests/cext/foreign_source.c -- TEST INFRASTRUCTURE: a tiny CPython extension that plays "another extension's video
 * source" (compiled by tests/test_pymodule.py).  Its type ForeignGradient is NOT derived from this library's classes; all
 * it offers is the reference's C protocol: the attribute `_video_frame_source_funcs`, a PyCapsule around its own static
 * vtable (/root/reference/include/pyframework.h:52, e.g. src/process/VideoMixFilter.c:90-98,120).  get_frame writes raw
 * half bits computed from the pixel position and the frame index, inside a fixed window.
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include "framework.h"

typedef struct { PyObject_HEAD box2i window; int calls; } ForeignGradient;

static uint16_t gradient_half( int x, int y, int frame, int channel ) {
    /* positive halves in [0, 1): 0x0000 .. 0x3BFF */
    return (uint16_t)( ((unsigned)( x * 131 + y * 71 + frame * 17 + channel * 977 )) % 0x3C00u );
}

static void gradient_get_frame( void *self, int frame_index, rgba_frame_f16 *frame ) {
    ForeignGradient *g = (ForeignGradient *) self;
    __sync_fetch_and_add( &g->calls, 1 );
    box2i_intersect( &frame->current_window, &g->window, &frame->full_window );
    if( box2i_is_empty( &frame->current_window ) ) { box2i_set_empty( &frame->current_window ); return; }
    for( int y = frame->current_window.min.y; y <= frame->current_window.max.y; y++ )
        for( int x = frame->current_window.min.x; x <= frame->current_window.max.x; x++ ) {
            rgba_f16 *px = video_get_pixel_f16( frame, x, y );
            px->r = gradient_half( x, y, frame_index, 0 ); px->g = gradient_half( x, y, frame_index, 1 );
            px->b = gradient_half( x, y, frame_index, 2 ); px->a = (uint16_t)( 0x3800 + (( x + y ) & 1) * 0x0400 );   /* 0.5 or 1.0 */
        }
}

static video_frame_source_funcs gradient_funcs = { .flags = 0, .get_frame = gradient_get_frame, .get_frame_32 = NULL, .get_frame_gl = NULL };

static int gradient_init( ForeignGradient *self, PyObject *args, PyObject *kw ) {
    (void) kw;
    self->calls = 0;
    return PyArg_ParseTuple( args, "iiii", &self->window.min.x, &self->window.min.y, &self->window.max.x, &self->window.max.y ) ? 0 : -1;
}
static PyObject *gradient_get_funcs( PyObject *self, void *closure ) {
    (void) self; (void) closure;
    return PyCapsule_New( &gradient_funcs, "_video_frame_source_funcs", NULL );
}
static PyObject *gradient_get_calls( PyObject *self, void *closure ) { (void) closure; return PyLong_FromLong( ((ForeignGradient *) self)->calls ); }
static PyGetSetDef gradient_getset[] = {
    { "_video_frame_source_funcs", gradient_get_funcs, NULL, NULL, NULL },
    { "calls", gradient_get_calls, NULL, NULL, NULL },
    { NULL }
};
static PyTypeObject gradient_type = {
    PyVarObject_HEAD_INIT( NULL, 0 )
    .tp_name = "foreign_source.ForeignGradient", .tp_basicsize = sizeof(ForeignGradient), .tp_flags = Py_TPFLAGS_DEFAULT,
    .tp_new = PyType_GenericNew, .tp_init = (initproc) gradient_init, .tp_getset = gradient_getset,
};

PyMODINIT_FUNC PyInit_foreign_source( void ) {
    static PyModuleDef mdef = { .m_base = PyModuleDef_HEAD_INIT, .m_name = "foreign_source", .m_size = -1 };
    if( PyType_Ready( &gradient_type ) < 0 ) return NULL;
    PyObject *m = PyModule_Create( &mdef );
    Py_INCREF( &gradient_type );
    PyModule_AddObject( m, "ForeignGradient", (PyObject *) &gradient_type );
    return m;
}
