/*
 * canvas_b200/csrc/py/py_process.c -- the CPython extension module `fluggo.media.process` for this library
 * (builds fluggo/media/process.<abi>.so; PyInit_process).
 *
 * What the reference's extension gives other C code is a PROTOCOL, not just Python classes
 * (/root/reference/include/pyframework.h:48-132, src/process/main.c:30-66, 359-417):
 *
 *   - every video source is an instance of the C type `VideoSource` and carries the attribute
 *     `_video_frame_source_funcs`: a PyCapsule (of that name) around its static video_frame_source_funcs vtable.  A C
 *     consumer -- an encoder, the pull queue, the widget, another extension -- calls py_video_take_source( obj, &src )
 *     and from then on pulls frames with video_get_frame_f16( src, ... ) from any thread, without the GIL;
 *   - coded image sources carry `_coded_image_source_funcs` (coded_image_source_funcs), frame functions
 *     `_frame_function_funcs` (FrameFunctionFuncs: get_values( self, count, frames, out[][4] )).
 *
 * This module implements that protocol in C and leaves the graph logic where it already is: the source, filter and
 * workspace classes of canvas_b200/process.py (which flatten a whole filter graph into ONE fused launch of
 * libcanvas_b200.so) derive from the three C base types defined here, so every one of them carries the capsules.
 * The vtables are trampolines: they take the GIL (the way src/process/CodedImageSource.c:120-125 does for sources
 * written in Python), ask the object for the frame -- rendered on the B200 -- and copy its window into the caller's
 * buffer.  In the other direction, pull_video_f16 / pull_video_f32 take ANY object with a valid capsule (a foreign
 * extension's source) through py_video_take_source and pull it with the GIL released, which is how process.py admits
 * foreign sources into its graphs.
 *
 * PyInit_process finally re-exports the public names of canvas_b200.process, so `from fluggo.media import process` sees
 * the reference's surface: SolidColorVideoSource, VideoWorkspace, LerpFunc, AnimationFunc, POINT_*, time_get_frame ...
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdbool.h>
#include <stdlib.h>
#include <string.h>

#include "framework.h"

#define VIDEO_FRAME_SOURCE_FUNCS "_video_frame_source_funcs"
#define CODED_IMAGE_SOURCE_FUNCS "_coded_image_source_funcs"
#define FRAME_FUNCTION_FUNCS "_frame_function_funcs"

/* pyframework.h:69-74 */
typedef void (*framefunc_get_values_func)( PyObject *self, Py_ssize_t count, double *frames, double (*out_values)[4] );
typedef struct { int flags; framefunc_get_values_func get_values; } FrameFunctionFuncs;

/* ------------------------------------------------------------------ helpers */

static PyObject *window_tuple( const box2i *b ) {
    return Py_BuildValue( "(iiii)", b->min.x, b->min.y, b->max.x, b->max.y );
}

static bool parse_window( PyObject *obj, box2i *b ) {
    /* (min_x, min_y, max_x, max_y) or ((min_x, min_y), (max_x, max_y)) -- both box2i spellings of fluggo.media.basetypes */
    if( PyArg_ParseTuple( obj, "iiii", &b->min.x, &b->min.y, &b->max.x, &b->max.y ) ) return true;
    PyErr_Clear();
    if( PyArg_ParseTuple( obj, "(ii)(ii)", &b->min.x, &b->min.y, &b->max.x, &b->max.y ) ) return true;
    PyErr_Clear();
    PyErr_SetString( PyExc_TypeError, "expected a box2i: four integers or two corner pairs" );
    return false;
}

/* Copies the rows of `current` out of a Python buffer that holds `full` (row-major, `px` bytes per pixel). */
static bool copy_window_from_buffer( PyObject *pixels, void *dst, const box2i *full, const box2i *current, size_t px ) {
    Py_buffer view;
    if( PyObject_GetBuffer( pixels, &view, PyBUF_C_CONTIGUOUS ) )
        return false;
    v2i size, wsize;
    box2i_get_size( full, &size );
    box2i_get_size( current, &wsize );
    const size_t need = (size_t) size.x * (size_t) size.y * px;
    bool ok = (size_t) view.len >= need;
    if( ok ) {
        for( int y = current->min.y; y <= current->max.y; y++ ) {
            const size_t offset = ((size_t)( y - full->min.y ) * (size_t) size.x + (size_t)( current->min.x - full->min.x )) * px;
            memcpy( (char *) dst + offset, (const char *) view.buf + offset, (size_t) wsize.x * px );
        }
    }
    else
        PyErr_SetString( PyExc_ValueError, "frame buffer smaller than its full_window" );
    PyBuffer_Release( &view );
    return ok;
}

/* ------------------------------------------------------------------ VideoSource: the vtable is a trampoline */

/* asks the Python object for a frame over frame->full_window: obj._pull_f16/_pull_f32( index, window ) ->
 * ( buffer holding full_window, (current_window) ) -- process.py renders it on the device */
static void pull_through_python( void *self, int frame_index, void *data, box2i *full, box2i *current, const char *method, size_t px ) {
    PyGILState_STATE gil = PyGILState_Ensure();
    box2i_set_empty( current );

    PyObject *window = window_tuple( full );
    PyObject *result = window ? PyObject_CallMethod( (PyObject *) self, method, "iO", frame_index, window ) : NULL;
    Py_XDECREF( window );

    if( result ) {
        PyObject *pixels = NULL, *cw = NULL;
        box2i got;
        if( PyArg_ParseTuple( result, "OO", &pixels, &cw ) && parse_window( cw, &got ) ) {
            box2i clipped;
            box2i_intersect( &clipped, &got, full );
            if( box2i_is_empty( &clipped ) || copy_window_from_buffer( pixels, data, full, &clipped, px ) )
                *current = box2i_is_empty( &clipped ) ? *current : clipped;
        }
        Py_DECREF( result );
    }
    if( PyErr_Occurred() ) {
        /* the reference's convention for callbacks into Python: report, produce nothing (CodedImageSource.c:127-130) */
        PyErr_Print();
        box2i_set_empty( current );
    }
    PyGILState_Release( gil );
}

static void source_get_frame( void *self, int frame_index, rgba_frame_f16 *frame ) {
    pull_through_python( self, frame_index, frame->data, &frame->full_window, &frame->current_window, "_pull_f16", sizeof(rgba_f16) );
}
static void source_get_frame_32( void *self, int frame_index, rgba_frame_f32 *frame ) {
    pull_through_python( self, frame_index, frame->data, &frame->full_window, &frame->current_window, "_pull_f32", sizeof(rgba_f32) );
}

static video_frame_source_funcs source_funcs = { .flags = 0, .get_frame = source_get_frame, .get_frame_32 = source_get_frame_32, .get_frame_gl = NULL };

static PyObject *source_get_funcs( PyObject *self, void *closure ) {
    (void) self; (void) closure;
    return PyCapsule_New( &source_funcs, VIDEO_FRAME_SOURCE_FUNCS, NULL );
}

static PyGetSetDef source_getsetters[] = {
    { VIDEO_FRAME_SOURCE_FUNCS, source_get_funcs, NULL, "Video frame source C API (a PyCapsule around video_frame_source_funcs).", NULL },
    { NULL }
};

PyTypeObject py_type_VideoSource = {
    PyVarObject_HEAD_INIT( NULL, 0 )
    .tp_name = "fluggo.media.process.VideoSource",
    .tp_basicsize = sizeof(PyObject),
    .tp_flags = Py_TPFLAGS_DEFAULT | Py_TPFLAGS_BASETYPE,
    .tp_doc = "Base of every video source: carries the " VIDEO_FRAME_SOURCE_FUNCS " capsule C consumers bind to.",
    .tp_getset = source_getsetters,
    .tp_new = PyType_GenericNew,
};

/* main.c:30-66: binds a C video_source to a Python object through its capsule; called with NULL / None (or again
 * with another object) it releases what *source held.  Exported for other extensions, as in the reference. */
typedef struct { video_source source; PyObject *csource; } py_video_source;

EXPORT bool py_video_take_source( PyObject *obj, video_source **source_ ) {
    py_video_source *held = (py_video_source *) *source_;
    if( held ) {
        Py_CLEAR( held->csource );
        PyObject *old = (PyObject *) held->source.obj;
        held->source.obj = NULL;
        Py_XDECREF( old );
        free( held );
        *source_ = NULL;
    }
    if( !obj || obj == Py_None )
        return true;

    PyObject *capsule = PyObject_GetAttrString( obj, VIDEO_FRAME_SOURCE_FUNCS );
    if( !capsule || !PyCapsule_IsValid( capsule, VIDEO_FRAME_SOURCE_FUNCS ) ) {
        Py_XDECREF( capsule );
        PyErr_SetString( PyExc_Exception, "The source didn't have an acceptable " VIDEO_FRAME_SOURCE_FUNCS " attribute." );
        return false;
    }
    held = (py_video_source *) calloc( 1, sizeof(*held) );
    if( !held ) { Py_DECREF( capsule ); PyErr_NoMemory(); return false; }
    Py_INCREF( obj );
    held->source.obj = obj;
    held->source.funcs = (video_frame_source_funcs *) PyCapsule_GetPointer( capsule, VIDEO_FRAME_SOURCE_FUNCS );
    held->csource = capsule;
    *source_ = &held->source;
    return true;
}

/* pull_video_f16( source, frame_index, window ) -> ( bytes of window-sized rgba_f16 pixels, (current_window) ): ANY
 * object with a valid capsule, pulled through the C interface with the GIL released */
static PyObject *pull_video( PyObject *args, int wide ) {
    PyObject *obj, *window_obj;
    int frame_index;
    box2i full;
    if( !PyArg_ParseTuple( args, "OiO", &obj, &frame_index, &window_obj ) || !parse_window( window_obj, &full ) )
        return NULL;

    video_source *source = NULL;
    if( !py_video_take_source( obj, &source ) )
        return NULL;

    v2i size;
    box2i_get_size( &full, &size );
    const size_t px = wide ? sizeof(rgba_f32) : sizeof(rgba_f16);
    PyObject *pixels = PyBytes_FromStringAndSize( NULL, (Py_ssize_t)( (size_t) size.x * (size_t) size.y * px ) );
    if( !pixels ) { py_video_take_source( NULL, &source ); return NULL; }
    memset( PyBytes_AS_STRING( pixels ), 0, (size_t) PyBytes_GET_SIZE( pixels ) );

    box2i current;
    box2i_set_empty( &current );
    if( source ) {
        if( wide ) {
            rgba_frame_f32 frame = { (rgba_f32 *) PyBytes_AS_STRING( pixels ), full, full };
            Py_BEGIN_ALLOW_THREADS
            video_get_frame_f32( source, frame_index, &frame );
            Py_END_ALLOW_THREADS
            current = frame.current_window;
        }
        else {
            rgba_frame_f16 frame = { (rgba_f16 *) PyBytes_AS_STRING( pixels ), full, full };
            Py_BEGIN_ALLOW_THREADS
            video_get_frame_f16( source, frame_index, &frame );
            Py_END_ALLOW_THREADS
            current = frame.current_window;
        }
    }
    py_video_take_source( NULL, &source );
    if( PyErr_Occurred() ) { Py_DECREF( pixels ); return NULL; }
    return Py_BuildValue( "(NN)", pixels, window_tuple( &current ) );
}

static PyObject *py_pull_video_f16( PyObject *self, PyObject *args ) { (void) self; return pull_video( args, 0 ); }
static PyObject *py_pull_video_f32( PyObject *self, PyObject *args ) { (void) self; return pull_video( args, 1 ); }

/* ------------------------------------------------------------------ CodedImageSource */

static void coded_image_release( void *image ) { free( image ); }

/* obj._pull_coded( frame ) -> None or a sequence of ( buffer, stride, line_count ) per plane; copied into ONE block
 * behind the coded_image (the way coded_image_alloc lays planes out, video_subsample.c:24-73) */
static coded_image *coded_get_frame( void *self, int frame, int quality_hint ) {
    (void) quality_hint;
    PyGILState_STATE gil = PyGILState_Ensure();
    coded_image *image = NULL;
    PyObject *result = PyObject_CallMethod( (PyObject *) self, "_pull_coded", "i", frame );
    PyObject *seq = (result && result != Py_None) ? PySequence_Fast( result, "expected a sequence of planes" ) : NULL;

    if( seq ) {
        const Py_ssize_t planes = PySequence_Fast_GET_SIZE( seq );
        Py_buffer views[CODED_IMAGE_MAX_PLANES];
        int strides[CODED_IMAGE_MAX_PLANES], lines[CODED_IMAGE_MAX_PLANES], taken = 0;
        size_t total = sizeof(coded_image);
        bool ok = planes <= CODED_IMAGE_MAX_PLANES;

        for( Py_ssize_t i = 0; ok && i < planes; i++ ) {
            PyObject *data;
            if( !PyArg_ParseTuple( PySequence_Fast_GET_ITEM( seq, i ), "Oii", &data, &strides[i], &lines[i] ) ||
                    PyObject_GetBuffer( data, &views[i], PyBUF_C_CONTIGUOUS ) ) { ok = false; break; }
            taken++;
            if( strides[i] < 0 || lines[i] < 0 || (size_t) views[i].len < (size_t) strides[i] * (size_t) lines[i] ) ok = false;
            total += (size_t) strides[i] * (size_t) lines[i];
        }
        if( ok && (image = (coded_image *) calloc( 1, total )) ) {
            uint8_t *cursor = (uint8_t *)( image + 1 );
            for( Py_ssize_t i = 0; i < planes; i++ ) {
                const size_t bytes = (size_t) strides[i] * (size_t) lines[i];
                memcpy( cursor, views[i].buf, bytes );
                image->data[i] = cursor; image->stride[i] = strides[i]; image->line_count[i] = lines[i];
                cursor += bytes;
            }
            image->free_func = coded_image_release;
        }
        for( int i = 0; i < taken; i++ ) PyBuffer_Release( &views[i] );
        Py_DECREF( seq );
    }
    Py_XDECREF( result );
    if( PyErr_Occurred() ) PyErr_Print();
    PyGILState_Release( gil );
    return image;
}

static coded_image_source_funcs coded_funcs = { .flags = 0, .getFrame = coded_get_frame };

static PyObject *coded_get_funcs( PyObject *self, void *closure ) {
    (void) self; (void) closure;
    return PyCapsule_New( &coded_funcs, CODED_IMAGE_SOURCE_FUNCS, NULL );
}
static PyGetSetDef coded_getsetters[] = {
    { CODED_IMAGE_SOURCE_FUNCS, coded_get_funcs, NULL, "Coded image source C API (a PyCapsule around coded_image_source_funcs).", NULL },
    { NULL }
};
PyTypeObject py_type_CodedImageSource = {
    PyVarObject_HEAD_INIT( NULL, 0 )
    .tp_name = "fluggo.media.process.CodedImageSource",
    .tp_basicsize = sizeof(PyObject),
    .tp_flags = Py_TPFLAGS_DEFAULT | Py_TPFLAGS_BASETYPE,
    .tp_doc = "Base of the coded image sources: carries the " CODED_IMAGE_SOURCE_FUNCS " capsule.",
    .tp_getset = coded_getsetters,
    .tp_new = PyType_GenericNew,
};

/* pull_coded( source, frame ) -> None or [ ( bytes, stride, line_count ) ... ] through the capsule, GIL released */
static PyObject *py_pull_coded( PyObject *self, PyObject *args ) {
    (void) self;
    PyObject *obj;
    int frame;
    if( !PyArg_ParseTuple( args, "Oi", &obj, &frame ) ) return NULL;
    PyObject *capsule = PyObject_GetAttrString( obj, CODED_IMAGE_SOURCE_FUNCS );
    if( !capsule || !PyCapsule_IsValid( capsule, CODED_IMAGE_SOURCE_FUNCS ) ) {
        Py_XDECREF( capsule );
        PyErr_SetString( PyExc_Exception, "The source didn't have an acceptable " CODED_IMAGE_SOURCE_FUNCS " attribute." );
        return NULL;
    }
    coded_image_source_funcs *funcs = (coded_image_source_funcs *) PyCapsule_GetPointer( capsule, CODED_IMAGE_SOURCE_FUNCS );
    coded_image *image = NULL;
    Py_BEGIN_ALLOW_THREADS
    image = funcs->getFrame ? funcs->getFrame( obj, frame, 0 ) : NULL;
    Py_END_ALLOW_THREADS
    Py_DECREF( capsule );
    if( !image ) Py_RETURN_NONE;

    PyObject *list = PyList_New( 0 );
    for( int i = 0; list && i < CODED_IMAGE_MAX_PLANES && image->data[i]; i++ ) {
        PyObject *plane = Py_BuildValue( "(y#ii)", (const char *) image->data[i], (Py_ssize_t)( (size_t) image->stride[i] * (size_t) image->line_count[i] ),
                                         image->stride[i], image->line_count[i] );
        if( !plane || PyList_Append( list, plane ) ) { Py_XDECREF( plane ); Py_CLEAR( list ); break; }
        Py_DECREF( plane );
    }
    if( image->free_func ) image->free_func( image );
    return list;
}

/* ------------------------------------------------------------------ FrameFunction */

/* FrameFunctionFuncs.get_values: obj.get_values( [frames] ) -> [ (a, b, c, d) ... ]; shorter tuples are zero-padded */
static void framefunc_get_values( PyObject *self, Py_ssize_t count, double *frames, double (*out)[4] ) {
    PyGILState_STATE gil = PyGILState_Ensure();
    for( Py_ssize_t i = 0; i < count; i++ ) out[i][0] = out[i][1] = out[i][2] = out[i][3] = 0.0;

    PyObject *list = PyList_New( count );
    for( Py_ssize_t i = 0; list && i < count; i++ )
        PyList_SET_ITEM( list, i, PyFloat_FromDouble( frames[i] ) );
    PyObject *result = list ? PyObject_CallMethod( self, "get_values", "O", list ) : NULL;
    PyObject *seq = result ? PySequence_Fast( result, "get_values must return a sequence" ) : NULL;
    if( seq ) {
        const Py_ssize_t n = PySequence_Fast_GET_SIZE( seq );
        for( Py_ssize_t i = 0; i < count && i < n; i++ ) {
            PyObject *values = PySequence_Fast( PySequence_Fast_GET_ITEM( seq, i ), "expected a tuple of values" );
            if( !values ) break;
            const Py_ssize_t m = PySequence_Fast_GET_SIZE( values );
            for( Py_ssize_t j = 0; j < 4 && j < m; j++ )
                out[i][j] = PyFloat_AsDouble( PySequence_Fast_GET_ITEM( values, j ) );
            Py_DECREF( values );
        }
        Py_DECREF( seq );
    }
    Py_XDECREF( result );
    Py_XDECREF( list );
    if( PyErr_Occurred() ) PyErr_Print();
    PyGILState_Release( gil );
}

static FrameFunctionFuncs framefunc_funcs = { .flags = 0, .get_values = framefunc_get_values };

static PyObject *framefunc_get_funcs( PyObject *self, void *closure ) {
    (void) self; (void) closure;
    return PyCapsule_New( &framefunc_funcs, FRAME_FUNCTION_FUNCS, NULL );
}
static PyGetSetDef framefunc_getsetters[] = {
    { FRAME_FUNCTION_FUNCS, framefunc_get_funcs, NULL, "Frame function C API (a PyCapsule around FrameFunctionFuncs).", NULL },
    { NULL }
};
PyTypeObject py_type_FrameFunction = {
    PyVarObject_HEAD_INIT( NULL, 0 )
    .tp_name = "fluggo.media.process.FrameFunction",
    .tp_basicsize = sizeof(PyObject),
    .tp_flags = Py_TPFLAGS_DEFAULT | Py_TPFLAGS_BASETYPE,
    .tp_doc = "Base of the frame functions: carries the " FRAME_FUNCTION_FUNCS " capsule.",
    .tp_getset = framefunc_getsetters,
    .tp_new = PyType_GenericNew,
};

/* framefunc_values( func, [frames] ) -> [ (a, b, c, d) ... ] through the capsule (what py_framefunc_take_source users do) */
static PyObject *py_framefunc_values( PyObject *self, PyObject *args ) {
    (void) self;
    PyObject *obj, *frames_obj;
    if( !PyArg_ParseTuple( args, "OO", &obj, &frames_obj ) ) return NULL;
    PyObject *capsule = PyObject_GetAttrString( obj, FRAME_FUNCTION_FUNCS );
    if( !capsule || !PyCapsule_IsValid( capsule, FRAME_FUNCTION_FUNCS ) ) {
        Py_XDECREF( capsule );
        PyErr_SetString( PyExc_Exception, "The source didn't have an acceptable " FRAME_FUNCTION_FUNCS " attribute." );
        return NULL;
    }
    FrameFunctionFuncs *funcs = (FrameFunctionFuncs *) PyCapsule_GetPointer( capsule, FRAME_FUNCTION_FUNCS );
    PyObject *seq = PySequence_Fast( frames_obj, "expected a sequence of frames" );
    if( !seq ) { Py_DECREF( capsule ); return NULL; }
    const Py_ssize_t count = PySequence_Fast_GET_SIZE( seq );
    double *frames = (double *) calloc( (size_t)( count ? count : 1 ), sizeof(double) );
    double (*out)[4] = (double (*)[4]) calloc( (size_t)( count ? count : 1 ), sizeof(double[4]) );
    for( Py_ssize_t i = 0; i < count; i++ ) frames[i] = PyFloat_AsDouble( PySequence_Fast_GET_ITEM( seq, i ) );
    Py_DECREF( seq );
    PyObject *list = NULL;
    if( !PyErr_Occurred() ) {
        Py_BEGIN_ALLOW_THREADS
        funcs->get_values( obj, count, frames, out );
        Py_END_ALLOW_THREADS
        list = PyList_New( count );
        for( Py_ssize_t i = 0; list && i < count; i++ )
            PyList_SET_ITEM( list, i, Py_BuildValue( "(dddd)", out[i][0], out[i][1], out[i][2], out[i][3] ) );
    }
    free( frames ); free( out );
    Py_DECREF( capsule );
    return list;
}

/* ------------------------------------------------------------------ module */

static PyMethodDef module_methods[] = {
    { "pull_video_f16", py_pull_video_f16, METH_VARARGS, "pull_video_f16(source, frame_index, window) -> (pixels, current_window): any object with a " VIDEO_FRAME_SOURCE_FUNCS " capsule, through the C vtable." },
    { "pull_video_f32", py_pull_video_f32, METH_VARARGS, "The same as f32 pixels." },
    { "pull_coded", py_pull_coded, METH_VARARGS, "pull_coded(source, frame) -> planes of a coded image through the " CODED_IMAGE_SOURCE_FUNCS " capsule." },
    { "framefunc_values", py_framefunc_values, METH_VARARGS, "framefunc_values(func, frames) through the " FRAME_FUNCTION_FUNCS " capsule." },
    { NULL, NULL, 0, NULL }
};

EXPORT PyMODINIT_FUNC PyInit_process( void ) {
    static PyModuleDef mdef = {
        .m_base = PyModuleDef_HEAD_INIT,
        .m_name = "process",
        .m_doc = "fluggo.media.process on the B200 path (canvas-b200): the reference's source/filter types, frames rendered by libcanvas_b200.so.",
        .m_size = -1,
        .m_methods = module_methods,
    };

    if( PyType_Ready( &py_type_VideoSource ) < 0 || PyType_Ready( &py_type_CodedImageSource ) < 0 || PyType_Ready( &py_type_FrameFunction ) < 0 )
        return NULL;
    PyObject *m = PyModule_Create( &mdef );
    if( !m ) return NULL;

    Py_INCREF( &py_type_VideoSource );
    Py_INCREF( &py_type_CodedImageSource );
    Py_INCREF( &py_type_FrameFunction );
    if( PyModule_AddObject( m, "VideoSource", (PyObject *) &py_type_VideoSource ) < 0 ||
        PyModule_AddObject( m, "CodedImageSource", (PyObject *) &py_type_CodedImageSource ) < 0 ||
        PyModule_AddObject( m, "FrameFunction", (PyObject *) &py_type_FrameFunction ) < 0 ||
        PyModule_AddObject( m, "NS_PER_SEC", PyLong_FromLongLong( NS_PER_SEC ) ) < 0 )
        return NULL;

    /* the C base types are registered with the implementation before it defines its classes on top of them, then
     * its public names become this module's (the reference's init_* calls, main.c:385-410, in one step) */
    PyObject *bases = PyImport_ImportModule( "canvas_b200._cbases" );
    if( !bases ) { Py_DECREF( m ); return NULL; }
    if( PyObject_SetAttrString( bases, "loading", Py_True ) < 0 ||
        PyObject_SetAttrString( bases, "VideoSource", (PyObject *) &py_type_VideoSource ) < 0 ||
        PyObject_SetAttrString( bases, "CodedImageSource", (PyObject *) &py_type_CodedImageSource ) < 0 ||
        PyObject_SetAttrString( bases, "FrameFunction", (PyObject *) &py_type_FrameFunction ) < 0 ||
        PyObject_SetAttrString( bases, "cmodule", m ) < 0 ) { Py_DECREF( bases ); Py_DECREF( m ); return NULL; }

    /* canvas_b200.process defines its classes on top of these types and, as its last statement, makes its public names
     * this module's (the reference's init_* calls, main.c:385-410, in one step).  When IT is what triggered this import,
     * it is half way through its own body here: the import below is then a no-op and the names arrive when it ends. */
    PyObject *impl = PyImport_ImportModule( "canvas_b200.process" );
    PyObject_SetAttrString( bases, "loading", Py_False );
    Py_DECREF( bases );
    if( !impl ) { Py_DECREF( m ); return NULL; }
    Py_DECREF( impl );
    return m;
}
