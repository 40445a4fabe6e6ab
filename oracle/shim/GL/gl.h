/* oracle/shim/GL/gl.h -- TEST INFRASTRUCTURE ONLY; everything is in glew.h. */
