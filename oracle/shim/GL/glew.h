/*
 * oracle/shim/GL/glew.h -- TEST INFRASTRUCTURE ONLY.
 * Declarations-only stand-in for GLEW/GL so the reference's C files compile;
 * the GL twins are never called on the CPU path (link stubs abort()).
 * There is no graphics stack on this image: unavailable on this box.
 */
#ifndef ORACLE_SHIM_GLEW_H
#define ORACLE_SHIM_GLEW_H

typedef unsigned int GLuint;
typedef int GLint;
typedef unsigned int GLenum;
typedef int GLsizei;
typedef float GLfloat;
typedef unsigned char GLboolean;
typedef struct GLEWContextStruct GLEWContext;

#define GL_FALSE 0
#define GL_TRUE 1
#define GL_FLOAT 0x1406
#define GL_UNSIGNED_BYTE 0x1401
#define GL_HALF_FLOAT_ARB 0x140B
#define GL_RGBA 0x1908
#define GL_LUMINANCE 0x1909
#define GL_LUMINANCE8 0x8040
#define GL_RGBA16F_ARB 0x881A
#define GL_TEXTURE0 0x84C0
#define GL_TEXTURE1 0x84C1
#define GL_TEXTURE2 0x84C2
#define GL_TEXTURE_RECTANGLE 0x84F5
#define GL_TEXTURE_RECTANGLE_ARB 0x84F5
#define GL_TRIANGLE_FAN 0x0006
#define GL_UNPACK_ROW_LENGTH 0x0CF2
#define GL_ARRAY_BUFFER 0x8892
#define GL_STATIC_DRAW 0x88E4
#define GL_FRAGMENT_SHADER 0x8B30
#define GL_VERTEX_SHADER 0x8B31
#define GL_FRAMEBUFFER_EXT 0x8D40
#define GL_COLOR_ATTACHMENT0_EXT 0x8CE0
#define GL_COLOR_ATTACHMENT1_EXT 0x8CE1
#define GL_MAX_COLOR_ATTACHMENTS 0x8CDF
#define GL_MAX_DRAW_BUFFERS 0x8824
#define GL_TEXTURE_MIN_FILTER 0x2801
#define GL_TEXTURE_MAG_FILTER 0x2800
#define GL_NEAREST 0x2600
#define GL_LINEAR 0x2601

/* Unprototyped on purpose: any argument list compiles (gnu99). */
int glActiveTexture(); int glBindBuffer(); int glBindFramebufferEXT(); int glBindTexture();
int glBufferData(); int glDeleteBuffers(); int glDeleteFramebuffersEXT(); int glDeleteProgram();
int glDeleteShader(); int glDeleteTextures(); int glDisable(); int glDisableVertexAttribArray();
int glDrawArrays(); int glDrawBuffers(); int glEnable(); int glEnableVertexAttribArray();
int glFramebufferTexture2DEXT(); int glGenBuffers(); int glGenFramebuffersEXT(); int glGenTextures();
int glGetAttribLocation(); int glGetIntegerv(); int glGetTexImage(); int glGetUniformLocation();
int glPixelStorei(); int glTexImage2D(); int glTexParameteri(); int glUniform1f(); int glUniform1i();
int glUniform2f(); int glUniform2i(); int glUniform2iv(); int glUniformMatrix3fv(); int glUseProgram();
int glVertexAttribPointer(); int glViewport();

#endif
