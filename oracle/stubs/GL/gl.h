/* Typedef-only stand-in for <GL/gl.h>: lets the reference's host files compile headless.
 * Test infrastructure only (oracle/_ref build). No GL entry point is ever reached: the
 * reference's Buffer class is exercised with CPU ownership only. */
#ifndef B2RT_ORACLE_STUB_GL_H
#define B2RT_ORACLE_STUB_GL_H
typedef unsigned int   GLenum;
typedef unsigned char  GLboolean;
typedef unsigned int   GLbitfield;
typedef void           GLvoid;
typedef signed char    GLbyte;
typedef short          GLshort;
typedef int            GLint;
typedef unsigned char  GLubyte;
typedef unsigned short GLushort;
typedef unsigned int   GLuint;
typedef int            GLsizei;
typedef float          GLfloat;
typedef float          GLclampf;
typedef double         GLdouble;
typedef double         GLclampd;
#ifdef __cplusplus
extern "C" {
#endif
void glGetIntegerv(GLenum pname, GLint* params);
#ifdef __cplusplus
}
#endif
#endif
