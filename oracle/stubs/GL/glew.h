/* Stand-in for <GL/glew.h> (see gl.h in this directory). Declares the buffer-object entry
 * points the reference's Buffer.cc names; stubs.c defines them as no-ops. */
#ifndef B2RT_ORACLE_STUB_GLEW_H
#define B2RT_ORACLE_STUB_GLEW_H
#include <stddef.h>
#include "gl.h"
#ifdef __cplusplus
extern "C" {
#endif
void glGenBuffers(GLsizei n, GLuint* buffers);
void glBindBuffer(GLenum target, GLuint buffer);
void glBufferData(GLenum target, ptrdiff_t size, const void* data, GLenum usage);
void glBufferSubData(GLenum target, ptrdiff_t offset, ptrdiff_t size, const void* data);
void glGetBufferSubData(GLenum target, ptrdiff_t offset, ptrdiff_t size, void* data);
void glGetBufferParameteriv(GLenum target, GLenum pname, GLint* params);
void glDeleteBuffers(GLsizei n, const GLuint* buffers);
#ifdef __cplusplus
}
#endif
#endif
