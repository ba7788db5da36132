/*
 * TEST INFRASTRUCTURE ONLY (oracle/): link shim for building the reference's own
 * CPU ray tracer headless.
 *
 * The reference's flyscene.cpp / BoundingBox.cpp pull in Tucano, whose Mesh/Shader
 * constructors reference GLEW function-pointer variables and a handful of GL 1.1
 * entry points.  None of them matter for Flyscene::traceRay (pure CPU arithmetic),
 * but the objects do not link without the symbols.  This file defines every such
 * symbol as a no-op.  It is NOT a graphics API or a software renderer: there is no
 * context, nothing is drawn, nothing is dlopen'ed.
 *
 * Symbol list = `nm -u` of the two reference objects.
 */
#include <stddef.h>

typedef unsigned int u32;

static long shim_noop(void) { return 0; }

/* glGetShaderiv / glGetProgramiv: (id, pname, int* out).  Tucano's shader.hpp
 * allocates `new char[*out]` for the info log, so the value must be initialised. */
static void shim_get_iv(u32 id, u32 pname, int *out) {
  (void)id; (void)pname;
  if (out) *out = 1;
}
/* glGen*(n, ids): give deterministic non-zero names. */
static void shim_gen(int n, u32 *ids) {
  static u32 next = 1;
  for (int i = 0; i < n; ++i) ids[i] = next++;
}
static void shim_info_log(u32 id, int maxlen, int *len, char *log) {
  (void)id;
  if (len) *len = 0;
  if (log && maxlen > 0) log[0] = 0;
}

#define NOOP_PTR(name) void *name = (void *)shim_noop;

NOOP_PTR(__glewActiveTexture)
NOOP_PTR(__glewAttachShader)
NOOP_PTR(__glewBindBuffer)
NOOP_PTR(__glewBindVertexArray)
NOOP_PTR(__glewBufferData)
NOOP_PTR(__glewCompileShader)
NOOP_PTR(__glewCreateProgram)
NOOP_PTR(__glewCreateShader)
NOOP_PTR(__glewDeleteBuffers)
NOOP_PTR(__glewDeleteProgram)
NOOP_PTR(__glewDeleteShader)
NOOP_PTR(__glewDeleteVertexArrays)
NOOP_PTR(__glewDetachShader)
NOOP_PTR(__glewDisableVertexAttribArray)
NOOP_PTR(__glewEnableVertexAttribArray)
void *__glewGenBuffers = (void *)shim_gen;
void *__glewGenVertexArrays = (void *)shim_gen;
NOOP_PTR(__glewGetAttribLocation)
void *__glewGetProgramInfoLog = (void *)shim_info_log;
void *__glewGetProgramiv = (void *)shim_get_iv;
void *__glewGetShaderInfoLog = (void *)shim_info_log;
void *__glewGetShaderiv = (void *)shim_get_iv;
NOOP_PTR(__glewGetUniformLocation)
NOOP_PTR(__glewLinkProgram)
NOOP_PTR(__glewPatchParameteri)
NOOP_PTR(__glewShaderSource)
NOOP_PTR(__glewTexImage2DMultisample)
NOOP_PTR(__glewTexImage3D)
NOOP_PTR(__glewUniform1f)
NOOP_PTR(__glewUniform1i)
NOOP_PTR(__glewUniform3f)
NOOP_PTR(__glewUniform4f)
NOOP_PTR(__glewUniformMatrix4fv)
NOOP_PTR(__glewUseProgram)
NOOP_PTR(__glewVertexAttribPointer)

void glBindTexture(u32 a, u32 b) { (void)a; (void)b; }
void glClear(u32 a) { (void)a; }
void glClearColor(float r, float g, float b, float a) { (void)r; (void)g; (void)b; (void)a; }
void glDeleteTextures(int n, const u32 *t) { (void)n; (void)t; }
void glDrawArrays(u32 m, int f, int c) { (void)m; (void)f; (void)c; }
void glDrawElements(u32 m, int c, u32 t, const void *i) { (void)m; (void)c; (void)t; (void)i; }
void glEnable(u32 c) { (void)c; }
void glGenTextures(int n, u32 *t) { shim_gen(n, t); }
u32 glGetError(void) { return 0; }
void glGetIntegerv(u32 p, int *d) { (void)p; if (d) *d = 0; }
void glTexImage1D(u32 a, int b, int c, int d, int e, u32 f, u32 g, const void *h) {
  (void)a; (void)b; (void)c; (void)d; (void)e; (void)f; (void)g; (void)h;
}
void glTexImage2D(u32 a, int b, int c, int d, int e, int f, u32 g, u32 h, const void *i) {
  (void)a; (void)b; (void)c; (void)d; (void)e; (void)f; (void)g; (void)h; (void)i;
}
void glTexParameteri(u32 a, u32 b, int c) { (void)a; (void)b; (void)c; }
void glViewport(int x, int y, int w, int h) { (void)x; (void)y; (void)w; (void)h; }
int glfwGetKey(void *w, int k) { (void)w; (void)k; return 0; }
