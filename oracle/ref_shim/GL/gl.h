/* Type-only stand-in for <GL/gl.h> (pulled in by cuda_gl_interop.h through the reference's
 * include/Global/Global.cuh:80-82).  Declarations only; no graphics API is brought up. */
#pragma once
#include <OGL/glad.h>
