// stub of FLTK <FL/gl.h> for the oracle build: scene/light.h:14 only needs the GLenum type
#pragma once
typedef unsigned int GLenum;
