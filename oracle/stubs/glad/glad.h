/* Stand-in for <glad/glad.h> (TEST INFRASTRUCTURE): the reference's camera.hpp:4 includes it for ONE type,
 * GLboolean (camera.hpp:93).  No OpenGL is declared, loaded or called. */
#ifndef RT_ORACLE_GLAD_STUB_H
#define RT_ORACLE_GLAD_STUB_H
typedef unsigned char GLboolean;
#endif
