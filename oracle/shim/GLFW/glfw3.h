// TEST INFRASTRUCTURE — stand-in for <GLFW/glfw3.h>; there is no graphics stack on this image.
// The reference's CPU back-end (src/update-cpu.cpp:1,124,134,136-137) uses GLFW only for a clock and
// OpenGL only to hand the finished float RGB buffer to a texture.  Here the clock is std::chrono and
// the "texture upload" records the pointer so the harness (oracle/ref_harness.cpp) can read the frame.
#ifndef RT_ORACLE_SHIM_GLFW3_H
#define RT_ORACLE_SHIM_GLFW3_H
#include <chrono>

#define GL_TEXTURE_2D 0x0DE1
#define GL_RGB 0x1907
#define GL_RGBA 0x1908
#define GL_FLOAT 0x1406

inline double glfwGetTime()
{
    using clock = std::chrono::steady_clock;
    static const clock::time_point t0 = clock::now();
    return std::chrono::duration<double>(clock::now() - t0).count();
}

struct rt_shim_texture_upload {
    unsigned int texture;
    int width, height;
    const void *pixels;
};
extern rt_shim_texture_upload g_rt_shim_last_upload;

inline void glBindTexture(unsigned int, unsigned int texture) { g_rt_shim_last_upload.texture = texture; }
inline void glTexImage2D(unsigned int, int, int, int width, int height, int, unsigned int, unsigned int,
                         const void *pixels)
{
    g_rt_shim_last_upload.width = width;
    g_rt_shim_last_upload.height = height;
    g_rt_shim_last_upload.pixels = pixels;
}
#endif
