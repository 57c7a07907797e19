// update_b200.h — the reference's back-end call surface (include/update.h:6-8) on top of librt_b200.so.
//
//     void  init_update(unsigned int texture, const Scene& scene);
//     float update(const glm::dmat4& camera_matrix);      // here: 16 doubles, GLM column-major (&m[0][0])
//     void  cleanup_update();
//
// Same names, same argument meaning, same state model (one scene per process, file-scope state, not re-entrant)
// and the same error behaviour: a failure prints the library's message and exits with EXIT_FAILURE
// (include/helper_cuda_opengl.h:13-44).  The GL texture is replaced by a host frame (RGBA8, row 0 = bottom)
// readable through update_frame(); `texture` is accepted and ignored (no graphics stack on this box).
// INTEGRATION.md shows the same adapter written against the reference's own headers (glm::dmat4, its Scene).
#ifndef UPDATE_B200_H
#define UPDATE_B200_H

#include <cstddef>

#include "rt_host_scene.h"

void init_update(unsigned int texture, const rt_host::Scene &scene);
float update(const double camera_matrix[16]);
void cleanup_update();

// The frame the last update() produced (what the GL texture would hold): width * height RGBA8 pixels.
const unsigned char *update_frame(size_t *bytes);
// Select devices before init_update (default: device 0, single GPU).
void update_set_option(const char *name, long long value);

#endif
