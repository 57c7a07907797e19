// rt_camera.h — the caller side of update(): fly-camera state -> camera matrix (reference src/ray-tracer.cpp:25-58).
//
// The reference keeps position / yaw / pitch (degrees), derives `direction` (ray-tracer.cpp:44-48) and hands
// inverse(lookAt(position, position - direction, up)) to update() every frame (ray-tracer.cpp:54-58).  lookAt is
// GLM's right-handed one; its inverse is written out in closed form here (rotation transposed, translation = eye),
// 16 doubles in GLM's column-major order — what rt_b200_update expects.
#ifndef RT_CAMERA_H
#define RT_CAMERA_H

#include <cmath>

namespace rt_host {

struct Camera {
    double position[3] = {0.0, 0.0, 0.0};   // ray-tracer.cpp:25-32: initial pose
    double yaw = 90.0, pitch = 0.0;         // degrees

    void direction(double d[3]) const   // update_direction(), ray-tracer.cpp:44-48
    {
        const double k = 0.01745329251994329576923690768489;   // glm::radians
        d[0] = std::cos(yaw * k) * std::cos(pitch * k);
        d[1] = std::sin(pitch * k);
        d[2] = std::sin(yaw * k) * std::cos(pitch * k);
    }

    // camera_matrix(), ray-tracer.cpp:54-58
    void matrix(double m[16]) const
    {
        double dir[3];
        direction(dir);
        // lookAt(eye, center = eye - dir, up = (0,1,0)): f = normalize(center - eye) = -dir normalised
        double f[3] = {-dir[0], -dir[1], -dir[2]};
        normalize(f);
        const double up[3] = {0.0, 1.0, 0.0};
        double s[3] = {f[1] * up[2] - f[2] * up[1], f[2] * up[0] - f[0] * up[2], f[0] * up[1] - f[1] * up[0]};
        normalize(s);
        const double u[3] = {s[1] * f[2] - s[2] * f[1], s[2] * f[0] - s[0] * f[2], s[0] * f[1] - s[1] * f[0]};
        // world-from-camera = inverse(view): columns are the camera axes (s, u, -f) and the eye position
        for (int r = 0; r < 3; r++) {
            m[0 + r] = s[r];
            m[4 + r] = u[r];
            m[8 + r] = -f[r];
            m[12 + r] = position[r];
        }
        m[3] = m[7] = m[11] = 0.0;
        m[15] = 1.0;
    }

private:
    static void normalize(double v[3])
    {
        const double inv = 1.0 / std::sqrt((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]);
        v[0] *= inv; v[1] *= inv; v[2] *= inv;
    }
};

} // namespace rt_host
#endif
