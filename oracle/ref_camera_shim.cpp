// TEST INFRASTRUCTURE (never the product): extern "C" access to the reference's OWN Camera::getCameras
// (src/Camera.cpp:33-58, compiled in place from the reference tree together with its glm and nlohmann/json) so that
// bench.py's reference arm generates its camera set without loading libgsb200.so.  Built into oracle/_ref/libgsrefcam.so
// by oracle/Makefile; host code only.
#include <vector>

#include "Camera.h"
#include "Project.h"

extern "C" {

// spheres: {count, distance, fovDeg, rotX, rotY} x 2.  Writes up to `cap` cameras as {x, y, z, fovDegY}; returns the count.
int gsrefcam_get_cameras(const float* sphere1, const float* sphere2, float* out4, int cap) {
    Project project;
    project.sphere1.count = (int)sphere1[0]; project.sphere1.distance = sphere1[1]; project.sphere1.fovDeg = sphere1[2];
    project.sphere1.rotX = sphere1[3]; project.sphere1.rotY = sphere1[4];
    project.sphere2.count = (int)sphere2[0]; project.sphere2.distance = sphere2[1]; project.sphere2.fovDeg = sphere2[2];
    project.sphere2.rotX = sphere2[3]; project.sphere2.rotY = sphere2[4];
    const std::vector<Camera> cams = Camera::getCameras(project);
    const int n = (int)cams.size();
    for (int i = 0; i < n && i < cap; i++) {
        out4[4 * i + 0] = cams[(size_t)i].location.x;
        out4[4 * i + 1] = cams[(size_t)i].location.y;
        out4[4 * i + 2] = cams[(size_t)i].location.z;
        out4[4 * i + 3] = cams[(size_t)i].fovDegY;
    }
    return n;
}

}  // extern "C"
