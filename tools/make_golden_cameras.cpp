// Generates tests/golden/cameras.json from the reference's OWN Camera.cpp / Project.h, compiled in place
// (nothing copied): run tools/make_golden_cameras.sh where /root/reference exists.
#include <cstdio>
#include <vector>
#include "Camera.h"
#include "Project.h"
#include <diff-gaussian-rasterization/third_party/glm/glm/gtc/type_ptr.hpp>

static void dump(FILE* f, const char* name, const float* p, int n, bool last = false) {
    fprintf(f, "    \"%s\": [", name);
    for (int i = 0; i < n; i++) fprintf(f, "%s%.9g", i ? ", " : "", p[i]);
    fprintf(f, "]%s\n", last ? "" : ",");
}

int main(int argc, char** argv) {
    FILE* f = fopen(argv[1], "w");
    struct Cfg { int n1, n2; float d1, d2, f1, f2, rx1, ry1, rx2, ry2; int W, H; };
    const Cfg cfgs[] = {{1, 0, 10, 10, 60, 30, 0, 0, 0, 0, 256, 256},   {8, 0, 10, 10, 60, 30, 0, 0, 0, 0, 512, 512},
                        {16, 16, 10, 10, 60, 20, 0, 0, 0, 0, 1024, 1024}, {5, 3, 7.5f, 12, 45, 25, 30, -20, 110, 15, 1920, 1080}};
    fprintf(f, "[\n");
    bool first = true;
    for (const Cfg& c : cfgs) {
        Project project;
        project.sphere1.count = c.n1; project.sphere1.distance = c.d1; project.sphere1.fovDeg = c.f1;
        project.sphere1.rotX = c.rx1; project.sphere1.rotY = c.ry1;
        project.sphere2.count = c.n2; project.sphere2.distance = c.d2; project.sphere2.fovDeg = c.f2;
        project.sphere2.rotX = c.rx2; project.sphere2.rotY = c.ry2;
        std::vector<Camera> cams = Camera::getCameras(project);
        for (size_t k = 0; k < cams.size(); k++) {
            const Camera& cam = cams[k];
            glm::mat4 view = cam.getView();
            glm::mat4 projview = cam.getProjection((float)c.W / (float)c.H) * view;   // src/Trainer.cu:321-322
            fprintf(f, "%s  {\n    \"n1\": %d, \"n2\": %d, \"d1\": %.9g, \"d2\": %.9g, \"f1\": %.9g, \"f2\": %.9g,\n", first ? "" : ",\n",
                    c.n1, c.n2, c.d1, c.d2, c.f1, c.f2);
            fprintf(f, "    \"rx1\": %.9g, \"ry1\": %.9g, \"rx2\": %.9g, \"ry2\": %.9g, \"W\": %d, \"H\": %d, \"index\": %zu, \"fov\": %.9g,\n",
                    c.rx1, c.ry1, c.rx2, c.ry2, c.W, c.H, k, cam.fovDegY);
            dump(f, "location", &cam.location[0], 3);
            dump(f, "view", glm::value_ptr(view), 16);
            dump(f, "projview", glm::value_ptr(projview), 16, true);
            fprintf(f, "  }");
            first = false;
        }
    }
    fprintf(f, "\n]\n");
    fclose(f);
    return 0;
}
