// TEST STAND-IN with the public interface of the application's src/Camera.h:12-24.
#pragma once
#include <vector>
#include "miniglm.h"
class Project;
class Camera {
public:
    static int getCamerasCount(const Project& project);
    static std::vector<Camera> getCameras(const Project& project);
    glm::vec3 location;
    glm::vec3 target;
    float fovDegY;
    Camera(const glm::vec3& location, const glm::vec3& target, float fovDegY) : location(location), target(target), fovDegY(fovDegY) {}
    glm::mat4 getView() const { return -glm::lookAt(location, target, glm::vec3{0.0f, 1.0f, 0.0f}); }
    glm::mat4 getProjection(float aspect) const { return glm::perspective(glm::radians(fovDegY), aspect, 0.1f, 100.0f); }
};
