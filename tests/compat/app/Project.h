// TEST STAND-IN with the fields of the application's src/Project.h:6-62 that the training path reads.
#pragma once
class Project {
public:
    struct CameraSphere { int count = 16; float distance = 10.0f; float fovDeg = 60.0f; float rotX = 0.0f; float rotY = 0.0f; } sphere1, sphere2;
    int rtSamples = 100;
    float lrLocation = 0.00005f, lrSh = 0.0001f, lrScale = 0.00002f, lrOpacity = 0.0001f, lrRotation = 0.000025f;
    float paramScaleMax = 0.3f;
    float paramCullOpacity = 0.005f, paramCullSize = 0.004f, paramDensifyVariance = 2.0f;
    float paramSplitSize = 0.04f, paramSplitDistance = 1.5f, paramSplitScale = 0.8f, paramCloneDistance = 1.6f;
    int iterations = 0, intervalCapture = 50, intervalDensify = 200;
};
