// TEST STAND-IN with the public interface of the application's src/ModelSplatsHost.h:8-39 (the subset used here).
#pragma once
#include <cuda_runtime.h>
class ModelSplatsDevice;
class ModelSplatsHost {
public:
    int capacity, shDegree, shCoeffs;
    int count = 0;
    float *locations = nullptr, *shs = nullptr, *scales = nullptr, *opacities = nullptr, *rotations = nullptr;
    ModelSplatsHost(int capacity, int shDegree, int shCoeffs) : capacity(capacity), shDegree(shDegree), shCoeffs(shCoeffs) {
        locations = new float[capacity * 3]; shs = new float[capacity * 3 * shCoeffs]; scales = new float[capacity * 3];
        opacities = new float[capacity]; rotations = new float[capacity * 4];
    }
    explicit inline ModelSplatsHost(const ModelSplatsDevice& device);
    ModelSplatsHost(const ModelSplatsHost&) = delete;
    ~ModelSplatsHost() { delete[] locations; delete[] shs; delete[] scales; delete[] opacities; delete[] rotations; }
};
