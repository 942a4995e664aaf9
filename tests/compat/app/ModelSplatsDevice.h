// TEST STAND-IN with the public interface of the application's src/ModelSplatsDevice.h:5-30.
#pragma once
#include <cuda_runtime.h>
#include "ModelSplatsHost.h"
class ModelSplatsDevice {
public:
    int capacity, shDegree, shCoeffs;
    int count = 0;
    float *devLocations = nullptr, *devShs = nullptr, *devScales = nullptr, *devOpacities = nullptr, *devRotations = nullptr;
    explicit ModelSplatsDevice(const ModelSplatsHost& host) : capacity(host.capacity), shDegree(host.shDegree), shCoeffs(host.shCoeffs) {
        count = host.count;
        cudaMalloc(&devLocations, count * 3 * sizeof(float)); cudaMalloc(&devShs, count * 3 * shCoeffs * sizeof(float));
        cudaMalloc(&devScales, count * 3 * sizeof(float)); cudaMalloc(&devOpacities, count * sizeof(float));
        cudaMalloc(&devRotations, count * 4 * sizeof(float));
        cudaMemcpy(devLocations, host.locations, count * 3 * sizeof(float), cudaMemcpyHostToDevice);
        cudaMemcpy(devShs, host.shs, count * 3 * shCoeffs * sizeof(float), cudaMemcpyHostToDevice);
        cudaMemcpy(devScales, host.scales, count * 3 * sizeof(float), cudaMemcpyHostToDevice);
        cudaMemcpy(devOpacities, host.opacities, count * sizeof(float), cudaMemcpyHostToDevice);
        cudaMemcpy(devRotations, host.rotations, count * 4 * sizeof(float), cudaMemcpyHostToDevice);
    }
    ModelSplatsDevice(const ModelSplatsDevice&) = delete;
    ~ModelSplatsDevice() { cudaFree(devLocations); cudaFree(devShs); cudaFree(devScales); cudaFree(devOpacities); cudaFree(devRotations); }
};
inline ModelSplatsHost::ModelSplatsHost(const ModelSplatsDevice& d) : ModelSplatsHost(d.capacity, d.shDegree, d.shCoeffs) {
    count = d.count;
    cudaMemcpy(locations, d.devLocations, count * 3 * sizeof(float), cudaMemcpyDeviceToHost);
    cudaMemcpy(shs, d.devShs, count * 3 * shCoeffs * sizeof(float), cudaMemcpyDeviceToHost);
    cudaMemcpy(scales, d.devScales, count * 3 * sizeof(float), cudaMemcpyDeviceToHost);
    cudaMemcpy(opacities, d.devOpacities, count * sizeof(float), cudaMemcpyDeviceToHost);
    cudaMemcpy(rotations, d.devRotations, count * 4 * sizeof(float), cudaMemcpyDeviceToHost);
}
