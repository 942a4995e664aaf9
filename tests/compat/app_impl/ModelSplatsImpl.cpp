// TEST INFRASTRUCTURE: bodies for the classes DECLARED by the reference's own src/ModelSplatsDevice.h and
// src/ModelSplatsHost.h (the demo includes those headers in place).  The reference's ModelSplatsDevice.cpp /
// ModelSplatsHost.cpp cannot be compiled here because they include ui/UiPanelViewInput.h (wxWidgets, OpenGL); what they
// do is restated: five arrays, host side sized to `capacity`, device side sized to `count`
// (src/ModelSplatsDevice.cpp:6-48, src/ModelSplatsHost.cpp:6-100).
#include <cstring>
#include <stdexcept>

#include <cuda_runtime.h>

#include "ModelSplatsDevice.h"
#include "ModelSplatsHost.h"

namespace {

// floats per splat of the five arrays, in member order: locations, shs, scales, opacities, rotations
struct Widths { size_t w[5]; };
Widths widths(int shCoeffs) { return {{3, 3 * (size_t)shCoeffs, 3, 1, 4}}; }

void device_fields(ModelSplatsDevice& d, float*** f) {
    f[0] = &d.devLocations; f[1] = &d.devShs; f[2] = &d.devScales; f[3] = &d.devOpacities; f[4] = &d.devRotations;
}
void host_fields(ModelSplatsHost& h, float*** f) {
    f[0] = &h.locations; f[1] = &h.shs; f[2] = &h.scales; f[3] = &h.opacities; f[4] = &h.rotations;
}

void device_alloc_and_fill(ModelSplatsDevice& d, float* const* src, cudaMemcpyKind kind) {
    float** f[5];
    device_fields(d, f);
    const Widths w = widths(d.shCoeffs);
    for (int k = 0; k < 5; k++) {
        const size_t bytes = (size_t)d.count * w.w[k] * sizeof(float);
        cudaMalloc(f[k], bytes);
        cudaMemcpy(*f[k], src[k], bytes, kind);
    }
}

void host_alloc(ModelSplatsHost& h) {
    float** f[5];
    host_fields(h, f);
    const Widths w = widths(h.shCoeffs);
    for (int k = 0; k < 5; k++) *f[k] = new float[(size_t)h.capacity * w.w[k]];
}

}  // namespace

ModelSplatsDevice::ModelSplatsDevice(const ModelSplatsDevice& device)
    : capacity(device.capacity), shDegree(device.shDegree), shCoeffs(device.shCoeffs) {
    count = device.count;
    float* const src[5] = {device.devLocations, device.devShs, device.devScales, device.devOpacities, device.devRotations};
    device_alloc_and_fill(*this, src, cudaMemcpyDeviceToDevice);
}

ModelSplatsDevice::ModelSplatsDevice(const ModelSplatsHost& host)
    : capacity(host.capacity), shDegree(host.shDegree), shCoeffs(host.shCoeffs) {
    count = host.count;
    float* const src[5] = {host.locations, host.shs, host.scales, host.opacities, host.rotations};
    device_alloc_and_fill(*this, src, cudaMemcpyHostToDevice);
}

ModelSplatsDevice::~ModelSplatsDevice() {
    float** f[5];
    device_fields(*this, f);
    for (int k = 0; k < 5; k++) cudaFree(*f[k]);
}

ModelSplatsHost::ModelSplatsHost(int capacity, int shDegree, int shCoeffs)
    : capacity(capacity), shDegree(shDegree), shCoeffs(shCoeffs) {
    host_alloc(*this);
}

ModelSplatsHost::ModelSplatsHost(const ModelSplatsDevice& device)
    : ModelSplatsHost(device.capacity, device.shDegree, device.shCoeffs) {
    count = device.count;
    float** f[5];
    host_fields(*this, f);
    float* const src[5] = {device.devLocations, device.devShs, device.devScales, device.devOpacities, device.devRotations};
    const Widths w = widths(shCoeffs);
    for (int k = 0; k < 5; k++) cudaMemcpy(*f[k], src[k], (size_t)count * w.w[k] * sizeof(float), cudaMemcpyDeviceToHost);
}

ModelSplatsHost::ModelSplatsHost(const std::vector<float>& locationsArg, const std::vector<float>& shsArg,
                                 const std::vector<float>& scalesArg, const std::vector<float>& opacitiesArg,
                                 const std::vector<float>& rotationsArg) {
    count = (int)(locationsArg.size() / 3);
    capacity = 1000000;                                    // grows by decades (src/ModelSplatsHost.cpp:31-32)
    while ((size_t)capacity < locationsArg.size() / 3) capacity *= 10;
    shCoeffs = count > 0 ? (int)shsArg.size() / (3 * count) : 1;
    shDegree = (shCoeffs - 1) / 3;
    const std::vector<float>* src[5] = {&locationsArg, &shsArg, &scalesArg, &opacitiesArg, &rotationsArg};
    const Widths w = widths(shCoeffs);
    for (int k = 0; k < 5; k++)
        if (src[k]->size() != (size_t)count * w.w[k])
            throw std::runtime_error("Inconsistent feature dimensions supplied when creating a host model!");
    host_alloc(*this);
    float** f[5];
    host_fields(*this, f);
    for (int k = 0; k < 5; k++) memcpy(*f[k], src[k]->data(), src[k]->size() * sizeof(float));
}

ModelSplatsHost::~ModelSplatsHost() {
    float** f[5];
    host_fields(*this, f);
    for (int k = 0; k < 5; k++) delete[] *f[k];
}

void ModelSplatsHost::pushBack(glm::vec3 location, std::vector<float> sh, glm::vec3 scale, float opacity, glm::quat rotation) {
    if (count >= capacity) throw std::runtime_error("Model ran out of capacity!");
    const size_t c = (size_t)count;
    for (int k = 0; k < 3; k++) { locations[c * 3 + k] = location[k]; scales[c * 3 + k] = scale[k]; }
    for (int k = 0; k < shCoeffs * 3; k++) shs[c * 3 * shCoeffs + k] = sh.at((size_t)k);
    opacities[c] = opacity;
    memcpy(&rotations[c * 4], &rotation, 4 * sizeof(float));   // glm::quat's in-memory component order
    count++;
}

void ModelSplatsHost::copy(int indexTo, int indexFrom) {
    if (indexTo < 0 || indexTo >= count || indexFrom < 0 || indexFrom >= count)
        throw std::runtime_error("Can't copy splat in model, incorrect bounds and/or no capacity!");
    float** f[5];
    host_fields(*this, f);
    const Widths w = widths(shCoeffs);
    for (int k = 0; k < 5; k++)
        memmove(*f[k] + (size_t)indexTo * w.w[k], *f[k] + (size_t)indexFrom * w.w[k], w.w[k] * sizeof(float));
}
