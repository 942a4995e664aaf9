// compat/Trainer.cuh -- the reference's `Trainer` (src/Trainer.cuh:9-75) over libgsb200.so.
//
// Same public members, same member-function signatures, so the wxWidgets callers
// (src/ui/UiFrame.cpp:292, src/ui/tools/UiPanelToolsTrain.cpp:166-169, src/ui/UiPanelViewOutput.cpp:55,
// src/ui/tools/UiPanelToolsView.cpp:251, src/ui/tools/UiPanelToolsTruth.cpp:187) compile unchanged.
// The implementation is compat/Trainer.cpp: build it in place of src/Trainer.cu.
// The reference's 21 private device buffers (src/Trainer.cuh:13-47) are gone: the library keeps its
// own arenas behind one opaque handle.
#pragma once

#include <cstdint>
#include <vector>

class ModelSplatsDevice;
class Camera;
class Project;
class RtxHost;
struct gsb_trainer;

class Trainer {

private:
    gsb_trainer* impl = nullptr;     // gsb_trainer_create(RENDER_RESOLUTION_X, RENDER_RESOLUTION_Y)
    int lastTruthCount = -1;         // as src/Trainer.cuh:16: re-allocate truth frames when the camera count changes

    // GSB_COMPAT_DEVICES=G (environment, read by the constructor): the truth cameras are sharded over G GPUs of this
    // process -- one replica trainer, one host thread and one NCCL communicator per extra device; `model`, the truth
    // frames and render() stay on the application's device exactly as before.  Null on a single GPU.
    struct Multi;
    Multi* multi = nullptr;
    void trainMulti(Project& project, bool densify);

public:
    ModelSplatsDevice* model;

    std::vector<uint32_t*> truthFrameBuffersW;   // truth frames rendered on white (device RGBA8)
    std::vector<uint32_t*> truthFrameBuffersB;   // truth frames rendered on black
    std::vector<Camera> truthCameras;

    Trainer();

    Trainer(const Trainer&) = delete;
    Trainer& operator=(const Trainer&) = delete;
    Trainer(Trainer&&) = delete;
    Trainer& operator=(Trainer&&) = delete;

    ~Trainer();

    // src/Trainer.cu:148-216
    void render(uint32_t* frameBuffer, int sizeX, int sizeY, float splatScale, const Camera& camera);

    // src/Trainer.cu:218-250 (needs the application's OptiX tracer; compiled out with GSB_COMPAT_NO_RTX)
    void captureTruths(const Project& project, RtxHost& rtx);

    // src/Trainer.cu:252-543
    void train(Project& project, bool densify);

};
