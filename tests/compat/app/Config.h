// TEST STAND-IN for the application's src/Config.h (only the constants compat/Trainer.cpp reads).
#pragma once
#define RENDER_RESOLUTION_X 256
#define RENDER_RESOLUTION_Y 256
#define SPLATS_LIMIT 1000000
#define SPLATS_SH_DEGREE 1
#define SPLATS_SH_COEF 4
