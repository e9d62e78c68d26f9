// Device-side include of the shared optimizer math (csrc/optim_math.h).
#pragma once
#include "common.cuh"
#include "optim_math.h"
