// encoder_kernels.cuh -- device code for the chunk-embedding path (filled in by the encoder milestone)
#pragma once
#include "common.cuh"
