// TEST INFRASTRUCTURE — stands in for <cuda_runtime.h> in the CUDA-on-CPU emulation
#pragma once
#include "cuda_emu.h"
