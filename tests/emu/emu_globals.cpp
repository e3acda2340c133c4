// TEST INFRASTRUCTURE ONLY: storage for the emulated CUDA built-ins (see emu_cuda.h).
#include "emu_cuda.h"
thread_local emu_dim threadIdx, blockIdx, blockDim, gridDim;
thread_local std::vector<double> emu_smem;
