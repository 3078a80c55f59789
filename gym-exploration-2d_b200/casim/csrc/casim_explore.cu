#include "casim_device.cuh"
cudaError_t casim_launch_explore(casim_handle*, const float*, const float*, float*, float*, float*, float*,
                                 unsigned long long*, cudaStream_t) { return cudaErrorNotSupported; }
