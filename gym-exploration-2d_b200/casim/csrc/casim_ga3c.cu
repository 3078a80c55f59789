#include "casim_device.cuh"
cudaError_t casim_launch_ga3c(casim_handle*, const float*, float*, int32_t*, float*, float*, int, cudaStream_t) { return cudaErrorNotSupported; }
cudaError_t casim_ga3c_upload(casim_handle*, const float*, int, cudaStream_t) { return cudaErrorNotSupported; }
