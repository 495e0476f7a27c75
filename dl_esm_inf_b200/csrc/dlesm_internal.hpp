// dlesm_internal.hpp -- what the other translation units of libdlesm_b200.so need from the
// runtime (dlesm_runtime.cu): the library stream and the registry of device arrays.
#pragma once

#include <cuda_runtime.h>

namespace dlesm {
// binds the device on first use; DLESM_OK or DLESM_ERR_CUDA (no CPU fallback)
int internal_ensure_device();
cudaStream_t internal_stream();
// pitch / rows of a device array allocated by dlesm_dev_field_alloc or dlesm_dev_grid_upload
int internal_pitch_of(const void *device_ptr, int &pitch, int &ny);
// a kernel outside the runtime wrote this array: its halos are out of date (dirty-flag halos); unknown pointers
// are ignored
void internal_mark_written(const void *device_ptr);
}  // namespace dlesm
