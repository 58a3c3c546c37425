// Internal definitions shared by the C-ABI translation units (host side).
#pragma once
#include <string>

#include "../../include/tokp.h"
#include "model.h"

struct DeviceModel;  // defined in device_model.h (CUDA side)

struct tokp_model {
  tokp::Model host;
  DeviceModel* dev = nullptr;  // lazily uploaded on first device use
};

namespace tokp {
void set_error(const std::string& msg);
// CUDA side (capi_cuda.cu)
void device_model_free(DeviceModel* dm);
}  // namespace tokp
