#pragma once
#include "device.hpp"
namespace ug {
void field_vec_op(DeviceCtx& ctx, int field, int op, const void* a, const void* b, void* out, uint64_t n);
}
