// Forwarding header: lets a translation unit written against the reference's
// dietgpu/ans/GpuANSCodec.h compile against libdietgpu_b200.so (add -I<repo>/include/compat -I<repo>/include).
#pragma once
#include "dietgpu_b200.hpp"
