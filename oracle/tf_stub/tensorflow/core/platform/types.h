#pragma once
#include <cstdint>
namespace tensorflow {
typedef int32_t int32;
typedef int64_t int64;
}  // namespace tensorflow
