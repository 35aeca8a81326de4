// Shim for the reference arm: the reference reaches NCCL through a dlopen table
// (platform::dynload::ncclXxx); here the same names forward to the linked library.
#pragma once
#include <nccl.h>
namespace paddle {
namespace platform {
namespace dynload {
using ::ncclAllGather;
using ::ncclAllReduce;
using ::ncclCommInitAll;
using ::ncclGroupEnd;
using ::ncclGroupStart;
}  // namespace dynload
}  // namespace platform
}  // namespace paddle
