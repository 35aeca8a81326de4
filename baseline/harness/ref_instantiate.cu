// The reference instantiates AccessorWrapper<CommonFeatureValueAccessor> (the CopyForPull /
// CopyForPush host functions and their kernels) at the end of feature_value.cu, but only under
// PADDLE_WITH_PSCORE — the build flavour that also links the brpc CPU parameter server.  This
// translation unit includes the unmodified file and states the same instantiation for the
// PSCORE-less build of the reference arm.
#include "paddle/fluid/framework/fleet/heter_ps/feature_value.cu"

namespace paddle {
namespace framework {
template class AccessorWrapper<CommonFeatureValueAccessor>;
}  // namespace framework
}  // namespace paddle
