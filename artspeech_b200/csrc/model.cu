// placeholder until the model kernels land (replaced in the next commit)
#include "asb_common.cuh"
struct asb_model { int dummy; };
extern "C" int asb_model_create(const asb_weights*, int, int, int, int, int, int, int, void*, asb_model**) {
  return asb::set_error(ASB_ERR_UNSUPPORTED, "model kernels not built yet");
}
extern "C" void asb_model_destroy(asb_model*) {}
extern "C" size_t asb_model_workspace_bytes(const asb_model*, int, int) { return 0; }
extern "C" int asb_model_forward(asb_model*, const int64_t*, int, const int32_t*, int, float*, void*, size_t, void*) {
  return asb::set_error(ASB_ERR_UNSUPPORTED, "model kernels not built yet");
}
