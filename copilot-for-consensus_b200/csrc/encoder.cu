// encoder.cu -- host side of the BERT-class encoder behind the C ABI (placeholder until the
// encoder milestone lands; every entry point reports B200_ERR_UNSUPPORTED).
#include "host_util.h"
#include "encoder_kernels.cuh"

using namespace b200;

struct b200_encoder { int hidden = 0; };

extern "C" {
int b200_encoder_create(const b200_encoder_config_t*, int, b200_encoder_t** out) {
  if (out) *out = nullptr;
  B200_FAIL(B200_ERR_UNSUPPORTED, "encoder not built yet");
}
int b200_encoder_destroy(b200_encoder_t* e) { delete e; return B200_OK; }
int b200_encoder_load_tensor(b200_encoder_t*, const char*, const float*, int64_t) { B200_FAIL(B200_ERR_UNSUPPORTED, "encoder not built yet"); }
int b200_encoder_finalize(b200_encoder_t*) { B200_FAIL(B200_ERR_UNSUPPORTED, "encoder not built yet"); }
int b200_encoder_hidden(const b200_encoder_t* e) { return e ? e->hidden : 0; }
int b200_encode(b200_encoder_t*, const int32_t*, const int32_t*, int, int, int, float*, int, void*) { B200_FAIL(B200_ERR_UNSUPPORTED, "encoder not built yet"); }
}
