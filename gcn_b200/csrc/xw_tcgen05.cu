// placeholder until the tcgen05 kernel lands
#include "common.h"
extern "C" int gcnb_xw(const float* X, int64_t ldx, int trans_x, const float* W, int64_t ldw,
                       int trans_w, float* D, int64_t ldd, int64_t M, int N, int K,
                       const float* bias, int epilogue, int precision, void* stream) {
  return gcnb::set_error(GCNB_ERR_UNSUPPORTED, "gcnb_xw: not built yet");
}
