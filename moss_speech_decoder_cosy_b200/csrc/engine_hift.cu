// HiFT vocoder sequence (placeholder until the kernels land in this file).
#include "engine.h"
namespace mossdec {
int Engine::hift(int, const int32_t*, const float*, const float*, int, const int32_t*, const float*, float*, float*,
                 uint64_t, cudaStream_t) {
    return fail("hift: not implemented yet");
}
int Engine::fade_in_out(float*, const float*, const float*, int, cudaStream_t) { return fail("fade: not implemented yet"); }
}  // namespace mossdec
