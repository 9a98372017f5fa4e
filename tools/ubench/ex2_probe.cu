// probe: what do the packed exponentials compile to on sm_100a?
#include <cstdint>
extern "C" __global__ void ex2_f16x2(const uint32_t* in, uint32_t* out) {
  uint32_t v = in[threadIdx.x], r;
  asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(r) : "r"(v));
  out[threadIdx.x] = r;
}
extern "C" __global__ void ex2_bf16x2(const uint32_t* in, uint32_t* out) {
  uint32_t v = in[threadIdx.x], r;
  asm volatile("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(r) : "r"(v));
  out[threadIdx.x] = r;
}
extern "C" __global__ void ex2_f32(const float* in, float* out) {
  float v = in[threadIdx.x], r;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  out[threadIdx.x] = r;
}
