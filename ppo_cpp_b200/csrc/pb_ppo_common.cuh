// pb_ppo_common.cuh — parameter layout of the ppo_continuous_action agent (src/ppo_continuous_action.cpp:120-169)
// in one flat fp32 buffer, in the order torch registers them (critic 0/2/4 weight+bias, actor_mean 0/2/4, logstd),
// plus the counter-based RNG used for action noise.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define PB_HID 64

struct AgentLayout {
  int obs, act;
  // offsets into the flat parameter / gradient / Adam-moment buffers
  int cW1, cb1, cW2, cb2, cW3, cb3, aW1, ab1, aW2, ab2, aW3, ab3, logstd, total;
  __host__ __device__ static AgentLayout make(int O, int A) {
    AgentLayout l;
    l.obs = O; l.act = A;
    int p = 0;
    l.cW1 = p; p += PB_HID * O; l.cb1 = p; p += PB_HID; l.cW2 = p; p += PB_HID * PB_HID; l.cb2 = p; p += PB_HID;
    l.cW3 = p; p += PB_HID; l.cb3 = p; p += 1;
    l.aW1 = p; p += PB_HID * O; l.ab1 = p; p += PB_HID; l.aW2 = p; p += PB_HID * PB_HID; l.ab2 = p; p += PB_HID;
    l.aW3 = p; p += A * PB_HID; l.ab3 = p; p += A;
    l.logstd = p; p += A;
    l.total = p;
    return l;
  }
};

// Philox4x32-10 (Salmon et al. 2011), counter-based: one call yields four 32-bit words for (counter, key).
__host__ __device__ inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; r++) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}
// two standard normals from two 32-bit words (Box-Muller on (0,1] x [0,1))
__device__ inline void box_muller(uint32_t a, uint32_t b, float *z0, float *z1) {
  float u1 = ((float)a + 1.0f) * 2.3283064365386963e-10f;  // (0, 1]
  float u2 = (float)b * 2.3283064365386963e-10f;
  float r = sqrtf(-2.0f * logf(u1)), sn, cs;
  sincospif(2.0f * u2, &sn, &cs);
  *z0 = r * cs; *z1 = r * sn;
}
