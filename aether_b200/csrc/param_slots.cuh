// Parameter slots of the step (state_dict order, see include/aether_b200.h).
#pragma once

namespace aether {

enum {
  P_OUT0_W = 0, P_OUT0_B, P_OUT3_W, P_OUT3_B, P_OUT6_W, P_OUT6_B,
  P_L1_M0_W = 6, P_L1_M0_B, P_L1_M2_W, P_L1_M2_B, P_L1_RES_W, P_L1_RES_B, P_L1_U0_W, P_L1_U0_B, P_L1_U2_W, P_L1_U2_B,
  P_L2_BASE = 16,
  P_F0_W = 40, P_F0_B, P_F2_W, P_F2_B, P_F4_W, P_F4_B, P_EMB = 46, P_NUM = 47
};
struct LayerSlots { int m0w, m0b, m2w, m2b, u0w, u0b, u2w, u2b; };
__host__ __device__ inline LayerSlots layer_slots(int l) {
  if (l == 1) return {P_L1_M0_W, P_L1_M0_B, P_L1_M2_W, P_L1_M2_B, P_L1_U0_W, P_L1_U0_B, P_L1_U2_W, P_L1_U2_B};
  const int b = P_L2_BASE + 8 * (l - 2);
  return {b, b + 1, b + 2, b + 3, b + 4, b + 5, b + 6, b + 7};
}

}  // namespace aether
