// swarm_emit.cuh — stage (4): observation emission from the packed resident grid.
//   FULL   obs[n][y*R+x] = grid value            (np.array(grid).T.flatten(), A2C/noVis_sim.py:168-170)
//   WINDOW win[n][i][b*W+a] = grid[cx-h+a][cy-h+b], pad outside, centre (floor(x+.5), floor(y+.5))  
// The packed flat order equals the obs order, so FULL is a streaming 2-bit -> element expansion with 128-bit
// (float) / 32-bit (uint8) coalesced stores.
#pragma once
#include "swarm_common.cuh"

namespace swarm {

template <typename OT> struct ObsCvt;
template <> struct ObsCvt<float> {
  float u;
  __device__ explicit ObsCvt(const DevParams& p) : u((float)p.U) {}
  __device__ __forceinline__ float one(uint32_t code) const { return code_to_f32(code, u); }
  __device__ __forceinline__ void store4(float* dst, uint32_t byte) const {   // 4 cells = 8 bits, dst 16 B aligned
    const float4 v = make_float4(one(byte & 3u), one((byte >> 2) & 3u), one((byte >> 4) & 3u), one((byte >> 6) & 3u));
    __stcs(reinterpret_cast<float4*>(dst), v);
  }
};
template <> struct ObsCvt<uint8_t> {
  uint32_t u;
  __device__ explicit ObsCvt(const DevParams& p) : u((uint32_t)p.U) {}
  __device__ __forceinline__ uint8_t one(uint32_t code) const { return (uint8_t)code_to_u8(code, u); }
  __device__ __forceinline__ void store4(uint8_t* dst, uint32_t byte) const {   // dst 4 B aligned
    const uint32_t v = code_to_u8(byte & 3u, u) | (code_to_u8((byte >> 2) & 3u, u) << 8) |
                       (code_to_u8((byte >> 4) & 3u, u) << 16) | (code_to_u8((byte >> 6) & 3u, u) << 24);
    __stcs(reinterpret_cast<uint32_t*>(dst), v);
  }
};

// Emits the obs of envs [env0, env0+nenv) with the whole warp.  Reads go to L2 (__ldcg) because the same kernel
// has just modified the grid / positions from other lanes.
template <typename OT>
__device__ __forceinline__ void emit_obs(const DevParams& p, int env0, int nenv, int lane, OT* __restrict__ obs) {
  const ObsCvt<OT> cvt(p);
  if (p.obs_mode == 0) {
    const int RC = p.RC;
    for (int e = 0; e < nenv; ++e) {
      const uint32_t* g = p.grid + (size_t)(env0 + e) * p.env_words;
      OT* o = obs + (size_t)(env0 + e) * RC;
      if ((RC & 3) == 0) {
        const int nq = RC >> 2;
#pragma unroll 4
        for (int q = lane; q < nq; q += 32) {
          const uint32_t w = ld_grid(g + (q >> 2));
          cvt.store4(o + 4 * q, (w >> ((q & 3) * 8)) & 0xffu);
        }
      } else {
        for (int c = lane; c < RC; c += 32) o[c] = cvt.one((ld_grid(g + (c >> 4)) >> ((c & 15) * 2)) & 3u);
      }
    }
  } else {
    const int W = p.W, WW = W * W, per = p.D * WW, h = (W - 1) >> 1;
    for (int e = 0; e < nenv; ++e) {
      const uint32_t* g = p.grid + (size_t)(env0 + e) * p.env_words;
      const float2* ps = p.pos + (size_t)(env0 + e) * p.D;
      OT* o = obs + (size_t)(env0 + e) * per;
      for (int f = lane; f < per; f += 32) {
        const int i = f / WW, k = f - i * WW, b = k / W, a = k - b * W;
        const float2 q = __ldcg(ps + i);
        const int gx = (int)floorf(q.x + 0.5f) - h + a, gy = (int)floorf(q.y + 0.5f) - h + b;
        uint32_t code = 3u;
        if (gx >= 0 && gx < p.R && gy >= 0 && gy < p.C) {
          const int cell = gy * p.R + gx;
          code = (ld_grid(g + (cell >> 4)) >> ((cell & 15) * 2)) & 3u;
        }
        o[f] = cvt.one(code);
      }
    }
  }
}

}  // namespace swarm
