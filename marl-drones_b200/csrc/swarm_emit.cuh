// swarm_emit.cuh — stage (4): observation emission from the packed resident grid.
//   FULL   obs[n][y*R+x] = grid value            (np.array(grid).T.flatten(), A2C/noVis_sim.py:168-170)
//   WINDOW win[n][i][b*W+a] = grid[cx-h+a][cy-h+b], pad outside, centre (floor(x+.5), floor(y+.5))  
// The packed flat order equals the obs order, so FULL is a streaming 2-bit -> element expansion with 128-bit stores.
//
// Code -> element conversion is one PRMT per element: the four possible float32 values {0, 1, U, -1} differ only in their
// two top bytes, so two 4-entry byte tables (`lt`, `ls`) indexed by the 2-bit code give the value.
#pragma once
#include "swarm_common.cuh"

namespace swarm {

template <typename OT> struct ObsCvt;
template <> struct ObsCvt<float> {
  uint32_t lt, ls;
  __device__ explicit ObsCvt(const DevParams& p) {
    const uint32_t fu = __float_as_uint((float)p.U);   // U <= 127: the low 16 bits of float(U) are zero
    lt = 0x00u | (0x3Fu << 8) | ((fu >> 24) << 16) | (0xBFu << 24);
    ls = 0x00u | (0x80u << 8) | (((fu >> 16) & 0xFFu) << 16) | (0x80u << 24);
  }
  __device__ __forceinline__ float one(uint32_t code) const {   // code in [0,3]
    return __uint_as_float(__byte_perm(lt, ls, code * 0x1100u + 0x0400u));
  }
  // element k of a packed 2-bit field
  __device__ __forceinline__ float sel(uint32_t field, int k) const { return one((field >> (2 * k)) & 3u); }
  // 4 cells = 8 bits -> one 128-bit store (dst 16 B aligned)
  __device__ __forceinline__ void store4(float* dst, uint32_t byte) const {
    __stcs(reinterpret_cast<float4*>(dst), make_float4(sel(byte, 0), sel(byte, 1), sel(byte, 2), sel(byte, 3)));
  }
};
template <> struct ObsCvt<uint8_t> {
  uint32_t lut;
  __device__ explicit ObsCvt(const DevParams& p) : lut(0u | (1u << 8) | ((uint32_t)p.U << 16) | (255u << 24)) {}
  __device__ __forceinline__ uint8_t one(uint32_t code) const { return (uint8_t)__byte_perm(lut, 0u, (code & 3u) | 0x4440u); }
  __device__ __forceinline__ uint8_t sel(uint32_t field, int k) const { return one((field >> (2 * k)) & 3u); }
  __device__ __forceinline__ uint32_t pack4(uint32_t byte) const {   // 4 codes -> 4 bytes
    const uint32_t s = (byte & 3u) | (((byte >> 2) & 3u) << 4) | (((byte >> 4) & 3u) << 8) | (((byte >> 6) & 3u) << 12);
    return __byte_perm(lut, 0u, s);
  }
  __device__ __forceinline__ void store4(uint8_t* dst, uint32_t byte) const {   // dst 4 B aligned
    __stcs(reinterpret_cast<uint32_t*>(dst), pack4(byte));
  }
};

// ---- CTA-wide emission straight from the global packed grid (swarm_cta.cuh: large grids / D > 32) --------------------
// Reads go to L2 (__ldcg): the same kernel has just modified the grid / positions from another warp of the CTA.
template <typename OT>
__device__ __forceinline__ void emit_obs_global(const DevParams& p, int env, int tid, int nthr, OT* __restrict__ obs) {
  const ObsCvt<OT> cvt(p);
  const uint32_t* g = p.grid + (size_t)env * p.env_words;
  if (p.obs_mode == 0) {
    const int RC = p.RC;
    OT* o = obs + (size_t)env * RC;
    if ((RC & 3) == 0 && ((uintptr_t)o & 15) == 0) {
      const int nq = RC >> 2;
#pragma unroll 4
      for (int q = tid; q < nq; q += nthr) {
        const uint32_t w = ld_grid(g + (q >> 2));
        cvt.store4(o + 4 * q, (w >> ((q & 3) * 8)) & 0xffu);
      }
    } else {
      for (int c = tid; c < RC; c += nthr) o[c] = cvt.one((ld_grid(g + (c >> 4)) >> ((c & 15) * 2)) & 3u);
    }
  } else {
    // one item = one window row of one drone: W consecutive output elements from <= 2 packed words
    const int W = p.W, WW = W * W, h = (W - 1) >> 1, nitem = p.D * W;
    const float2* ps = p.pos + (size_t)env * p.D;
    OT* o = obs + (size_t)env * p.D * WW;
    for (int it = tid; it < nitem; it += nthr) {
      const int i = it / W, b = it - i * W;
      const float2 q = __ldcg(ps + i);
      const int cx = (int)floorf(q.x + 0.5f), gy = (int)floorf(q.y + 0.5f) - h + b;
      OT* dst = o + (size_t)i * WW + b * W;
      if ((unsigned)gy >= (unsigned)p.C) {
        for (int a = 0; a < W; ++a) dst[a] = cvt.one(3u);
        continue;
      }
      for (int a = 0; a < W; ++a) {
        const int gx = cx - h + a;
        uint32_t code = 3u;
        if ((unsigned)gx < (unsigned)p.R) {
          const int cell = gy * p.R + gx;
          code = (ld_grid(g + (cell >> 4)) >> ((cell & 15) * 2)) & 3u;
        }
        dst[a] = cvt.one(code);
      }
    }
  }
}

}  // namespace swarm
