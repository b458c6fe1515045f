// swarm_emit.cuh — stage (4): observation emission from the resident coverage map.
//   FULL   obs[n][y*R+x] = grid value            (np.array(grid).T.flatten(), A2C/noVis_sim.py:168-170)
//   WINDOW win[n][i][b*W+a] = grid[cx-h+a][cy-h+b], pad outside, centre (floor(x+.5), floor(y+.5))  
// The flat bit order of the coverage plane equals the obs order, so FULL is a streaming bit -> element expansion with
// 128-bit stores, followed by a patch of the (few) anomalous cells that are still unmapped.
//
// Code -> element conversion (codes: 0 mapped -> 0, 1 normal -> 1, 2 pad -> -1 / 255, 3 anomalous -> U) is one PRMT per
// element: the four possible float32 values differ only in their two top bytes, so two 4-entry byte tables (`lt`, `ls`)
// indexed by the 2-bit code give the value.
#pragma once
#include "swarm_common.cuh"

namespace swarm {

template <typename OT> struct ObsCvt;
template <> struct ObsCvt<float> {
  uint32_t lt, ls;
  __device__ explicit ObsCvt(const DevParams& p) {
    const uint32_t fu = __float_as_uint((float)p.U);   // U <= 127: the low 16 bits of float(U) are zero
    // byte k of lt / ls = top / second byte of the value of code k:  0 -> 0x0000, 1 -> 0x3F80, 2 -> 0xBF80 (-1), 3 -> float(U)
    lt = 0x00u | (0x3Fu << 8) | (0xBFu << 16) | ((fu >> 24) << 24);
    ls = 0x00u | (0x80u << 8) | (0x80u << 16) | (((fu >> 16) & 0xFFu) << 24);
  }
  __device__ __forceinline__ float one(uint32_t code) const {   // code in [0,3]
    return __uint_as_float(__byte_perm(lt, ls, code * 0x1100u + 0x0400u));
  }
  // 4 cells given as a nibble of `unmapped` bits and a nibble of `anomalous-or-pad` bits -> one 128-bit value
  __device__ __forceinline__ float4 quad(uint32_t lo4, uint32_t hi4) const {
    return make_float4(one((lo4 & 1u) | ((hi4 & 1u) << 1)), one(((lo4 >> 1) & 1u) | (((hi4 >> 1) & 1u) << 1)),
                       one(((lo4 >> 2) & 1u) | (((hi4 >> 2) & 1u) << 1)), one(((lo4 >> 3) & 1u) | (((hi4 >> 3) & 1u) << 1)));
  }
  __device__ __forceinline__ void store4(float* dst, uint32_t lo4, uint32_t hi4) const {   // dst 16 B aligned
    __stcs(reinterpret_cast<float4*>(dst), quad(lo4, hi4));
  }
};
template <> struct ObsCvt<uint8_t> {
  uint32_t lut;
  __device__ explicit ObsCvt(const DevParams& p) : lut(0u | (1u << 8) | (255u << 16) | ((uint32_t)p.U << 24)) {}
  __device__ __forceinline__ uint8_t one(uint32_t code) const { return (uint8_t)__byte_perm(lut, 0u, (code & 3u) | 0x4440u); }
  __device__ __forceinline__ uint32_t quad(uint32_t lo4, uint32_t hi4) const {   // 4 codes -> 4 bytes
    const uint32_t c0 = (lo4 & 1u) | ((hi4 & 1u) << 1), c1 = ((lo4 >> 1) & 1u) | (((hi4 >> 1) & 1u) << 1);
    const uint32_t c2 = ((lo4 >> 2) & 1u) | (((hi4 >> 2) & 1u) << 1), c3 = ((lo4 >> 3) & 1u) | (((hi4 >> 3) & 1u) << 1);
    return __byte_perm(lut, 0u, c0 | (c1 << 4) | (c2 << 8) | (c3 << 12));
  }
  __device__ __forceinline__ void store4(uint8_t* dst, uint32_t lo4, uint32_t hi4) const {   // dst 4 B aligned
    __stcs(reinterpret_cast<uint32_t*>(dst), quad(lo4, hi4));
  }
};

// Packed 2-bit windows carry the codes themselves: nothing to convert (the FULL-mode emitters are never instantiated for it).
template <> struct ObsCvt<Packed2> {
  __device__ explicit ObsCvt(const DevParams&) {}
};

// ---- FULL obs of `nenv` consecutive envs whose coverage planes sit in shared memory (`sc`, env stride env_words words;
// `sa`: the anomaly planes in PLANE mode, else null) by `nthr` cooperating threads (a warp or a CTA).
// LIST mode: expand bits to 0 / 1, then — after a barrier supplied by the caller between the two calls — patch the listed
// cells that are still unmapped with U (emit_full_patch).
template <typename OT>
__device__ __forceinline__ void emit_full_bits(const DevParams& p, const ObsCvt<OT>& cvt, const uint32_t* sc, const uint32_t* sa, int nenv, int tid,
                                               int nthr, OT* __restrict__ o) {
  const int RC = p.RC;
  if ((RC & 3) == 0 && ((uintptr_t)o & 15) == 0) {
    const int nq = RC >> 2, total = nenv * nq;
#pragma unroll 2
    for (int v = tid; v < total; v += nthr) {
      const int e = p.nq_magic ? (int)__umulhi((uint32_t)v, p.nq_magic) : v;   // nq == 1: one quad per env
      const int qd = v - e * nq;
      const uint32_t sh = (uint32_t)(qd & 7) * 4u;
      const uint32_t lo = (sc[e * p.env_words + (qd >> 3)] >> sh) & 15u;
      const uint32_t hi = sa ? ((sa[e * p.env_words + (qd >> 3)] >> sh) & 15u & lo) : 0u;
      cvt.store4(o + (size_t)e * RC + 4 * qd, lo, hi);
    }
  } else {
    for (int e = 0; e < nenv; ++e)
      for (int c = tid; c < RC; c += nthr) {
        const uint32_t lo = (sc[e * p.env_words + (c >> 5)] >> (c & 31)) & 1u;
        const uint32_t hi = sa ? ((sa[e * p.env_words + (c >> 5)] >> (c & 31)) & 1u & lo) : 0u;
        o[(size_t)e * RC + c] = cvt.one(lo | (hi << 1));
      }
  }
}
// `sl`: the envs' anomaly lists (8 x uint32 per env) in shared or global memory
template <typename OT>
__device__ __forceinline__ void emit_full_patch(const DevParams& p, const ObsCvt<OT>& cvt, const uint32_t* sc, const uint32_t* sl, int nenv, int tid,
                                                int nthr, OT* __restrict__ o) {
  const int na = p.n_anom;
  for (int v = tid; v < nenv * na; v += nthr) {
    const int e = v / na, k = v - e * na;
    const uint32_t key = sl[e * 8 + k];
    const int x = (int)(key & 0xffffu), y = (int)(key >> 16);
    if (x >= p.R || y >= p.C) continue;   // LIST_EMPTY: an injected state with no anomalous cell to name (set_state_kernel)
    const int cell = y * p.R + x;
    if ((sc[e * p.env_words + (cell >> 5)] >> (cell & 31)) & 1u) o[(size_t)e * p.RC + cell] = cvt.one(CODE_ANOM);
  }
}

// ---- CTA-wide emission straight from the HBM-resident state (swarm_cta.cuh fallback for grids too large to stage) ------
// Reads go to L2 (__ldcg): the same kernel has just modified the planes / positions from another warp of the CTA.
template <typename OT>
__device__ __forceinline__ void emit_obs_global(const DevParams& p, int env, int tid, int nthr, OT* __restrict__ obs) {
  const ObsCvt<OT> cvt(p);
  if (p.obs_mode == 0) {
    OT* o = obs + (size_t)env * p.RC;
    for (int c = tid; c < p.RC; c += nthr) o[c] = cvt.one(cell_code_global(p, env, c));
  } else {
    const int W = p.W, WW = W * W, h = (W - 1) >> 1, nitem = p.D * WW;
    const float2* ps = p.pos + (size_t)env * p.D;
    OT* o = obs + (size_t)env * p.D * WW;
    for (int it = tid; it < nitem; it += nthr) {
      const int i = it / WW, r = it - i * WW, b = r / W, a = r - b * W;
      const float2 q = __ldcg(ps + i);
      const int gx = (int)floorf(q.x + 0.5f) - h + a, gy = (int)floorf(q.y + 0.5f) - h + b;
      uint32_t code = CODE_PAD;
      if ((unsigned)gx < (unsigned)p.R && (unsigned)gy < (unsigned)p.C) code = cell_code_global(p, env, gy * p.R + gx);
      o[it] = cvt.one(code);
    }
  }
}

}  // namespace swarm
