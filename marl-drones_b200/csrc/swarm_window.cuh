// swarm_window.cuh — stage (4), WINDOW mode: the W x W local observation of one drone from a coverage plane
// staged in shared memory.  Shared by the warp kernel (swarm_warp.cuh) and the CTA kernel (swarm_cta.cuh).
#pragma once
#include "swarm_common.cuh"
#include "swarm_emit.cuh"

namespace swarm {

// ---- window helpers (W == 5 fast path) --------------------------------------------------------------------------------
// Anomalous cells of the env that fall inside the 5x5 window centred on (cx, cy): bit dy*5+dx.  `sl_s` = shared address of
// the env's list (8 x uint32, x | y << 16); n (<= 8) is warp-uniform.  Both coordinates are tested with one packed
// subtraction: u = e + 0x00020002 - (cx | cy << 16) holds dx in its low and dy in its high half (a borrow out of the low
// half can only happen when dx < 0, i.e. when the entry is rejected anyway); v = 0x00040004 - u; inside <=> no half of u or
// v is negative.  bit = (u * 0x00010005) >> 16 = dx + 5 dy.
__device__ __forceinline__ uint32_t anomaly_bit5(uint32_t e, uint32_t cxy) {
  const uint32_t u = e + 0x00020002u - cxy, v = 0x00040004u - u;
  return ((u | v) & 0x80008000u) ? 0u : (1u << ((u * 0x00010005u) >> 16));
}
__device__ __forceinline__ uint32_t window_anomalies5(uint32_t sl_s, int n, int cx, int cy) {
  if (n == 0) return 0u;
  const uint32_t cxy = (uint32_t)cx | ((uint32_t)cy << 16);
  uint4 a, b = make_uint4(0u, 0u, 0u, 0u);
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "r"(sl_s));
  uint32_t A = anomaly_bit5(a.x, cxy);
  if (n > 1) A |= anomaly_bit5(a.y, cxy);
  if (n > 2) A |= anomaly_bit5(a.z, cxy);
  if (n > 3) A |= anomaly_bit5(a.w, cxy);
  if (n > 4) {
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4+16];" : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "r"(sl_s));
    A |= anomaly_bit5(b.x, cxy);
    if (n > 5) A |= anomaly_bit5(b.y, cxy);
    if (n > 6) A |= anomaly_bit5(b.z, cxy);
    if (n > 7) A |= anomaly_bit5(b.w, cxy);
  }
  return A;
}

// The drone's 5x5 window as float / uint8 values in st[0 .. 25): U = unmapped bits gathered from the coverage plane `sc_s`
// (32-bit shared address of the env's plane; rows outside the grid read neighbouring shared memory, which the pad mask
// overrides), P = pad mask from the per-coordinate tables, A = anomalies (window_anomalies5).  Codes (swarm_common.cuh):
// lo = unmapped & !pad, hi = (anomalous & unmapped) | pad.  `lut_s` = shared address of the 256-entry LUT indexed lo4 | hi4 << 4.
template <typename OT>
__device__ __forceinline__ void build_window5(int R, uint32_t sc_s, uint32_t lut_s, uint32_t P, uint32_t A, int wx0, int wy0, OT* st) {
  const int sb = wy0 * R + wx0;   // bit offset of the window's first row in the plane (may be negative)
  uint32_t w0[5], w1[5];
  uint32_t U = 0u;
  asm volatile("" ::: "memory");   // the plane reads below stay behind the __syncwarp that publishes this step's clears
  if ((R & 31) == 0) {   // rows are whole words: one address, one shift for all five rows (warp-uniform branch)
    const uint32_t a = sc_s + (uint32_t)((sb >> 5) << 2), rb = (uint32_t)R >> 3;
#pragma unroll
    for (int b = 0; b < 5; ++b) {   // ten independent loads, issued back to back
      asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w0[b]) : "r"(a + b * rb));
      asm volatile("ld.shared.b32 %0, [%1+4];" : "=r"(w1[b]) : "r"(a + b * rb));
    }
#pragma unroll
    for (int b = 0; b < 5; ++b) U |= (__funnelshift_r(w0[b], w1[b], (uint32_t)sb) & 31u) << (5 * b);
  } else {
#pragma unroll
    for (int b = 0; b < 5; ++b) {
      const uint32_t a = sc_s + (uint32_t)(((sb + b * R) >> 5) << 2);
      asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w0[b]) : "r"(a));
      asm volatile("ld.shared.b32 %0, [%1+4];" : "=r"(w1[b]) : "r"(a));
    }
#pragma unroll
    for (int b = 0; b < 5; ++b) U |= (__funnelshift_r(w0[b], w1[b], (uint32_t)(sb + b * R)) & 31u) << (5 * b);
  }
  const uint32_t LO = U & ~P, HI = (A & U) | P;
  if constexpr (sizeof(OT) == 8) {   // packed 2-bit format: the two planes ARE the observation
    *st = (OT)LO | ((OT)HI << 32);
  } else {
  uint32_t off[7];
#pragma unroll
  for (int j = 0; j < 7; ++j) {
    const uint32_t lo = (j == 0 ? (LO << 4) : (LO >> (4 * j - 4))) & 0xf0u;
    const uint32_t hi = (j < 2 ? (HI << (8 - 4 * j)) : (HI >> (4 * j - 8))) & 0xf00u;
    off[j] = lo + hi;   // 16 * (lo4 | hi4 << 4)
  }
  if constexpr (sizeof(OT) == 4) {
    float4 v[6];
    float v6;
#pragma unroll
    for (int j = 0; j < 6; ++j)   // the LUT is immutable after the prologue: seven independent loads, then 25 stores
      asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v[j].x), "=f"(v[j].y), "=f"(v[j].z), "=f"(v[j].w) : "r"(lut_s + off[j]));
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v6) : "r"(lut_s + off[6]));
    float* o = reinterpret_cast<float*>(st);
#pragma unroll
    for (int j = 0; j < 6; ++j) { o[4 * j] = v[j].x; o[4 * j + 1] = v[j].y; o[4 * j + 2] = v[j].z; o[4 * j + 3] = v[j].w; }
    o[24] = v6;
  } else {
    uint32_t v[7];
#pragma unroll
    for (int j = 0; j < 7; ++j) asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v[j]) : "r"(lut_s + (off[j] >> 2)));
    uint8_t* o = reinterpret_cast<uint8_t*>(st);
#pragma unroll
    for (int j = 0; j < 6; ++j) { o[4 * j] = (uint8_t)v[j]; o[4 * j + 1] = (uint8_t)(v[j] >> 8); o[4 * j + 2] = (uint8_t)(v[j] >> 16); o[4 * j + 3] = (uint8_t)(v[j] >> 24); }
    o[24] = (uint8_t)v[6];
  }
  }   // not packed
}

// Generic W (any odd size): per-cell, bounds checked.  `sc`: the env's plane, `sl`: its list (shared memory).
template <typename OT>
__device__ __forceinline__ void build_window_generic(const DevParams& p, const ObsCvt<OT>& cvt, const uint32_t* sc, const uint32_t* sl, int cx, int cy,
                                                     OT* st) {
  const int W = p.W, h = (W - 1) >> 1;
  for (int b = 0; b < W; ++b)
    for (int a = 0; a < W; ++a) {
      const int gx = cx - h + a, gy = cy - h + b;
      uint32_t code = CODE_PAD;
      if ((unsigned)gx < (unsigned)p.R && (unsigned)gy < (unsigned)p.C) {
        const int cell = gy * p.R + gx;
        code = (sc[cell >> 5] >> (cell & 31)) & 1u;
        if (code) {
          const uint32_t key = (uint32_t)gx | ((uint32_t)gy << 16);
          for (int k = 0; k < p.n_anom; ++k)
            if (sl[k] == key) code = CODE_ANOM;
        }
      }
      st[b * W + a] = cvt.one(code);
    }
}

// The tables of the W == 5 fast path, built once per CTA: `lut` = 256 entries (lo4 | hi4 << 4 -> four cells as float4 / four
// bytes; `n_lut` = 16 builds the FULL-mode nibble table only), padx / pady = pad masks by centre column / row.
template <typename OT>
__device__ __forceinline__ void build_window_tables(const DevParams& p, const ObsCvt<OT>& cvt, uint32_t* lut, int n_lut, uint32_t* padx, uint32_t* pady,
                                                    int tid, int nthr) {
  if constexpr (sizeof(OT) != 8)   // (the packed format needs no value table)
    for (int b = tid; b < n_lut; b += nthr) {
      if constexpr (sizeof(OT) == 4) reinterpret_cast<float4*>(lut)[b] = cvt.quad(b & 15u, b >> 4);
      else lut[b] = cvt.quad(b & 15u, b >> 4);
    }
  if (padx)
    for (int c = tid; c < 256; c += nthr) {
      uint32_t mx = 0u, my = 0u;
#pragma unroll
      for (int a = 0; a < 5; ++a) {
        if ((unsigned)(c - 2 + a) >= (unsigned)p.R) mx |= 0x108421u << a;      // column a of every row
        if ((unsigned)(c - 2 + a) >= (unsigned)p.C) my |= 31u << (5 * a);       // row a
      }
      padx[c] = mx;
      pady[c] = my;
    }
}

}  // namespace swarm
