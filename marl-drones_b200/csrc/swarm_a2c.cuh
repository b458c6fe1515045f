// swarm_a2c.cuh — the device steps of the batched A2C rollout loop that sit either side of the env step.
//   sample_actions_kernel   softmax + Categorical(probs).sample() of A2C/train.py:66-73 for every (env, drone) pair,
//                           written straight in the env's action layout uint8 [N][D]
//   nstep_returns_kernel    process_rollout's reverse scan, A2C/train.py:160-183:
//                               returns = rewards + returns * gamma * masks        (float32, this operation order, no FMA)
//                           for every (drone, env) column of a [T] rollout, plus the product of gamma*mask factors that
//                           the backward pass needs (d returns_t / d last_values).
//   gru_gates_fwd / bwd     the GRU cell of A2C/models.py:104 (`_, h = self.gru(x.unsqueeze(0))`, one step from h0 = 0) after
//                           the input-side GEMM, for all 2D stacked nets: ten element-wise torch kernels (and their
//                           autograd slices) folded into one pass each way.
// All are streaming kernels (bound by HBM); none is on the env-step path.
#pragma once
#include "swarm_common.cuh"

namespace swarm {

// logits float32 [D][N][A] (the stacked policy heads' output), actions uint8 [N][D].
// One thread per env: for every drone it reads the A logits (one 128-bit load when A == 4; consecutive threads read
// consecutive envs, so the loads coalesce), draws u from Philox4x32-10 keyed by (seed; global env id, step, drone / 4) and
// inverts the CDF of softmax(logits):  a = #{k : cum_k <= u * sum}  with cum_k = sum_{j<=k} exp(l_j - max).
template <int A>
__global__ void __launch_bounds__(256) sample_actions_kernel(const float* __restrict__ logits, int D, long long N, int Art, uint32_t seed_lo,
                                                             uint32_t seed_hi, uint32_t step, long long env_id_base,
                                                             uint8_t* __restrict__ actions) {
  const int na = A ? A : Art;
  for (long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x; n < N; n += (long long)gridDim.x * blockDim.x) {
    const unsigned long long gid = (unsigned long long)(env_id_base + n);
    uint4 r = make_uint4(0, 0, 0, 0);
    uint32_t word = 0;
    for (int d = 0; d < D; ++d) {
      if ((d & 3) == 0) r = philox4x32_10(make_uint4((uint32_t)gid, (uint32_t)(gid >> 32), step, (uint32_t)(d >> 2)), seed_lo, seed_hi);
      const uint32_t bits = (d & 3) == 0 ? r.x : ((d & 3) == 1 ? r.y : ((d & 3) == 2 ? r.z : r.w));
      const float u = (float)(bits >> 8) * (1.0f / 16777216.0f);   // [0, 1)
      const float* l = logits + ((size_t)d * (size_t)N + (size_t)n) * na;
      int a = 0;
      if constexpr (A == 4) {
        const float4 v = __ldcs(reinterpret_cast<const float4*>(l));
        const float m = fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w));
        const float e0 = expf(v.x - m), e1 = expf(v.y - m), e2 = expf(v.z - m), e3 = expf(v.w - m);
        const float c0 = e0, c1 = __fadd_rn(c0, e1), c2 = __fadd_rn(c1, e2), c3 = __fadd_rn(c2, e3);
        const float thr = __fmul_rn(u, c3);
        a = (c0 <= thr) + (c1 <= thr) + (c2 <= thr);
      } else {
        float m = l[0];
        for (int k = 1; k < na; ++k) m = fmaxf(m, l[k]);
        float sum = 0.f;
        for (int k = 0; k < na; ++k) sum = __fadd_rn(sum, expf(l[k] - m));
        const float thr = __fmul_rn(u, sum);
        float cum = 0.f;
        for (int k = 0; k < na - 1; ++k) {
          cum = __fadd_rn(cum, expf(l[k] - m));
          a += (cum <= thr);
        }
      }
      if ((D & 3) == 0) {   // four actions per 32-bit store (N*D is then a multiple of 4 and the row is 4 B aligned)
        word |= (uint32_t)a << (8 * (d & 3));
        if ((d & 3) == 3) {
          *reinterpret_cast<uint32_t*>(actions + (size_t)n * D + (d - 3)) = word;
          word = 0;
        }
      } else {
        actions[(size_t)n * D + d] = (uint8_t)a;
      }
    }
  }
}

// rewards float32 [T][N], dones uint8 [T][N] (mask = 1 - done, A2C/train.py:95-97), last_values float32 [D][N] (the
// bootstrap values, train.py:109-119), returns float32 [T][D][N], coef float32 [T][N] = prod_{s >= t} gamma * mask_s.
// One thread per (drone, env) column, env fastest: every access of a warp is a contiguous 128 B line.
__global__ void __launch_bounds__(256) nstep_returns_kernel(const float* __restrict__ rewards, const uint8_t* __restrict__ dones,
                                                            const float* __restrict__ last_values, int T, long long N, int D, float gamma,
                                                            float* __restrict__ returns, float* __restrict__ coef) {
  const long long total = (long long)D * N;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long d = i / N, n = i - d * N;
    float ret = last_values[i];
    float c = 1.0f;
    for (int t = T - 1; t >= 0; --t) {
      const float mask = 1.0f - (float)dones[(size_t)t * N + n];
      // train.py:174  returns = rewards + returns * args.gamma * masks   (left to right, each product rounded to float32)
      ret = __fadd_rn(rewards[(size_t)t * N + n], __fmul_rn(__fmul_rn(ret, gamma), mask));
      __stcs(returns + ((size_t)t * D + d) * N + n, ret);
      if (d == 0 && coef) {
        c = __fmul_rn(__fmul_rn(c, gamma), mask);
        coef[(size_t)t * N + n] = c;
      }
    }
  }
}

// GRU cell, one step from h0 = 0 (torch.nn.GRU gate order r, z, n; the hidden-side products vanish, bias_hh survives):
//   r = sigmoid(gi_r + bh_r)   z = sigmoid(gi_z + bh_z)   n = tanh(gi_n + r * bh_n)   h = (1 - z) * n
// gi float32 [N][G][3H] (x W_ih^T + b_ih, env-major as the GEMM leaves it), bhh float32 [G][3H], h float32 [G][N][H]
// (net-major, contiguous: the layout the per-net batched GEMMs downstream want).  One thread per (n, g, j), j fastest:
// reads and writes are contiguous runs of H floats.
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

__global__ void __launch_bounds__(256) gru_gates_fwd_kernel(const float* __restrict__ gi, const float* __restrict__ bhh, long long N, int G, int H,
                                                            float* __restrict__ h) {
  if ((H & 3) == 0) {   // four consecutive hidden units per thread: 128-bit accesses throughout
    const int H4 = H >> 2;
    const long long total4 = N * G * H4;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (long long)gridDim.x * blockDim.x) {
      const int j = (int)(i % H4) * 4;
      const long long ng = i / H4;
      const int g = (int)(ng % G);
      const long long n = ng / G;
      const float* gp = gi + ng * 3 * H + j;
      const float* bp = bhh + (size_t)g * 3 * H + j;
      const float4 gr = __ldcs(reinterpret_cast<const float4*>(gp)), gz = __ldcs(reinterpret_cast<const float4*>(gp + H));
      const float4 gn = __ldcs(reinterpret_cast<const float4*>(gp + 2 * H));
      const float4 br = __ldg(reinterpret_cast<const float4*>(bp)), bz = __ldg(reinterpret_cast<const float4*>(bp + H));
      const float4 bn = __ldg(reinterpret_cast<const float4*>(bp + 2 * H));
      float4 o;
      o.x = (1.0f - sigmoidf_(gz.x + bz.x)) * tanhf(gn.x + sigmoidf_(gr.x + br.x) * bn.x);
      o.y = (1.0f - sigmoidf_(gz.y + bz.y)) * tanhf(gn.y + sigmoidf_(gr.y + br.y) * bn.y);
      o.z = (1.0f - sigmoidf_(gz.z + bz.z)) * tanhf(gn.z + sigmoidf_(gr.z + br.z) * bn.z);
      o.w = (1.0f - sigmoidf_(gz.w + bz.w)) * tanhf(gn.w + sigmoidf_(gr.w + br.w) * bn.w);
      *reinterpret_cast<float4*>(h + ((size_t)g * N + n) * H + j) = o;
    }
    return;
  }
  const long long total = N * G * H;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(i % H);
    const long long ng = i / H;
    const int g = (int)(ng % G);
    const long long n = ng / G;
    const float* gp = gi + ng * 3 * H + j;
    const float* bp = bhh + (size_t)g * 3 * H + j;
    const float r = sigmoidf_(__ldcs(gp) + __ldg(bp));
    const float z = sigmoidf_(__ldcs(gp + H) + __ldg(bp + H));
    const float c = tanhf(__ldcs(gp + 2 * H) + r * __ldg(bp + 2 * H));
    h[((size_t)g * N + n) * H + j] = (1.0f - z) * c;
  }
}

// Backward of the above: dh [G][N][H] -> dgi [N][G][3H] (= the gradient of the pre-activations, which is also what b_ih and
// the r / z thirds of b_hh accumulate) and dnr [N][G][H] = d(pre_n) * r, whose sum over n is the gradient of b_hh's n third.
// The gates are recomputed from gi (cheaper than saving three activations per rollout step).
__device__ __forceinline__ void gru_gate_grad(float gr, float gz, float gn, float br, float bz, float bn, float d, float& dpr, float& dpz,
                                              float& dpn, float& dnr) {
  const float r = sigmoidf_(gr + br), z = sigmoidf_(gz + bz), c = tanhf(gn + r * bn);
  dpn = d * (1.0f - z) * (1.0f - c * c);     // d pre_n
  dpr = dpn * bn * r * (1.0f - r);            // d pre_r
  dpz = -d * c * z * (1.0f - z);              // d pre_z
  dnr = dpn * r;
}

__global__ void __launch_bounds__(256) gru_gates_bwd_kernel(const float* __restrict__ gi, const float* __restrict__ bhh, const float* __restrict__ dh,
                                                            long long N, int G, int H, float* __restrict__ dgi, float* __restrict__ dnr) {
  if ((H & 3) == 0) {
    const int H4 = H >> 2;
    const long long total4 = N * G * H4;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (long long)gridDim.x * blockDim.x) {
      const int j = (int)(i % H4) * 4;
      const long long ng = i / H4;
      const int g = (int)(ng % G);
      const long long n = ng / G;
      const float* gp = gi + ng * 3 * H + j;
      const float* bp = bhh + (size_t)g * 3 * H + j;
      const float4 gr = __ldcs(reinterpret_cast<const float4*>(gp)), gz = __ldcs(reinterpret_cast<const float4*>(gp + H));
      const float4 gn = __ldcs(reinterpret_cast<const float4*>(gp + 2 * H));
      const float4 br = __ldg(reinterpret_cast<const float4*>(bp)), bz = __ldg(reinterpret_cast<const float4*>(bp + H));
      const float4 bn = __ldg(reinterpret_cast<const float4*>(bp + 2 * H));
      const float4 d = __ldcs(reinterpret_cast<const float4*>(dh + ((size_t)g * N + n) * H + j));
      float4 o_r, o_z, o_n, o_a;
      gru_gate_grad(gr.x, gz.x, gn.x, br.x, bz.x, bn.x, d.x, o_r.x, o_z.x, o_n.x, o_a.x);
      gru_gate_grad(gr.y, gz.y, gn.y, br.y, bz.y, bn.y, d.y, o_r.y, o_z.y, o_n.y, o_a.y);
      gru_gate_grad(gr.z, gz.z, gn.z, br.z, bz.z, bn.z, d.z, o_r.z, o_z.z, o_n.z, o_a.z);
      gru_gate_grad(gr.w, gz.w, gn.w, br.w, bz.w, bn.w, d.w, o_r.w, o_z.w, o_n.w, o_a.w);
      float* op = dgi + ng * 3 * H + j;
      *reinterpret_cast<float4*>(op) = o_r;
      *reinterpret_cast<float4*>(op + H) = o_z;
      *reinterpret_cast<float4*>(op + 2 * H) = o_n;
      *reinterpret_cast<float4*>(dnr + ng * H + j) = o_a;
    }
    return;
  }
  const long long total = N * G * H;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(i % H);
    const long long ng = i / H;
    const int g = (int)(ng % G);
    const long long n = ng / G;
    const float* gp = gi + ng * 3 * H + j;
    const float* bp = bhh + (size_t)g * 3 * H + j;
    float* op = dgi + ng * 3 * H + j;
    gru_gate_grad(__ldcs(gp), __ldcs(gp + H), __ldcs(gp + 2 * H), __ldg(bp), __ldg(bp + H), __ldg(bp + 2 * H),
                  __ldcs(dh + ((size_t)g * N + n) * H + j), op[0], op[H], op[2 * H], dnr[i]);
  }
}

// y[g][n][:] += bias[g][:], optionally followed by relu, in place: the epilogue of the stacked linears (nn.Linear + F.relu of
// A2C/models.py:106-116) after a plain batched GEMM — one read and one write of y instead of a broadcast-bias fill, the
// GEMM's read of it, and a separate relu pass.
__global__ void __launch_bounds__(256) bias_act_kernel(float* __restrict__ y, const float* __restrict__ bias, int G, long long N, int F, int act) {
  if ((F & 3) == 0) {
    const int F4 = F >> 2;
    const long long per_net = N * F4, total = per_net * G;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
      const int g = (int)(i / per_net), f = (int)(i % F4) * 4;
      float4 v = reinterpret_cast<float4*>(y)[i];
      const float4 b = __ldg(reinterpret_cast<const float4*>(bias + (size_t)g * F + f));
      v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
      if (act == 1) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
      else if (act == 2) { v.x = tanhf(v.x); v.y = tanhf(v.y); v.z = tanhf(v.z); v.w = tanhf(v.w); }
      reinterpret_cast<float4*>(y)[i] = v;
    }
    return;
  }
  const long long per_net = N * F, total = per_net * G;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const float v = y[i] + __ldg(bias + (size_t)(i / per_net) * F + (i % F));
    y[i] = act == 1 ? fmaxf(v, 0.f) : (act == 2 ? tanhf(v) : v);
  }
}

// relu followed by F.normalize(dim=-1) (A2C/models.py:107-108: `F.relu(self.state_input(...))`, `F.normalize(...)`):
//   y = max(x, 0)   o = y / max(||y||_2, eps)
// x, o float32 [rows][F], one warp per row (a row is read twice; the second read hits L1).  F % 4 == 0 takes 128-bit accesses.
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(256) relu_l2norm_fwd_kernel(const float* __restrict__ x, long long rows, int F, float eps, float* __restrict__ o) {
  const int lane = threadIdx.x & 31;
  const long long w0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = w0; r < rows; r += nw) {
    const float* xr = x + (size_t)r * F;
    float* orow = o + (size_t)r * F;
    float ss = 0.f;
    if ((F & 3) == 0) {
      for (int k = lane * 4; k < F; k += 128) {
        const float4 v = *reinterpret_cast<const float4*>(xr + k);
        const float a = fmaxf(v.x, 0.f), b = fmaxf(v.y, 0.f), c = fmaxf(v.z, 0.f), d = fmaxf(v.w, 0.f);
        ss += a * a + b * b + c * c + d * d;
      }
    } else {
      for (int k = lane; k < F; k += 32) { const float a = fmaxf(xr[k], 0.f); ss += a * a; }
    }
    const float inv = 1.0f / fmaxf(sqrtf(warp_sum(ss)), eps);
    if ((F & 3) == 0) {
      for (int k = lane * 4; k < F; k += 128) {
        const float4 v = *reinterpret_cast<const float4*>(xr + k);
        __stcs(reinterpret_cast<float4*>(orow + k), make_float4(fmaxf(v.x, 0.f) * inv, fmaxf(v.y, 0.f) * inv, fmaxf(v.z, 0.f) * inv, fmaxf(v.w, 0.f) * inv));
      }
    } else {
      for (int k = lane; k < F; k += 32) orow[k] = fmaxf(xr[k], 0.f) * inv;
    }
  }
}

// Backward: with d = max(||y||, eps) and g = d loss / d o:
//   ||y|| >  eps:  dy = g / d - y * (g . y) / d^3        ||y|| <= eps:  dy = g / eps   (the clamp blocks the norm's path)
//   dx = dy where x > 0, else 0
__global__ void __launch_bounds__(256) relu_l2norm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ g, long long rows, int F, float eps,
                                                              float* __restrict__ dx) {
  const int lane = threadIdx.x & 31;
  const long long w0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = w0; r < rows; r += nw) {
    const float* xr = x + (size_t)r * F;
    const float* gr = g + (size_t)r * F;
    float* dr = dx + (size_t)r * F;
    float ss = 0.f, gy = 0.f;
    if ((F & 3) == 0) {
      for (int k = lane * 4; k < F; k += 128) {
        const float4 v = *reinterpret_cast<const float4*>(xr + k);
        const float4 u = *reinterpret_cast<const float4*>(gr + k);
        const float a = fmaxf(v.x, 0.f), b = fmaxf(v.y, 0.f), c = fmaxf(v.z, 0.f), d = fmaxf(v.w, 0.f);
        ss += a * a + b * b + c * c + d * d;
        gy += u.x * a + u.y * b + u.z * c + u.w * d;
      }
    } else {
      for (int k = lane; k < F; k += 32) { const float a = fmaxf(xr[k], 0.f); ss += a * a; gy += gr[k] * a; }
    }
    const float nrm = sqrtf(warp_sum(ss));
    gy = warp_sum(gy);
    const float d = fmaxf(nrm, eps), inv = 1.0f / d;
    const float c3 = nrm > eps ? gy * inv * inv * inv : 0.f;
    if ((F & 3) == 0) {
      for (int k = lane * 4; k < F; k += 128) {
        const float4 v = *reinterpret_cast<const float4*>(xr + k);
        const float4 u = *reinterpret_cast<const float4*>(gr + k);
        float4 t;
        t.x = v.x > 0.f ? u.x * inv - v.x * c3 : 0.f;
        t.y = v.y > 0.f ? u.y * inv - v.y * c3 : 0.f;
        t.z = v.z > 0.f ? u.z * inv - v.z * c3 : 0.f;
        t.w = v.w > 0.f ? u.w * inv - v.w * c3 : 0.f;
        __stcs(reinterpret_cast<float4*>(dr + k), t);
      }
    } else {
      for (int k = lane; k < F; k += 32) dr[k] = xr[k] > 0.f ? gr[k] * inv - xr[k] * c3 : 0.f;
    }
  }
}

// The deep-Q nets' embedding (DQL_model/DQLearning.py:43-46): out = F.normalize(cat(tanh(state_input(s)), tanh(drone_input(p))))
// for all stacked nets in one pass.  a float32 [N][D][fa], b float32 [N][D][fb] are the two shared input-layer products (bias
// included), out float32 [D][N][fa + fb] is the layout the stacked dense layers consume.  One warp per (env row, net);
// fa, fb multiples of 4, fa + fb <= 256 (tanh values stay in registers between the norm and the store).
// element k..k+3 of cat(a_row + bias_a, b_row + bias_b); the bias rows may be null (bias already inside a / b)
__device__ __forceinline__ float4 ld_cat4(const float* ar, const float* br, const float* ba, const float* bb, int fa, int k) {
  const bool first = k < fa;
  float4 v = *reinterpret_cast<const float4*>(first ? ar + k : br + (k - fa));
  const float* bp = first ? ba : bb;
  if (bp) {
    const float4 c = __ldg(reinterpret_cast<const float4*>(bp + (first ? k : k - fa)));
    v.x += c.x; v.y += c.y; v.z += c.z; v.w += c.w;
  }
  return v;
}

__global__ void __launch_bounds__(256) embed_tanh_l2norm_fwd_kernel(const float* __restrict__ a, const float* __restrict__ b, const float* __restrict__ bias_a,
                                                                    const float* __restrict__ bias_b, long long N, int D, int fa, int fb,
                                                                    float eps, float* __restrict__ out) {
  const int lane = threadIdx.x & 31, F = fa + fb;
  const long long rows = N * D;
  const long long w0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = w0; r < rows; r += nw) {
    const float* ar = a + (size_t)r * fa;
    const float* br = b + (size_t)r * fb;
    const int net = (int)(r % D);
    const float* ba = bias_a ? bias_a + (size_t)net * fa : nullptr;
    const float* bb = bias_b ? bias_b + (size_t)net * fb : nullptr;
    float* orow = out + ((size_t)net * N + (size_t)(r / D)) * F;
    float4 y[2];
    float ss = 0.f;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int k = lane * 4 + j * 128;
      if (k < F) {
        const float4 v = ld_cat4(ar, br, ba, bb, fa, k);
        y[j] = make_float4(tanhf(v.x), tanhf(v.y), tanhf(v.z), tanhf(v.w));
        ss += y[j].x * y[j].x + y[j].y * y[j].y + y[j].z * y[j].z + y[j].w * y[j].w;
      }
    }
    const float inv = 1.0f / fmaxf(sqrtf(warp_sum(ss)), eps);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int k = lane * 4 + j * 128;
      if (k < F) *reinterpret_cast<float4*>(orow + k) = make_float4(y[j].x * inv, y[j].y * inv, y[j].z * inv, y[j].w * inv);
    }
  }
}

// Backward: y = tanh(pre) recomputed, d = max(||y||, eps), g = d loss / d out [D][N][F]:
//   dy = g / d - y * (g . y) / d^3 (second term only when ||y|| > eps);   dpre = dy * (1 - y^2)  ->  da [N][D][fa], db [N][D][fb]
__global__ void __launch_bounds__(256) embed_tanh_l2norm_bwd_kernel(const float* __restrict__ a, const float* __restrict__ b, const float* __restrict__ bias_a,
                                                                    const float* __restrict__ bias_b, const float* __restrict__ g, long long N, int D, int fa, int fb, float eps, float* __restrict__ da, float* __restrict__ db) {
  const int lane = threadIdx.x & 31, F = fa + fb;
  const long long rows = N * D;
  const long long w0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = w0; r < rows; r += nw) {
    const float* ar = a + (size_t)r * fa;
    const float* br = b + (size_t)r * fb;
    const int net = (int)(r % D);
    const float* ba = bias_a ? bias_a + (size_t)net * fa : nullptr;
    const float* bb = bias_b ? bias_b + (size_t)net * fb : nullptr;
    const float* gr = g + ((size_t)net * N + (size_t)(r / D)) * F;
    float4 y[2], u[2];
    float ss = 0.f, gy = 0.f;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int k = lane * 4 + j * 128;
      if (k < F) {
        const float4 v = ld_cat4(ar, br, ba, bb, fa, k);
        u[j] = *reinterpret_cast<const float4*>(gr + k);
        y[j] = make_float4(tanhf(v.x), tanhf(v.y), tanhf(v.z), tanhf(v.w));
        ss += y[j].x * y[j].x + y[j].y * y[j].y + y[j].z * y[j].z + y[j].w * y[j].w;
        gy += u[j].x * y[j].x + u[j].y * y[j].y + u[j].z * y[j].z + u[j].w * y[j].w;
      }
    }
    const float nrm = sqrtf(warp_sum(ss));
    gy = warp_sum(gy);
    const float d = fmaxf(nrm, eps), inv = 1.0f / d;
    const float c3 = nrm > eps ? gy * inv * inv * inv : 0.f;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int k = lane * 4 + j * 128;
      if (k < F) {
        float4 t;
        t.x = (u[j].x * inv - y[j].x * c3) * (1.f - y[j].x * y[j].x);
        t.y = (u[j].y * inv - y[j].y * c3) * (1.f - y[j].y * y[j].y);
        t.z = (u[j].z * inv - y[j].z * c3) * (1.f - y[j].z * y[j].z);
        t.w = (u[j].w * inv - y[j].w * c3) * (1.f - y[j].w * y[j].w);
        float* dst = k < fa ? da + (size_t)r * fa + k : db + (size_t)r * fb + (k - fa);
        *reinterpret_cast<float4*>(dst) = t;
      }
    }
  }
}

// (x - mean(x)) / std(x) per row, unbiased std (torch.Tensor.std()): the input normalisation of both policy families at batch 1
// (A2C/models.py:101-103, DQL_model/DQLearning.py:40-42).  x, out float32 [rows][F]; one warp per row, three passes over a row
// that stays in L1.  F == 1 gives NaN, as torch does.
__global__ void __launch_bounds__(256) row_standardize_kernel(const float* __restrict__ x, long long rows, int F, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const long long w0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = w0; r < rows; r += nw) {
    const float* xr = x + (size_t)r * F;
    float* orow = out + (size_t)r * F;
    float sum = 0.f;
    for (int k = lane; k < F; k += 32) sum += xr[k];
    const float mean = warp_sum(sum) / (float)F;
    float ss = 0.f;
    for (int k = lane; k < F; k += 32) { const float d = xr[k] - mean; ss += d * d; }
    const float sd = sqrtf(warp_sum(ss) / (float)(F - 1));
    for (int k = lane; k < F; k += 32) orow[k] = (xr[k] - mean) / sd;
  }
}

}  // namespace swarm
