#!/usr/bin/env python
"""In-kernel timeline of the staged step kernel at the cfg3 shape (instrumented builds only):
  * CTA 0: per iteration and warp, cycles from loop top to "inputs landed", compute, release;
  * every CTA: globaltimer at start / end, SM id, cycles warp 1 spent waiting for the TMA input — who the stragglers are;
  * the per-step time of the instrumented build, eager and under CUDA-graph replay (not a bench value).
Build and run on the GPU box:
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -shared -Xcompiler -fPIC -ldl -DSWARM_TRACE=1 \
       -o marl-drones_b200/swarm_b200/libswarm_trace.so marl-drones_b200/csrc/swarm_api.cu
  SWARM_B200_SO=$PWD/marl-drones_b200/swarm_b200/libswarm_trace.so python tools/trace.py"""
import ctypes as C, os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "marl-drones_b200")); sys.path.insert(0, REPO)
import torch
from swarm_b200 import BatchedDroneEnv, _lib
envs = [BatchedDroneEnv(65536, row_count=64, col_count=64, n_drones=16, step_size=1.0, obs_mode="window", auto_reset=True, seed=1, env_id_base=s * 65536) for s in range(4)]
obs = [torch.empty(e.obs_shape, device="cuda") for e in envs]
half = torch.randint(0, 4, (16, 65536, 16), dtype=torch.uint8, device="cuda")
acts = torch.cat([half, (half + 2) % 4])
for rep in range(40):
    for k in range(32):
        envs[k % 4].step(acts[k], out=obs[k % 4])
torch.cuda.synchronize()
buf = np.zeros(16 * 8 * 8, np.uint64)
L = _lib.lib()
if not hasattr(L, "swarm_debug_trace"): sys.exit("this library was built without -DSWARM_TRACE=1 (see the docstring)")
L.swarm_debug_trace.argtypes = [C.c_void_p]
assert L.swarm_debug_trace(buf.ctypes.data) == 0
t = buf.reshape(16, 8, 8).astype(np.int64)
t0 = t[0, :, 0].min()
names = ["top", "prod", "b4wait", "ready", "b4win", "b4arr", "done"]
print("iteration rows; per warp: top->ready wait(cycles), compute ready->b4arr, total top->done")
for it in range(10):
    if t[it, 0, 0] == 0: break
    w = t[it, :, 3] - t[it, :, 2]; c = t[it, :, 5] - t[it, :, 3]; tot = t[it, :, 6] - t[it, :, 0]
    print(f"it{it}: start {t[it,:,0].min()-t0:7d}  prod(w0) {t[it,0,1]-t[it,0,0]:5d}  wait min/max {w.min():5d}/{w.max():5d}  compute min/max {c.min():5d}/{c.max():5d}  win {(t[it,:,5]-t[it,:,4]).mean():6.0f} (stage drain wait {(t[it,:,7]-t[it,:,4]).mean():5.0f})  total {tot.mean():6.0f}")
print("kernel span (cycles):", t[:, :, 6].max() - t0)
cb = np.zeros(4096, np.uint64); L.swarm_debug_cta.argtypes = [C.c_void_p]; assert L.swarm_debug_cta(cb.ctypes.data) == 0
smid = cb[3072:3072 + 444].astype(np.int64); waitc = cb[2048:2048 + 444].astype(np.int64); cb = cb.reshape(2048, 2)[:444].astype(np.int64); st = cb[:, 0] - cb[:, 0].min(); en = cb[:, 1] - cb[:, 0].min()
print("CTA start ns: min %d max %d | end ns: min %d median %d p90 %d max %d" % (st.min(), st.max(), en.min(), np.median(en), np.percentile(en, 90), en.max()))
slow = en > 1.3 * np.median(en); print("slow CTAs:", int(slow.sum()), "on SMs", sorted(smid[slow].tolist()))
print("CTAs per SM of slow SMs:", {int(m): int((smid == m).sum()) for m in set(smid[slow].tolist())})
print("input-wait cycles (warp 1, whole kernel): normal CTAs median %d, slow CTAs median %d (kernel ~%d cycles)" % (np.median(waitc[~slow]), np.median(waitc[slow]) if slow.any() else 0, int(np.median(en) * 1.965)))
late = np.argsort(en)[-8:]; print("latest CTAs", late.tolist(), en[late].tolist(), "their starts", st[late].tolist())
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
torch.cuda.synchronize()
ev[0].record()
for rep in range(8):
    for k in range(32):
        envs[k % 4].step(acts[k], out=obs[k % 4])
ev[1].record(); torch.cuda.synchronize()
print("eager back-to-back: %.1f us per step" % (ev[0].elapsed_time(ev[1]) * 1e3 / 256))
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for k in range(32):
        envs[k % 4].step(acts[k], out=obs[k % 4])
g.replay(); torch.cuda.synchronize()
ev[0].record()
for rep in range(8):
    g.replay()
ev[1].record(); torch.cuda.synchronize()
print("graph replay: %.1f us per step" % (ev[0].elapsed_time(ev[1]) * 1e3 / 256))
