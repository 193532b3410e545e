"""GPU debugging helper: wall/device time of consecutive render calls under bench-like conditions."""
import os, sys, time, subprocess
import numpy as np
import gpuspectral_b200 as g

W, H, S = 1920, 1080, 32
mode = sys.argv[1] if len(sys.argv) > 1 else "plain"
if "torch" in mode:
    import torch
    torch.cuda.set_device(0)
    torch.zeros(1, device="cuda")
proc = None
if "smi" in mode:
    proc = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-lms", "100", "-i", "0"],
                            stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
scene = g.Scene.builtin("cornell")
with g.Context(0) as ctx:
    ctx.upload(scene)
    ctx.set_params(max_depth=8)
    ctx.render(W, H, 0, 0)
    ctx.set_profiling("prof" in mode)
    ctx.clear()
    out = []
    for k in range(6):
        t0 = time.time()
        ctx.render(W, H, k * S, S)
        ctx.synchronize()
        out.append(((time.time() - t0) * 1e3, ctx.last_render_ms()["total"]))
    print(mode, " ".join(f"{w:.0f}/{d:.0f}" for w, d in out))
if proc:
    proc.terminate()
