#!/usr/bin/env python
"""Where a render call's time goes outside its kernels: wall clock vs torch events vs the library's own CUDA-event
time, with and without per-stage events (flags bit1) and an L2 flush before the call; plus the small renders the host
loop dominates (cbox as shipped: 512x512, 4 spp; sponza at 8 spp) with their launch and iteration counts."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from lajolla_public_b200 import LajollaGPU, RenderParams, RNG_SAMPLE_STREAMS  # noqa: E402


def blob(name):
    for d in ("scenes_flat", "tests/golden"):
        p = os.path.join(ROOT, d, name + ".ljs")
        if os.path.exists(p):
            return p
    raise SystemExit(name + " missing")


def main():
    gpu = LajollaGPU(0)
    out = {}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
    for name, spp in (("sponza", 64), ("sponza", 8), ("cbox", 4), ("veach_mi", 16), ("matpreview", 64)):
        sc = gpu.load_scene(blob(name))
        sums = torch.zeros(sc.height * sc.width * 3, dtype=torch.float64, device="cuda:0")
        for flags in (0, 2):
            for do_flush in (False, True):
                rp = RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, flags=flags)
                sc.render_into_device(sums.data_ptr(), rp)
                walls, evs, libs = [], [], []
                for _ in range(5):
                    if do_flush:
                        flush.fill_(1)
                    torch.cuda.synchronize()
                    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                    t0 = time.perf_counter()
                    e0.record()
                    st = sc.render_into_device(sums.data_ptr(), rp)
                    e1.record(); torch.cuda.synchronize()
                    walls.append((time.perf_counter() - t0) * 1e3); evs.append(e0.elapsed_time(e1)); libs.append(st.render_ms)
                key = f"{name}_{spp}spp_flags{flags}_{'flush' if do_flush else 'noflush'}"
                out[key] = {"wall_ms": min(walls), "torch_event_ms": min(evs), "lib_ms": min(libs), "wall_all": walls,
                            "iterations": int(st.iterations), "launches": int(st.kernel_launches),
                            "stage_sum_ms": st.extend_ms + st.shade_ms + st.shadow_ms + st.generate_ms,
                            "msamples_per_s": sc.width * sc.height * spp / min(libs) * 1e-3}
                print(key, json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in out[key].items() if k != "wall_all"}), flush=True)
        sc.close()
    gpu.close()
    tag = os.environ.get("LJ_OVERHEAD_TAG", "")
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"step_overhead{tag}.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
