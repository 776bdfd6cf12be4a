#!/usr/bin/env python
"""bench.py -- headline benchmark of the render hot path (BASELINE.json: Mrays/s and
Msamples/s on the sponza path-tracing render, 768x575, per B200 and for N GPUs).

One "step" = one render pass of SPP_PER_STEP samples per pixel of scenes/sponza/sponza.xml
(the XML's own film size and integrator settings) through the C ABI.
  value : device-resident throughput (scene + BVH already in HBM, CUDA-event timed, max over ranks)
  e2e   : the same metric through the host-buffer call a user of render() makes: per step the
          flattened scene is uploaded from host memory (H2D), the BVH is built, the pass is
          rendered and the f32 image is read back (D2H)
  --impl reference : the reference's own CPU implementation (oracle/_ref/ref_driver = unmodified
          reference sources + BVH shim; the Embree binary is missing from the reference tree)
          on all host cores, on a bounded spp sample of the same workload.
N > 1 (torchrun): the scene is replicated, rank r renders the spp slice [r*S, (r+1)*S) of every
pixel (weak scaling: per-GPU work fixed) and the f64 radiance sums are all-reduced over NCCL.
The same line carries a `strong` block: FIXED total work (64 and 1024 spp of the same scene) split over the N GPUs by
the library's own multi-device context (lj_init_devices), with the 1-GPU time of the same workload beside it.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SCENE_NAME = "sponza"
SPP_PER_STEP = 64           # the XML's sampleCount
BYTES_RAY_FIXED = 32 + 20   # ray (8 x f32) + closest-hit record


def find_blob():
    for d in ("scenes_flat", os.path.join("tests", "golden")):
        p = os.path.join(ROOT, d, SCENE_NAME + ".ljs")
        if os.path.exists(p):
            return p
    raise SystemExit(f"{SCENE_NAME}.ljs not found: run tools/make_goldens.py in the authoring container")


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons sampled DURING the timed region (NVML; nvidia-smi as a fallback)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.sm, self.mx, self.reasons, self.stop_flag = index, [], [], set(), False
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
        except Exception:
            self.nvml = None

    def sample_nvml(self):
        n = self.nvml
        self.sm.append(int(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
        self.mx.append(int(n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)))
        r = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
            else int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
        for name, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40)):
            if r & bit:
                self.reasons.add(name)

    def sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                             stdout=subprocess.PIPE, text=True, timeout=5).stdout.strip()
        f = [x.strip() for x in out.split(",")]
        if f and f[0].isdigit():
            self.sm.append(int(f[0]))
        if len(f) > 1 and f[1].isdigit():
            self.mx.append(int(f[1]))
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for i in range(4):
            if len(f) >= 7 and f[3 + i].lower().startswith("active"):
                self.reasons.add(names[i])

    def run(self):
        while not self.stop_flag:
            try:
                self.sample_nvml() if self.nvml else self.sample_smi()
            except Exception:
                pass
            time.sleep(0.02 if self.nvml else 0.2)

    def summary(self):
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": sorted(self.reasons), "samples": len(sm), "source": "nvml" if self.nvml else "nvidia-smi"}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference_sample(spp, threads):
    """Time the oracle (reference code + BVH shim) on sponza at `spp` samples per pixel."""
    drv = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
    xml = os.path.join(ROOT, "oracle", "_ref", "scenes", "sponza", "sponza.xml")
    if not (os.path.exists(drv) and os.path.exists(xml)):
        return None
    out = subprocess.run([drv, "render", xml, "-o", "/tmp/lj_bench_ref.pfm", "--mode", "A", "--spp", str(spp), "-t", str(threads)],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    lines = [l for l in out.splitlines() if l.startswith("{")]
    return json.loads(lines[-1]) if lines else None


def reference_arm(args):
    """`--impl reference`: rank 0 only; each step = one bounded-spp render on all host cores."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = host_cores()
    probe = run_reference_sample(1, cores)
    if probe is None:
        emit(({"impl": "reference", "unavailable": "oracle/_ref/ref_driver or its scene copy is missing (build() in the authoring container creates both)"}))
        return
    # size each step to ~8 s of wall time
    spp = max(1, min(SPP_PER_STEP, int(round(8.0 / max(probe["render_s"], 1e-3)))))
    for _ in range(args.warmup if args.warmup < 2 else 1):
        run_reference_sample(spp, cores)
    rays = samples = secs = 0.0
    for _ in range(args.steps):
        r = run_reference_sample(spp, cores)
        rays += r["intersect_rays"] + r["occluded_rays"]; samples += r["samples"]; secs += r["render_s"]
    v = rays / secs * 1e-6
    line = {"impl": "reference", "metric": "sponza_path_tracing_mrays_per_s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "scenes/sponza/sponza.xml (reference asset), CPU",
            "msamples_per_s": samples / secs * 1e-6,
            "config": {"workload": f"sponza.xml 768x575 path integrator, {spp} spp per step (bounded sample of the 64-spp pass)"},
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": "reference",
                             "sample": f"{spp} spp of sponza 768x575 per step, reference sources + BVH shim (not Embree)"},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def gpu_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from lajolla_public_b200 import LajollaGPU, RenderParams, RNG_SAMPLE_STREAMS, ENGINE_WAVEFRONT, ENGINE_MEGAKERNEL
    from lajolla_public_b200.api import HostSceneDesc

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the render path has no CPU fallback")
    torch.cuda.set_device(local)
    cpu_group = None
    if world > 1:
        # (NCCL_DEBUG is left as the launcher set it: main() has already moved fd 1 to stderr, so NCCL's banner and
        #  rank lines cannot reach the JSON line)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        cpu_group = dist.new_group(backend="gloo")      # host-side barrier that leaves the GPUs idle (strong block)
    dev = torch.device("cuda", local)

    blob = find_blob()
    gpu = LajollaGPU(local)
    desc = HostSceneDesc(blob, pin=True)
    scene = gpu.scene_from_desc(desc)
    w, h = scene.width, scene.height
    S = args.spp
    engine = ENGINE_MEGAKERNEL if args.engine == "megakernel" else ENGINE_WAVEFRONT
    # flags bit1: CUDA-event intervals around every stage launch (the roofline's kernel time)
    rp = RenderParams(spp=S * world, rng_mode=RNG_SAMPLE_STREAMS, engine=engine, sample_begin=rank * S, sample_end=(rank + 1) * S, flags=2)
    sums = torch.zeros(h * w * 3, dtype=torch.float64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step():
        st = scene.render_into_device(sums.data_ptr(), rp)
        if world > 1:
            dist.all_reduce(sums)
        return st

    for _ in range(args.warmup):
        one_step()
    sampler = ClockSampler(local); sampler.start()
    barrier()
    total_ms = 0.0; rays = samples = launches = iters = 0
    rtypes = {"primary": 0, "extension": 0, "shadow": 0}
    stage = {"extend_ms": 0.0, "shadow_ms": 0.0, "shade_ms": 0.0, "generate_ms": 0.0}
    for _ in range(args.steps):
        flush.fill_(1)                       # L2 flush between timed iterations (not timed)
        barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        st = one_step()
        e1.record(); torch.cuda.synchronize()
        total_ms += e0.elapsed_time(e1)
        rays += st.rays; samples += st.samples; launches += st.kernel_launches; iters += st.iterations
        rtypes["primary"] += st.rays_primary; rtypes["extension"] += st.rays_extension; rtypes["shadow"] += st.rays_shadow
        for k in stage:
            stage[k] += getattr(st, k)
    barrier()
    sampler.stop_flag = True; sampler.join()
    t = torch.tensor([total_ms, float(rays), float(samples)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t)                   # sums of rays / samples over ranks
        total_ms = float(tmax[0])
    all_rays, all_samples = float(t[1]), float(t[2])
    value = all_rays / (total_ms * 1e-3) * 1e-6

    # ---- end to end: host scene -> upload -> BVH build -> render -> host image, every step ----
    # (the call a user makes: no per-stage timing flag, so the library may overlap its stages)
    rp_user = RenderParams(spp=S * world, rng_mode=RNG_SAMPLE_STREAMS, engine=engine, sample_begin=rank * S, sample_end=(rank + 1) * S)
    img = np.empty((h, w, 3), dtype=np.float32)
    e2e_s = 0.0; e2e_rays = 0
    barrier()
    for i in range(max(1, min(args.steps, 3)) + 1):
        barrier()
        t0 = time.perf_counter()
        sc2 = gpu.scene_from_desc(desc)
        if world > 1:
            st2 = sc2.render_into_device(sums.data_ptr(), rp_user); dist.all_reduce(sums)
            img_t = (sums / float(S * world)).to(torch.float32).cpu()
        else:
            _, st2 = sc2.render(rp_user, out=img)
        sc2.close()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if i > 0:                            # first pass = warm-up
            e2e_s += dt; e2e_rays += st2.rays
    te = torch.tensor([e2e_s, float(e2e_rays)], dtype=torch.float64, device=dev)
    if world > 1:
        tm = te.clone(); dist.all_reduce(tm, op=dist.ReduceOp.MAX); dist.all_reduce(te)
        e2e_s = float(tm[0])
    e2e_value = float(te[1]) / e2e_s * 1e-6

    # ---- fixed-work (strong) scaling through the library's own multi-device context, rank 0 drives all N GPUs ----
    strong = None
    torch.cuda.synchronize()
    if cpu_group is not None:
        dist.barrier(group=cpu_group)
    if rank == 0 and not args.no_strong:
        strong = strong_block(gpu, scene, desc, world, args)
    if cpu_group is not None:
        dist.barrier(group=cpu_group)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        roof = roofline_block(scene, rp, stage, rtypes, iters, peak, "measured" if peaks else "fallback", sums.data_ptr())
        line = {"metric": "sponza_path_tracing_mrays_per_s", "value": value, "unit": "Mrays/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "scenes/sponza/sponza.xml (reference asset, flattened blob)",
                "msamples_per_s": all_samples / (total_ms * 1e-3) * 1e-6,
                "config": {"workload": f"sponza.xml 768x575 path integrator (maxDepth -1, rr 5), {S} spp per GPU per step",
                           "engine": args.engine, "rng": "per-(pixel,sample) pcg32 streams", "shading": "f64", "traversal": "f32",
                           "l2": "256 MiB buffer written between timed steps (L2 flush)", "partition": f"spp slices x{world}, NCCL all-reduce of f64 sums"},
                "clocks": sampler.summary(), "gpu_launches": int(launches),
                "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": desc.nbytes, "d2h_bytes_per_step": h * w * 3 * 4},
                "roofline": roof, "strong": strong}
        if world == 1 and not args.no_cpu_baseline:
            cores = host_cores()
            probe = run_reference_sample(1, cores)
            if probe:
                spp = max(1, min(SPP_PER_STEP, int(round(12.0 / max(probe["render_s"], 1e-3)))))
                r = run_reference_sample(spp, cores) if spp > 1 else probe
                line["cpu_baseline"] = {"value": (r["intersect_rays"] + r["occluded_rays"]) / r["render_s"] * 1e-6, "unit": "Mrays/s",
                                        "cores": cores, "kind": "reference", "msamples_per_s": r["msamples_per_s"],
                                        "sample": f"{spp} spp of sponza 768x575 ({r['render_s']:.1f} s), reference sources + BVH shim (not Embree)"}
        emit(line)
    scene.close(); gpu.close()
    if world > 1:
        dist.destroy_process_group()


def strong_block(gpu1, scene1, desc, n, args):
    """Fixed TOTAL work split over the N GPUs of the box by the library itself (lj_init_devices: per-device host threads,
    tile or sample-slice partition, one ncclReduce -- lajolla_public_b200/csrc/lj_multi.cu): sponza at the shipped 64 spp
    and at 1024 spp.  Every entry: wall-clock ms of the synchronous lj_render_device call (best of `reps`), the device
    time the library reports (slowest device's CUDA-event time + the reduce), the 1-GPU time of the same workload
    measured in this run, and e2e_ms = scene upload + BVH build on every device + render + host image.
    image_check = max |N-way f32 image - 1-GPU f32 image| on identical streams (0 for tiles by construction)."""
    import numpy as np
    import torch
    from lajolla_public_b200 import LajollaGPU, RenderParams, RNG_SAMPLE_STREAMS, PARTITION_TILES, PARTITION_SPP
    w, h = scene1.width, scene1.height
    sums = torch.zeros(h * w * 3, dtype=torch.float64, device="cuda:0")
    out = {"n_gpus": n, "driver": "lj_init_devices (one process, one host thread per GPU, ncclReduce)", "runs": []}
    multi = LajollaGPU(list(range(n))) if n > 1 else gpu1
    sceneN = multi.scene_from_desc(desc) if n > 1 else scene1
    try:
        for total_spp, reps in ((64, 3), (args.strong_spp, 1)):
            def timed(sc, rp, reps=reps):
                best, st = None, None
                for _ in range(reps):
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    st = sc.render_into_device(sums.data_ptr(), rp)
                    dt = (time.perf_counter() - t0) * 1e3
                    best = dt if best is None else min(best, dt)
                return best, st
            rp1 = RenderParams(spp=total_spp, rng_mode=RNG_SAMPLE_STREAMS)
            timed(scene1, rp1, 1)                                  # warm the pool for this spp
            ms1, st1 = timed(scene1, rp1)
            ref_img, _ = scene1.render(rp1)
            for part, pname in ((PARTITION_TILES, "tiles"), (PARTITION_SPP, "spp")):
                if n == 1 and part == PARTITION_SPP:
                    continue
                rp = RenderParams(spp=total_spp, rng_mode=RNG_SAMPLE_STREAMS, partition=part)
                timed(sceneN, rp, 1)
                ms, st = timed(sceneN, rp)
                img, _ = sceneN.render(rp)
                t0 = time.perf_counter()
                sc2 = multi.scene_from_desc(desc)
                img2, st2 = sc2.render(rp)
                sc2.close()
                e2e_ms = (time.perf_counter() - t0) * 1e3
                out["runs"].append({"total_spp": total_spp, "partition": pname if n > 1 else "none", "ms": ms, "device_ms": st.render_ms,
                                    "reduce_ms": st.reduce_ms, "device_ms_min": st.device_ms_min, "device_ms_max": st.device_ms_max,
                                    "one_gpu_ms": ms1, "speedup": ms1 / ms, "efficiency": ms1 / ms / n,
                                    "mrays_per_s": st.rays / ms * 1e-3, "msamples_per_s": w * h * total_spp / ms * 1e-3,
                                    "e2e_ms": e2e_ms, "bvh_build_ms": st2.bvh_build_ms, "upload_ms": st2.upload_ms,
                                    "iterations": int(st.iterations), "launches": int(st.kernel_launches),
                                    "rays": int(st.rays), "rays_one_gpu": int(st1.rays),
                                    "image_check": float(np.abs(img.astype(np.float64) - ref_img).max()),
                                    "image_bit_identical": bool(np.array_equal(img.view(np.uint32), ref_img.view(np.uint32)))})
    finally:
        if n > 1:
            sceneN.close(); multi.close()
    return out


def roofline_block(scene, rp, stage, rays_by_type, iterations, peak, which, scratch_ptr):
    """Roofline of the dominant kernel, k_wf_extend (closest-hit traversal of the 8-wide BVH).
    ALGORITHMIC bytes per ray (DESIGN.md "Roofline"):
        32 (ray, 8 x f32) + 20 (hit record) + 80 * nodes/ray + 48 * prims/ray
    nodes/ray and prims/ray come from one untimed counting pass of the same workload (flags bit0);
    the kernel time is the sum of CUDA-event intervals around every k_wf_extend launch inside the
    timed steps (flags bit1, recorded on the stream the library launches on)."""
    from lajolla_public_b200 import RenderParams
    cp = RenderParams(spp=rp.spp, rng_mode=rp.rng_mode, engine=rp.engine, sample_begin=rp.sample_begin,
                      sample_end=rp.sample_end, flags=1)
    ct = scene.render_into_device(scratch_ptr, cp)
    closest = max(ct.rays_primary + ct.rays_extension, 1)
    nodes_per_ray = ct.nodes_visited / closest
    prims_per_ray = ct.prims_tested / closest
    bytes_ray = BYTES_RAY_FIXED + 80.0 * nodes_per_ray + 48.0 * prims_per_ray
    n_closest = rays_by_type["primary"] + rays_by_type["extension"]
    ext_s = stage["extend_ms"] * 1e-3
    achieved = n_closest * bytes_ray / ext_s * 1e-9 if ext_s > 0 else None
    sh = max(ct.rays_shadow, 1)
    step_ms = sum(stage.values())
    issue = None
    traffic = None
    try:   # dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture of this kernel (profiles/)
        # (the captured launch is an early, larger one: its bytes are scaled to the average launch by the ratio of the
        #  launch-list average duration to the captured launch's duration -- tools/summarize_profiles.py)
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["k_wf_extend<0>"]
        traffic = t.get("dram_bytes_per_avg_launch", t["dram_bytes_per_launch"])
        if t.get("warp_inst_per_launch") and t.get("launch_ms"):
            # the ceiling that actually binds this kernel (the BVH is L2-resident): warp instructions issued per second
            # against SMs x 4 schedulers x clock, and how many of the 32 lanes an issued instruction keeps busy
            sm_count, clock_hz = 148, 1.965e9
            per_s = t["warp_inst_per_launch"] / (t["launch_ms"] * 1e-3)
            issue = {"warp_inst_per_s": per_s, "peak_warp_inst_per_s": sm_count * 4 * clock_hz, "frac": per_s / (sm_count * 4 * clock_hz),
                     "lanes_per_inst": t.get("lanes_per_inst"), "lane_frac": (t.get("lanes_per_inst") or 0) / 32.0,
                     "useful_lane_issue_frac": per_s / (sm_count * 4 * clock_hz) * (t.get("lanes_per_inst") or 0) / 32.0,
                     "issue_active_pct": t.get("issue_active_pct"), "l2_hit_pct": t.get("l2_hit_pct"), "l1_hit_pct": t.get("l1_hit_pct"),
                     "registers": t.get("registers"), "source": "profiles/" + t.get("capture", "") + " (one ncu --set full launch, under the profiler's serialisation)"}
    except Exception:
        pass
    return {"bound": "hbm", "kernel": "k_wf_extend<false>", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak if achieved else None, "traffic": traffic, "issue": issue, "peak_source": which,
            "bytes_per_ray": bytes_ray, "nodes_per_ray": nodes_per_ray, "prims_per_ray": prims_per_ray,
            "launches": int(iterations), "avg_launch_ms": stage["extend_ms"] / max(iterations, 1),
            "bytes_per_launch": n_closest * bytes_ray / max(iterations, 1),
            "share_of_step": stage["extend_ms"] / step_ms if step_ms > 0 else None,
            "stage_ms": dict(stage),
            "shadow_nodes_per_ray": ct.shadow_nodes_visited / sh, "shadow_prims_per_ray": ct.shadow_prims_tested / sh,
            "note": "scene (BVH 8-wide nodes + triangles) is L2-resident; the HBM roofline is the SURVEY 8(d) convention, "
                    "the practical limiter is dependent-load latency + divergence (see profiles/)"}


_OUT = None


def emit(line):
    """the ONE JSON line, on the process's real stdout"""
    print(json.dumps(line), file=_OUT if _OUT is not None else sys.stdout, flush=True)


def main():
    # stdout carries exactly one JSON line: anything libraries print on fd 1 (NCCL's version banner, ...) goes to stderr
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--spp", type=int, default=SPP_PER_STEP)
    ap.add_argument("--engine", default="wavefront")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the fixed-work (strong scaling) block")
    ap.add_argument("--strong-spp", type=int, default=1024)
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
