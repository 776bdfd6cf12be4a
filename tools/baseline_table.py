#!/usr/bin/env python
"""Markdown rows of BASELINE.md section 4 from a GPU visit's outputs:
tools/baseline_table.py gpurun_out/scenes.json gpurun_out/pytest_gpu.log [peak GB/s]"""
import json
import re
import sys

scenes = json.load(open(sys.argv[1]))
log = open(sys.argv[2]).read() if len(sys.argv) > 2 else ""
peak = float(sys.argv[3]) if len(sys.argv) > 3 else 6555.2
rmse = {m.group(1): (m.group(2), float(m.group(3)), float(m.group(4)), float(m.group(5)))
        for m in re.finditer(r"(\w+) (\d+x\d+) 1024 spp: RMSE vs reference render ([\d.e+-]+) \(floor ([\d.e+-]+), bound ([\d.e+-]+)\)", log)}
match = {}
for m in re.finditer(r"(\w+) mode A: pixel match ([\d.]+), tile match ([\d.]+)", log):
    match[m.group(1)] = (float(m.group(2)), float(m.group(3)))
print("| scene | film, spp | render ms | Msamples/s | Mrays/s | rays: primary / extension / shadow (M) | N̄_node, N̄_prim closest (shadow) | B/ray closest (shadow) | algorithmic GB/s, fraction of %.0f | 1-spp match pixels / tiles (mode A) | RMSE@1024 vs reference render (floor, bound) | CPU oracle Msamples/s, Mrays/s (threads) |" % peak)
print("|---|---|---|---|---|---|---|---|---|---|---|---|")
for name, e in scenes.items():
    w, c, cpu = e["wavefront"], e.get("counting"), e.get("cpu_oracle")
    if not c:
        continue
    r = rmse.get(name)
    mt = match.get(name)
    print(f"| {name} | {e['width']}×{e['height']}, {e['spp']} | {w['render_ms']:.1f} | {w['msamples_per_s']:.1f} | {w['mrays_per_s']:.0f} | "
          f"{c['rays_primary'] / 1e6:.1f} / {c['rays_extension'] / 1e6:.1f} / {c['rays_shadow'] / 1e6:.1f} | "
          f"{c['nodes_per_ray']:.2f}, {c['prims_per_ray']:.2f} ({c['shadow_nodes_per_ray']:.2f}, {c['shadow_prims_per_ray']:.2f}) | "
          f"{c['bytes_per_ray']:.0f} ({c['shadow_bytes_per_ray']:.0f}) | {c['algorithmic_gbps_whole_render']:.0f}, {c['algorithmic_gbps_whole_render'] / peak:.3f} | "
          + (f"{mt[0]:.5f} / {mt[1]:.4f}" if mt else "—") + " | "
          + (f"{r[1]:.5f} ({r[2]:.5f}, {r[3]:.5f}) at {r[0]}" if r else "—") + " | "
          + (f"{cpu['msamples_per_s']:.2f}, {cpu['mrays_per_s']:.1f} ({cpu['threads']})" if cpu else "—") + " |")
