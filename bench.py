#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 path-tracing backend (contract: see the task statement).

Metric (BASELINE.json): Msamples/s path tracing on the C3 workload (1920x1080, 4M-triangle scene with
diffuse/roughconductor/dielectric BSDFs, area + envmap emitters, max_depth 6), spp-sharded over N GPUs.
A "step" = one pass of the wavefront path tracer over 1920x1080 x --spp-per-step samples per GPU
(weak scaling: rank r renders its own disjoint sample-index range) + one NCCL all-reduce of the film.

  python bench.py --gpus 1 --steps 3 --warmup 3
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
  python bench.py --impl reference ...      (CPU oracle = C restatement of the reference path, all host threads)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    "c3": dict(desc="C3 path tracing 1920x1080, 4M-triangle displaced sphere (roughconductor) + textured diffuse ground + "
                    "100k-triangle dielectric sphere + area light + envmap, max_depth 6, rr_depth 5, box filter, rgb",
               width=1920, height=1080),
    "c3-small": dict(desc="C3 scene at reduced size (smoke)", width=480, height=270),
}


def build_scene(name):
    from mitsuba2_b200 import scenes
    w = WORKLOADS[name]
    if name == "c3":
        return scenes.c3_scene(width=w["width"], height=w["height"])
    return scenes.c3_scene(n_theta=100, n_phi=200, width=w["width"], height=w["height"], small=(32, 64), tex_res=64)


def side_metrics(mi, local):
    """BASELINE.json's other two metrics, measured briefly on rank 0 at N=1 and reported under `extra`:
    C2 intersection throughput (1M-triangle displaced sphere, 2.07M coherent / incoherent rays, closest and any hit) and
    the C4 gradient pass (512^2 x 64 spp, 512^2 x 3 albedo texels)."""
    import torch
    from mitsuba2_b200 import scenes
    out = {}
    sd = scenes.c2_scene()
    sc = mi.Scene(sd, device=local)
    for kind in ("coherent", "incoherent"):
        om, dm = scenes.c2_rays(sd, kind)
        o_t, d_t = torch.from_numpy(om).cuda(local), torch.from_numpy(dm).cuda(local)
        _, dms = scenes.c2_rays(sd, kind + "_shadow")
        ds_t = torch.from_numpy(dms).cuda(local)
        for mode, dd in (("closest", d_t), ("any", ds_t)):
            fn = sc.intersect_closest_raw if mode == "closest" else sc.intersect_any_raw
            ms = []
            for i in range(3 + 20):
                fn(o_t, dd)
                if i >= 3:
                    ms.append(sc.last_stats["ms_total"])
            out[f"c2_{kind}_{mode}_mrays_per_s"] = om.shape[0] / (float(np.median(ms)) * 1e-3) / 1e6
    del sc
    sd4, tex = scenes.c4_scene(os.path.join(ROOT, "tests", "golden", "cbox_scene.npz"))
    sc4 = mi.Scene(sd4, device=local)
    dl = torch.randn(512 * 512 * 3, device=f"cuda:{local}")
    ms = []
    for i in range(2 + 3):
        sc4.render_grad(dl, spp=64, seed=i, tex_id=tex)
        if i >= 2:
            ms.append(sc4.last_stats["ms_total"])
    out["c4_gradient_pass_ms"] = float(np.median(ms))
    out["c4_gradient_pass_msamples_per_s"] = 512 * 512 * 64 / (out["c4_gradient_pass_ms"] * 1e-3) / 1e6
    # the full C4 parameter set: albedo texture + the vertex positions of the 20k-triangle sphere in one pass
    # (interior term + edge-sampling boundary term at 0.25 edge samples per path sample)
    mid = next(i for i, m in enumerate(sd4.meshes) if m.name == "c4_sphere")
    ms, stages = [], None
    sc4.set_profiling(True)
    for i in range(2 + 3):
        sc4.render_grad(dl, spp=64, seed=i, tex_id=tex, mesh_id=mid)
        if i >= 2:
            ms.append(sc4.last_stats["ms_total"])
            stages = {k: sc4.last_stats[k] for k in ("ms_closest", "ms_shade", "ms_any", "ms_raygen")}
    out["c4_gradient_pass_texture_and_vertices_ms"] = float(np.median(ms))
    out["c4_gradient_pass_texture_and_vertices_stage_ms"] = stages
    out["c4_sphere_triangles"] = int(len(sd4.meshes[mid].faces))
    del sc4
    out.update(spectral_metric(mi, local))
    out["parity_gates"] = parity_gates(mi, local)
    return out


def multi_gpu_side_metrics(mi, local, rank, world, dist, torch, dev):
    """The other configurations at N > 1 (every rank runs this; device times are the max over ranks):
    C3 in the spectral mode, samples sharded like the headline (weak scaling, film all-reduced once);
    the C4 gradient pass with the samples of one optimisation step sharded over the ranks (strong scaling: 64 spp in total)
    and the texture + vertex gradients all-reduced -- the data-parallel step of C5 (one view per rank works the same way)."""
    from mitsuba2_b200 import scenes, distributed
    A = mi._abi
    out = {}

    def rmax(v):
        t = torch.tensor([v], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- C3 spectral, weak scaling
    sd = scenes.c3_scene()
    sd.mode = "spectral"
    sc = mi.Scene(sd, device=local)
    spp, steps, warm = 16, 3, 2
    npix = sd.sensors[0]["width"] * sd.sensors[0]["height"]
    film = torch.empty(npix * 5, dtype=torch.float32, device=dev)
    ms = 0.0
    for i in range(warm + steps):
        if i == warm:
            sc.film_clear()
            dist.barrier(); torch.cuda.synchronize()
        st = sc.render_accumulate(spp=spp, seed=1, spp_begin=(i * world + rank) * spp, spp_total=(warm + steps) * world * spp)
        if i >= warm:
            ms += st["ms_total"]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    A.check(A.lib().mtsb_film_read(sc._h, 0, film.data_ptr(), A.MEM_DEVICE))
    dist.all_reduce(film)
    e1.record(); e1.synchronize()
    ms = rmax(ms + e0.elapsed_time(e1))
    out["c3_spectral_msamples_per_s"] = npix * spp * steps * world / (ms * 1e-3) / 1e6
    out["c3_spectral_scaling"] = "weak (16 spp per GPU per step, film all-reduced once per job)"
    del sc
    # ---- C4 gradient pass, 64 spp shared by the ranks, gradients all-reduced
    sd4, tex = scenes.c4_scene(os.path.join(ROOT, "tests", "golden", "cbox_scene.npz"))
    mid = next(i for i, m in enumerate(sd4.meshes) if m.name == "c4_sphere")
    sc4 = mi.Scene(sd4, device=local)
    dl = torch.randn(512 * 512 * 3, device=dev)
    dist.broadcast(dl, 0)
    lo, cnt = distributed.sample_share(64, rank, world)
    for name, kw in (("texture", {}), ("texture_and_vertices", {"mesh_id": mid})):
        t_all = []
        for i in range(2 + 3):
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            gt, gv = sc4.render_grad(dl, spp=cnt, seed=i, tex_id=tex, spp_begin=lo, spp_total=64, **kw)
            dist.all_reduce(gt)
            if gv is not None:
                dist.all_reduce(gv)
            e1.record(); e1.synchronize()
            if i >= 2:
                t_all.append(rmax(e0.elapsed_time(e1)))
        out[f"c4_gradient_pass_{name}_ms"] = float(np.median(t_all))
    out["c4_scaling"] = f"strong (64 spp per step split over {world} GPUs, texture + vertex gradients all-reduced inside the timed region)"
    return out


def spectral_metric(mi, local, spp=16, steps=3, warmup=2):
    """C3 in the spectral mode (`cuda_b200_spectral`: 4 wavelengths per path, same scene authored in RGB + Cu optical
    constants): Msamples/s on one GPU, inputs resident, same timing as the headline."""
    import time
    from mitsuba2_b200 import scenes
    sd = scenes.c3_scene()
    sd.mode = "spectral"
    t0 = time.time()
    sc = mi.Scene(sd, device=local)
    out = {"c3_spectral_scene_setup_s": time.time() - t0}
    sc.set_profiling(True)
    ms, stages = [], {}
    for i in range(warmup + steps):
        sc.film_clear()
        st = sc.render_accumulate(spp=spp, seed=i, spp_begin=0, spp_total=spp)
        if i >= warmup:
            ms.append(st["ms_total"])
            for k in ("ms_closest", "ms_shade", "ms_any"):
                stages[k] = stages.get(k, 0.0) + st[k] / steps
    n = sd.sensors[0]["width"] * sd.sensors[0]["height"] * spp
    out["c3_spectral_msamples_per_s"] = n / (float(np.median(ms)) * 1e-3) / 1e6
    out["c3_spectral_stage_ms_per_step"] = stages
    out["c3_spectral_samples_per_step"] = n
    return out


def parity_gates(mi, local):
    """the parity numbers reported next to every perf number, against the CPU oracle (checker only):
    hit-id mismatch rate, image agreement at identical sample streams, gradient cosine vs finite differences."""
    import helpers as H
    g = {}
    # at the sizes the numbers are quoted on (same functions as tests/test_gpu_bench_parity.py): C1 cbox 256^2 x 64 spp
    # (vs the oracle, the committed reference-CLI render and a live reference render), C2 1M-triangle hit ids (vs the
    # oracle and the reference kd-tree), C3 4.1M-triangle film crop (same streams + z-test vs the oracle)
    try:
        import test_gpu_bench_parity as BP
        g["c1_cbox_256x256x64spp"] = BP.c1_parity(mi, local)
        g["c2_1M_triangles"] = BP.c2_parity(mi, device=local)
        g["c3_4M_triangles_crop"] = BP.c3_parity(mi, local, run_spp=32, gpu_runs=32, cpu_runs=8)
        g["c4_512x512_primal"] = BP.c4_parity(mi, local)
        g["c3_4M_triangles_crop_spectral"] = BP.c3_spectral_parity(mi, local)
    except Exception as e:
        g["bench_size_error"] = repr(e)
    # C4 at its own size (512^2, 512^2 x 3 texels, 20k-triangle sphere): adjoint vs central finite differences of the CUDA
    # renderer at identical seeds (tools/c4_fd_check.py; the full protocol -- 64 texel blocks, 32 vertex coordinates, epsilon
    # sweep, 4096 spp -- is run by hand and tabulated in DESIGN.md section 4b; this is the bounded version)
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import c4_fd_check
        r = c4_fd_check.run(res=512, spp=64, passes=1, fd_spp=1024, n_tex=16, n_vtx=0, eps_list=(0.25, 0.5), device=local, same_samples=True)
        g["c4_512x512_fd"] = {"texture_rel_l2": r["texture"]["rel_l2"], "texture_cosine": r["texture"]["cosine"],
                              "vertex_fields_rel_l2": r.get("vertex_fields_rel_l2_richardson", r["vertex_fields_rel_l2"]),
                              "vertex_fields": {k: {"adjoint": v["adjoint"], "fd": v.get("fd_richardson", v["fd_by_eps"]["0.5"][0])} for k, v in r["vertex_fields"].items()},
                              "fd": "central differences at 0.25 and 0.5 mean edge lengths, Richardson-extrapolated",
                              "fd_spp": 1024, "same_samples": True}
    except Exception as e:
        g["c4_fd_error"] = repr(e)
    sd = H.sphere_scene(n_theta=128, n_phi=256)
    sc, orc = mi.Scene(sd, device=local), H.Oracle(sd)
    om, dm = H.random_rays(200000, 5, radius=1.5)
    a = np.ascontiguousarray(sc.intersect_closest_raw(om, dm)[:, 3]).view(np.uint32)
    b = np.ascontiguousarray(orc.intersect_closest(om, dm)[:, 3]).view(np.uint32)
    g["hit_id_mismatch_rate"] = float(np.mean(a != b))
    sdc = H.cbox_scene(res=64, spp=16, rfilter="box")
    scc, orcc = mi.Scene(sdc, device=local), H.Oracle(sdc)
    ia, ib = scc.render(spp=16, seed=3), orcc.develop(orcc.render_xyzaw(spp=16, seed=3))
    g["cbox_same_streams_rel_mse"] = float(np.mean((ia - ib) ** 2) / np.mean(ib ** 2))
    g["cbox_same_streams_max_abs"] = float(np.abs(ia - ib).max())
    sdt, tex = H.textured_cbox(res=16, spp=32, tex_res=5)
    sct, orct = mi.Scene(sdt, device=local), H.Oracle(sdt)
    rng = np.random.default_rng(0)
    wgt = rng.normal(size=(16, 16, 3)).astype(np.float32)
    gt, _ = sct.render_grad(wgt, spp=32, seed=5, tex_id=tex)
    gt = gt.reshape(sdt.textures[tex]["data"].shape)
    base = sdt.textures[tex]["data"].copy()

    def loss(data):
        orct.update_texture(tex, data)
        return float(np.sum(np.float64(wgt) * orct.develop(orct.render_xyzaw(seed=5))))

    fd, an = [], []
    for y, x, c in zip(rng.integers(0, 5, 16), rng.integers(0, 5, 16), rng.integers(0, 3, 16)):
        d = base.copy()
        d[y, x, c] += 0.02
        lp = loss(d)
        d[y, x, c] -= 0.04
        fd.append((lp - loss(d)) / 0.04)
        an.append(gt[y, x, c])
    fd, an = np.array(fd), np.array(an)
    g["texture_grad_cosine_vs_fd"] = float(np.dot(fd, an) / (np.linalg.norm(fd) * np.linalg.norm(an)))
    g["texture_grad_rel_l2_vs_fd"] = float(np.linalg.norm(fd - an) / np.linalg.norm(fd))
    # vertex-position adjoint vs central finite differences of the oracle along global displacement fields:
    # interior term (plane under a point light: same estimator sample by sample) and boundary term (quad on a background)
    for kind, fields, eps, fd_spp, name in (("plane", ("tz", "tilt"), 2e-3, 32, "interior"), ("quadenv", ("tx", "ty", "scale"), 0.02, 1024, "boundary"),
                                        ("quadshadow", ("tx", "tz", "scale"), 0.02, 1024, "shadow_boundary")):
        sdv = H.vgrad_scene(kind, "ggx" if kind == "plane" else "diffuse", res=16, spp=32 if kind == "plane" else 128)
        yy, xx = np.mgrid[0:16, 0:16]
        w = np.stack([xx / 16, yy / 16, 0.5 + 0 * xx], -1).astype(np.float32)
        scv = mi.Scene(sdv, device=local)
        scv.grad_config(boundary=(0 if kind == "plane" else 2), edge_fraction=1.0)
        _, gv = scv.render_grad(w, seed=3, mesh_id=0)
        P0 = np.asarray(sdv.meshes[0].positions, np.float32).reshape(-1, 3)
        F = H.displacement_fields(P0)
        fd = np.array([H.oracle_fd_vertices(sdv, 0, w, F[n].astype(np.float32), eps, fd_spp, 3) for n in fields])
        an = np.array([float((gv * F[n]).sum()) for n in fields])
        g[f"vertex_grad_{name}_rel_l2_vs_fd"] = float(np.linalg.norm(fd - an) / np.linalg.norm(fd))
    return g


def ncu_profile(kernel="k_trace<0>"):
    """Per-launch ncu numbers of `kernel` from the committed capture of this same command (profiles/r2_kernel_profile.json,
    recipe in profiles/README.md; falls back to the round-1 file): DRAM bytes per launch (`roofline.traffic`), thread
    instructions per ray and active threads per instruction -- the two numbers that actually bound this kernel."""
    for name in ("r2_kernel_profile.json", "r1_kernel_traffic.json"):
        try:
            d = json.load(open(os.path.join(ROOT, "profiles", name)))[kernel]
            d = dict(d)
            d["source"] = f"ncu per k_trace<closest> launch, profiles/{name}"
            return d
        except Exception:
            continue
    return {}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return json.load(open(path))["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        inside = [r for t, r in self.rows if (t0 is None or t >= t0) and (t1 is None or t <= t1 + 0.25)]
        self.rows = inside if inside else [r for _, r in self.rows[-3:]]
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


REF_SPP_PER_STEP = 2          # reference arm: one step = the full 1920x1080 film at 2 spp (4.1 M samples, ~2.5 s on 16 cores)


def reference_render_seconds(sd, spp, reps, threads, workdir):
    """Times `Integrator::render` of the UNMODIFIED reference (oracle/_ref/mitsuba, built by oracle/build_reference.sh) on
    the exported scene: oracle/ref_query.cpp loads the XML once with the reference's loader and calls the scene's `path`
    integrator `reps` times (src/librender/integrator.cpp:52-176; scene loading and kd-tree build are outside, like the
    CLI's "Rendering finished" timer).  -> list of seconds per repetition, or None when the build is absent / fails."""
    import helpers
    if not helpers.RefBuild.available(query=True):
        return None
    try:
        ref = helpers.RefBuild(sd, workdir, spp=spp)
        r = subprocess.run([ref.QUERY, ref.xml, "render", "/dev/null", os.path.join(workdir, "render_out.bin"), str(threads), str(reps)],
                           env=ref._env(), capture_output=True, text=True, timeout=3000)
        import re
        secs = [float(x) for x in re.findall(r"render_seconds ([0-9.]+)", r.stdout)]
        return secs if r.returncode == 0 and len(secs) == reps else None
    except Exception as e:      # e.g. a host CPU without the instruction set the reference was compiled for
        print(f"[bench] reference build unusable here ({e!r}); falling back to the oracle port", file=sys.stderr)
        return None


def run_reference(args):
    """Reference arm: the reference's own CPU implementation of the path on this box's host cores -- the unmodified
    `scalar_rgb` build (kind "reference") where oracle/_ref/mitsuba travelled with the repo, else the C oracle (kind "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import tempfile
    import helpers
    sd = build_scene(args.workload)
    cores = os.cpu_count() or 1
    s = sd.sensors[0]
    W, K = args.warmup, args.steps
    with tempfile.TemporaryDirectory(prefix="mtsb_ref_") as tmp:
        secs = reference_render_seconds(sd, REF_SPP_PER_STEP, W + K, cores, tmp)
    if secs is not None:
        samples = s["width"] * s["height"] * REF_SPP_PER_STEP
        dt = sum(secs[W:])
        kind = "reference"
        sample = (f"full {s['width']}x{s['height']} film x {REF_SPP_PER_STEP} spp per step ({samples} samples/step), unmodified Mitsuba 2 "
                  f"scalar_rgb `path` integrator (kd-tree), TBB on {cores} threads; Msamples/s does not depend on spp")
    else:
        # bounded sample of the same workload: same scene and camera, 1/4-resolution film
        from mitsuba2_b200.scenedata import perspective_sample_to_camera
        s["width"], s["height"] = max(16, s["width"] // 4), max(9, s["height"] // 4)
        s["sample_to_camera"] = perspective_sample_to_camera(s["fov_x"], s["near"], s["far"], s["width"] / s["height"]).reshape(16)
        orc = helpers.Oracle(sd, threads=cores)
        SPP = 8
        samples = s["width"] * s["height"] * SPP
        total = (W + K) * SPP
        for i in range(W):
            orc.render_xyzaw(spp=SPP, seed=1, spp_begin=i * SPP, spp_total=total)
        t0 = time.perf_counter()
        for i in range(K):
            orc.render_xyzaw(spp=SPP, seed=1, spp_begin=(W + i) * SPP, spp_total=total)
        dt = time.perf_counter() - t0
        kind = "port"
        sample = f"{s['width']}x{s['height']} x {SPP} spp per step of the same scene/camera ({samples} samples/step), C oracle (reference build absent)"
    value = samples * K / dt / 1e6
    print(json.dumps({
        "impl": "reference", "metric": "path_tracing_msamples_per_s", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": K, "warmup": W, "ms_per_step": dt / K * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload]["desc"], "sample": sample},
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def cpu_baseline(sd_full, budget_s=20.0):
    """CPU baseline timed on this box's host cores on a bounded sample of the workload: the unmodified reference
    (kind "reference": full film, 2 spp, 1 warm-up + 3 timed renders) where its build is present, else the oracle port."""
    import copy
    import tempfile
    import helpers
    from mitsuba2_b200.scenedata import perspective_sample_to_camera
    cores = os.cpu_count() or 1
    s0 = sd_full.sensors[0]
    with tempfile.TemporaryDirectory(prefix="mtsb_ref_") as tmp:
        secs = reference_render_seconds(sd_full, REF_SPP_PER_STEP, 4, cores, tmp)
    if secs is not None:
        samples = s0["width"] * s0["height"] * REF_SPP_PER_STEP
        return {"value": samples * 3 / sum(secs[1:]) / 1e6, "unit": "Msamples/s", "cores": cores, "kind": "reference",
                "sample": f"full {s0['width']}x{s0['height']} film x {REF_SPP_PER_STEP} spp, 3 renders after 1 warm-up ({3 * samples} samples, "
                          f"{sum(secs[1:]):.1f} s), unmodified Mitsuba 2 scalar_rgb path integrator on {cores} TBB threads"}
    sd = copy.copy(sd_full)
    sd.sensors = [dict(sd_full.sensors[0])]
    s = sd.sensors[0]
    s["width"], s["height"] = max(16, s["width"] // 4), max(9, s["height"] // 4)
    s["sample_to_camera"] = perspective_sample_to_camera(s["fov_x"], s["near"], s["far"], s["width"] / s["height"]).reshape(16)
    orc = helpers.Oracle(sd, threads=cores)
    orc.render_xyzaw(spp=1, seed=1, spp_begin=0, spp_total=64)          # warm-up
    n, t0 = 0, time.perf_counter()
    while True:
        orc.render_xyzaw(spp=1, seed=1, spp_begin=1 + n, spp_total=64)
        n += 1
        if time.perf_counter() - t0 > budget_s / 2 or n >= 32:
            break
    dt = time.perf_counter() - t0
    samples = s["width"] * s["height"] * n
    return {"value": samples / dt / 1e6, "unit": "Msamples/s", "cores": cores, "kind": "port",
            "sample": f"{s['width']}x{s['height']} film of the same scene/camera, {n} spp ({samples} samples, {dt:.1f} s), C oracle"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=list(WORKLOADS))
    ap.add_argument("--spp-per-step", type=int, default=16)
    ap.add_argument("--strong-spp", type=int, default=256, help="fixed total spp of the strong-scaling side measurement (0 = skip)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-side-metrics", action="store_true", help="skip the C2 intersect / C4 gradient side measurements")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import mitsuba2_b200 as mi
    from mitsuba2_b200 import distributed as D
    A = mi._abi
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL writes its version banner / warnings to stdout; NCCL_DEBUG_FILE is honoured above the VERSION level only.
        # Keep stdout to the one JSON line: raise VERSION to WARN (which still prints the banner) and send it all to stderr.
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    mi.set_variant("cuda_b200_rgb")
    t_build = time.perf_counter()
    sd = build_scene(args.workload)
    scene = mi.Scene(sd, device=local)
    build = scene.build_stats()
    t_build = time.perf_counter() - t_build
    s = sd.sensors[0]
    npix = s["width"] * s["height"]
    spp, W, K = args.spp_per_step, args.warmup, args.steps
    spp_total = (2 * (W + K) + 2) * world * spp
    film_t = [torch.empty(npix * 5, dtype=torch.float32, device=dev) for _ in range(2)]
    ev = lambda: torch.cuda.Event(enable_timing=True)      # noqa: E731

    def render_step(i):
        """one bench step on this rank: this rank's disjoint sample-index range of the job, accumulated into its device film"""
        return scene.render_accumulate(spp=spp, seed=1, spp_begin=(i * world + rank) * spp, spp_total=spp_total)

    def film_to(t):
        A.check(A.lib().mtsb_film_read(scene._h, 0, t.data_ptr(), A.MEM_DEVICE))      # k_film_pack straight into t (1 launch)

    def reduce_film(t):
        """the job's ONE film reduce: pack this rank's XYZAW sums and all-reduce them; -> device ms"""
        e0, e1 = ev(), ev()
        e0.record()
        film_to(t)
        if world > 1:
            dist.all_reduce(t)
        e1.record()
        e1.synchronize()
        return e0.elapsed_time(e1)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(*vals):
        if world == 1:
            return list(vals)
        t = torch.tensor(vals, device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()

    scene.set_profiling(True)
    scene.film_clear()
    clocks = ClockSampler(local)
    clocks.start()
    for i in range(W):
        render_step(i)
    reduce_film(film_t[0])
    scene.film_clear()
    barrier()
    # ---- timed region: K steps of the job (film accumulates on the device) + the job's one film reduce
    t0 = time.perf_counter()
    dev_ms, launches, agg = 0.0, 0, {}
    for i in range(W, W + K):
        st = render_step(i)
        dev_ms += st["ms_total"]
        launches += st["kernel_launches"]
        for k in ("samples", "closest_rays", "shadow_rays", "segments", "ms_closest", "ms_any", "ms_shade", "ms_raygen", "ms_film"):
            agg[k] = agg.get(k, 0) + st[k]
    reduce_ms = reduce_film(film_t[0])
    dev_ms += reduce_ms
    launches += 1 + (1 if world > 1 else 0)
    barrier()
    t1 = time.perf_counter()
    wall_ms = (t1 - t0) * 1e3
    clk = clocks.stop(t0, t1)
    dev_ms, wall_ms, reduce_ms = max_over_ranks(dev_ms, wall_ms, reduce_ms)
    samples_all = npix * spp * K * world
    value = samples_all / (dev_ms * 1e-3) / 1e6

    # ---- N-rank film == 1-rank film of the union of the sample indices (one extra 1-spp step; checked inside the run
    # because the driver's GPU test box has one GPU)
    mg_parity = None
    if world > 1:
        scene.film_clear()
        scene.render_accumulate(spp=1, seed=5, spp_begin=rank, spp_total=world)
        reduce_film(film_t[0])
        if rank == 0:
            scene.film_clear()
            scene.render_accumulate(spp=world, seed=5, spp_begin=0, spp_total=world)
            film_to(film_t[1])
            a, b = film_t[0], film_t[1]
            mg_parity = {"max_abs": float((a - b).abs().max()), "max_abs_over_max": float((a - b).abs().max() / b.abs().max()),
                         "rel_l1": float((a - b).abs().sum() / b.abs().sum()), "spp": world}
        barrier()

    # ---- strong scaling: ONE fixed job (full film x --strong-spp), sample indices split by distributed.sample_share,
    # one film reduce at the end; the same job at every N, so time(1) / (N time(N)) is the strong-scaling efficiency
    strong = None
    if args.strong_spp > 0:
        begin, count = D.sample_share(args.strong_spp, rank, world)
        scene.film_clear()
        barrier()
        ms = 0.0
        if count > 0:
            ms = scene.render_accumulate(spp=count, seed=2, spp_begin=begin, spp_total=args.strong_spp)["ms_total"]
        ms += reduce_film(film_t[0])
        (ms,) = max_over_ranks(ms)
        strong = {"job": f"{s['width']}x{s['height']} x {args.strong_spp} spp, spp-sharded over {world} GPU(s), one film all-reduce",
                  "ms": ms, "msamples_per_s": npix * args.strong_spp / (ms * 1e-3) / 1e6, "spp_per_rank": count}

    # ---- e2e: the user's call with HOST buffers.  Every step: upload the albedo texture from pinned memory (every rank),
    # clear, render this rank's share, reduce the film over the ranks and read the step's film back into pinned host memory
    # (rank 0).  The reduce + read-back of step i run on a side stream under the render of step i+1 (double-buffered film copy);
    # the timed region ends when the last read-back has landed in host memory.
    tex_id = next(i for i, t in enumerate(sd.textures) if t["type"] == A.TEX_BITMAP)
    tex_host = torch.from_numpy(sd.textures[tex_id]["data"].copy()).pin_memory()
    film_host = [torch.empty(npix * 5, dtype=torch.float32).pin_memory() for _ in range(2)]
    side = torch.cuda.Stream(device=dev)
    done = [None, None]

    def e2e_step(n, i):
        k = n & 1
        scene.update_texture(tex_id, tex_host.numpy())
        scene.film_clear()
        render_step(i)
        if done[k] is not None:
            done[k].synchronize()          # buffer k is free again
        film_to(film_t[k])                  # returns after the pack kernel finished (library stream synchronised)
        with torch.cuda.stream(side):
            if world > 1:
                dist.all_reduce(film_t[k])
            if rank == 0:
                film_host[k].copy_(film_t[k], non_blocking=True)
            done[k] = torch.cuda.Event()
            done[k].record(side)

    for n in range(2):                      # untimed: first use of the side stream, the pinned buffers and the NCCL channel on it
        e2e_step(n, W + K + n)
    side.synchronize()
    barrier()
    t0 = time.perf_counter()
    for n, i in enumerate(range(W + K, W + 2 * K)):
        e2e_step(n, i)
    side.synchronize()
    barrier()
    e2e_s = time.perf_counter() - t0
    (e2e_s,) = max_over_ranks(e2e_s)
    e2e = {"value": samples_all / e2e_s / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": int(tex_host.numel() * 4) * world,
           "d2h_bytes_per_step": int(film_host[0].numel() * 4),
           "note": "per step: pinned texture upload on every rank, render, film all-reduce (N > 1), rank 0 reads the film back into pinned host memory; the read-back of step i overlaps the render of step i+1"}

    bvh_update = None
    if rank == 0 and not args.no_side_metrics:
        # build and refit timed separately -- the big mesh's own positions pushed through update_vertices
        # (device pointer; vertex normals recomputed on the GPU), once as a full rebuild and once as a refit
        try:
            p = torch.from_numpy(np.ascontiguousarray(sd.meshes[0].positions, np.float32)).to(dev)
            bvh_update = {"triangles": int(sum(len(m.faces) for m in sd.meshes))}
            for name, refit in (("rebuild", False), ("refit", True)):
                scene.update_config(refit=refit)
                scene.update_vertices(0, p); torch.cuda.synchronize()
                t0 = time.perf_counter()
                scene.update_vertices(0, p); torch.cuda.synchronize()
                bvh_update[name + "_call_ms"] = (time.perf_counter() - t0) * 1e3
                bvh_update[name + "_bvh_ms"] = scene.build_stats()["ms_build"]
            scene.update_config(refit=False)
        except Exception as e:
            bvh_update = {"error": repr(e)}

    mg_side = None
    if world > 1 and not args.no_side_metrics:
        del scene
        try:
            mg_side = multi_gpu_side_metrics(mi, local, rank, world, dist, torch, dev)
        except Exception as e:
            mg_side = {"error": repr(e)}

    if rank == 0:
        peak, peak_src = peaks()
        prof = ncu_profile()
        # dominant kernel = closest-hit traversal; algorithmic bytes = 48 B per ray (32 B ray in + 16 B hit out)
        ms_c = max(agg["ms_closest"], 1e-9)
        pool = int(os.environ.get("MTSB_POOL", 1 << 25))
        n_trace_launches = K * (-(-npix * spp // pool)) * sd.max_depth      # one closest-hit launch per batch and depth
        achieved = agg["closest_rays"] * 48 / (ms_c * 1e-3) / 1e9
        stage_ms = {k: agg[k] / K for k in ("ms_raygen", "ms_closest", "ms_shade", "ms_any", "ms_film")}
        bytes_per_sample = (20 * agg["samples"] + 224 * agg["segments"] + 96 * agg["shadow_rays"]) / max(1, agg["samples"])
        out = {
            "metric": "path_tracing_msamples_per_s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload]["desc"], "envmap": getattr(sd, "envmap_name", "synthetic sky 1024x512"),
                       "spp_per_step_per_gpu": spp, "samples_per_step": npix * spp * world,
                       "parallelism": f"spp-sharded x{world}; the K-step job accumulates on each GPU and its film is all-reduced ONCE, inside the timed region",
                       "l2": "queues per step (>1 GB) exceed L2; no flush needed", "timing": "CUDA events on the launching stream, max over ranks"},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clk,
            "roofline": {"bound": "hbm", "kernel": "k_trace<closest>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": prof.get("dram_bytes_per_launch"), "peak_source": peak_src,
                         "algorithmic_bytes": "48 B/ray x rays per launch",
                         "algorithmic_bytes_per_launch": agg["closest_rays"] * 48 / max(1, n_trace_launches),
                         "traffic_source": prof.get("source"),
                         "limiter": "SM instruction issue / ALU pipe, not HBM (the kernel walks L1/L2-resident BVH nodes); see profiles/",
                         "thread_instructions_per_ray": prof.get("thread_inst_per_ray"),
                         "active_threads_per_instruction": prof.get("threads_per_inst"),
                         "issue_slots_busy_pct": prof.get("issue_slots_busy_pct"),
                         "mrays_per_s": agg["closest_rays"] / (ms_c * 1e-3) / 1e6},
            "extra": {"wall_ms_per_step": wall_ms / K, "stage_ms_per_step": stage_ms, "film_reduce_ms_per_job": reduce_ms,
                      "rays_per_sample": agg["closest_rays"] / agg["samples"],
                      "shadow_rays_per_sample": agg["shadow_rays"] / agg["samples"], "algorithmic_bytes_per_sample": bytes_per_sample,
                      "whole_step_gbs": bytes_per_sample * agg["samples"] / (sum(stage_ms.values()) * K * 1e-3) / 1e9,
                      "any_hit_mrays_per_s": agg["shadow_rays"] / max(agg["ms_any"], 1e-9) / 1e3,
                      "bvh": {"ms_build": build["ms_build"], "nodes": build["bvh_nodes"], "depth": build["bvh_depth"], "update_vertices": bvh_update},
                      "scene_setup_s": t_build, "strong_scaling": strong, "multi_gpu_parity": mg_parity},
        }
        if mg_side is not None:
            out["extra"]["side_metrics_multi_gpu"] = mg_side
        if world == 1 and not args.no_side_metrics:
            del scene
            try:
                out["extra"]["side_metrics"] = side_metrics(mi, local)
            except Exception as e:      # the headline line must still be printed
                out["extra"]["side_metrics"] = {"error": repr(e)}
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(sd)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
